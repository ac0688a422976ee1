"""Every BASELINE.json config through the public API at its native size: GPU step time (with the nvidia-smi clocks
sampled while it runs) next to the CPU oracle's step time for the SAME problem on the box's host cores.  One JSON line
per config among the progress chatter of fit() (grep "^{": profiles/r02_configs.txt).  python tools/gpu_configs.py [--no-cpu]"""
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import tensordiffeq_b200 as tdq
from tensordiffeq_b200 import tf
from tensordiffeq_b200.boundaries import *
from tensordiffeq_b200.models import CollocationSolverND, DiscoveryModel
from tests.test_jet_host import f_ac, f_burgers, f_poisson, f_ch
from tests.gpu_util import deriv_model

NO_CPU = "--no-cpu" in sys.argv


def oracle_terms(model):
    from oracle import pinn_oracle as po
    terms = []
    for bc in model.bcs:
        if bc.isPeriodic:
            terms.append(po.Term("periodic", upper=np.asarray(bc.upper, np.float64), lower=np.asarray(bc.lower, np.float64), var=["x"]))
        else:
            val = np.asarray(bc.val.numpy() if hasattr(bc.val, "numpy") else bc.val, np.float32)
            X = np.asarray(bc.input, np.float64)
            val = np.broadcast_to(val.reshape(-1, 1) if val.ndim else val, (len(X), 1)).astype(np.float32)
            terms.append(po.Term("ic" if bc.isInit else "dirichlet", X=X, val=val))
    return terms


def cpu_collocation(model, residual, steps=3, lam=None, lam_bcs=None):
    """the oracle's Adam step (nested autograd + Keras Adam) on the same points / terms, all host cores"""
    import torch
    from oracle import pinn_oracle as po
    torch.set_num_threads(os.cpu_count() or 1)
    prob = po.Problem(model.layer_sizes, residual, np.asarray(model.domain.X_f, np.float64), oracle_terms(model))
    opt = po.KerasAdam()
    w = model.u_model.get_flat().copy()
    ts = []
    for i in range(steps + 1):
        t0 = time.perf_counter()
        r = po.loss_and_grads(prob, w, lam, lam_bcs)
        opt.begin_step()
        w = opt.apply("w", w, r["grad"])
        if i:
            ts.append(time.perf_counter() - t0)
    return float(np.mean(ts))


def gpu_timed(run, n, sync):
    run(20)
    sync()
    sampler = bench.ClockSampler(int(os.environ.get("TDQ_DEVICE", "0")))
    time.sleep(0.3)
    s0 = sampler.count()
    t0 = time.perf_counter()
    reps = 0
    while True:                                   # at least 0.5 s under load so nvidia-smi sees it
        run(n)
        sync()
        reps += 1
        if time.perf_counter() - t0 > 0.5:
            break
    dt = (time.perf_counter() - t0) / (n * reps)
    return dt, sampler.stop(skip=s0)


def report(name, npts, dt, clocks, kernel, cpu_s, extra=None):
    line = {"config": name, "n_points": npts, "gpu": {"ms_per_step": dt * 1e3, "points_per_s": npts / dt, "kernel": kernel}, "clocks": clocks}
    if cpu_s is not None:
        line["cpu_baseline"] = {"s_per_step": cpu_s, "points_per_s": npts / cpu_s, "cores": os.cpu_count(), "kind": "port",
                                "sample": "native size, 3 timed steps of the oracle (torch CPU nested autograd + Keras Adam)"}
        line["gpu_over_cpu"] = cpu_s / dt
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)


def colloc(name, model, npts, residual, n=100, lam_cpu=None, **kw):
    model.verbose = False
    dt, clocks = gpu_timed(lambda k: model.fit(tf_iter=k, **kw), n, lambda: model._engine().sync())
    eng = model._engine()
    kernel = eng.term_kernel(model._terms[-1]["tid"])
    cpu = None
    if not NO_CPU:
        lam, lam_bcs = (None, None)
        if lam_cpu:
            lam = np.asarray(model.lambdas[0].numpy(), np.float32)
            lam_bcs = {0: np.asarray(model.lambdas[1].numpy(), np.float32)}
        cpu = cpu_collocation(model, residual, 3, lam, lam_bcs)
    report(name, npts, dt, clocks, kernel, cpu)
    return model


def ac(N_f, sa):
    D = DomainND(["x", "t"], time_var="t"); D.add("x", [-1.0, 1.0], 512); D.add("t", [0.0, 1.0], 201)
    D.generate_collocation_points(N_f)
    init = IC(D, [lambda x: x ** 2 * np.cos(math.pi * x)], var=[["x"]])
    per = periodicBC(D, ["x"], [deriv_model])
    m = CollocationSolverND(verbose=False)
    if sa:
        m.compile([2, 128, 128, 128, 128, 1], f_ac, D, [init, per], isAdaptive=True,
                  dict_adaptive={"residual": [True], "BCs": [True, False]},
                  init_weights={"residual": [tf.random.uniform([N_f, 1])], "BCs": [100 * tf.random.uniform([512, 1]), None]})
    else:
        m.compile([2, 128, 128, 128, 128, 1], f_ac, D, [init, per])
    return m


if __name__ == "__main__":
    # config 1: Burgers
    D = DomainND(["x", "t"], time_var="t"); D.add("x", [-1.0, 1.0], 256); D.add("t", [0.0, 1.0], 100)
    D.generate_collocation_points(10000)
    bcs = [IC(D, [lambda x: -np.sin(x * math.pi)], var=[["x"]]), dirichletBC(D, val=0.0, var="x", target="upper"),
           dirichletBC(D, val=0.0, var="x", target="lower")]
    m = CollocationSolverND(verbose=False); m.compile([2] + [20] * 8 + [1], f_burgers, D, bcs)
    emit = print
    if True:
        colloc("1 Burgers [2,20x8,1] N_f=10000 Adam", m, 10000, "burgers", 200)
        t0 = time.perf_counter(); m.fit(tf_iter=0, newton_iter=200); m._engine().sync()
        emit(json.dumps({"config": "1 Burgers eager L-BFGS", "ms_per_iter": (time.perf_counter() - t0) * 5, "iterations": 200, "P": 3021}))
        t0 = time.perf_counter(); m.fit(tf_iter=0, newton_iter=50, newton_eager=False); m._engine().sync()
        emit(json.dumps({"config": "1 Burgers graph-mode L-BFGS (Hager-Zhang)", "ms_per_iter": (time.perf_counter() - t0) * 20, "iterations": 50,
                         "evaluations": m.lbfgs_n_eval, "loss": m.lbfgs_history[-1]}))
        m2 = colloc("2 AC-SA [2,128x4,1] N_f=50000 Adam (self-adaptive)", ac(50000, True), 50000, "allen_cahn", 100, lam_cpu=True)
        t0 = time.perf_counter(); m2.fit(tf_iter=0, newton_iter=200); m2._engine().sync()
        emit(json.dumps({"config": "2 AC-SA eager L-BFGS", "ms_per_iter": (time.perf_counter() - t0) * 5, "iterations": 200, "P": 50049}))
        # config 3: Poisson
        D = DomainND(["x", "y"]); D.add("x", [0, 1.0], 11); D.add("y", [0, 1.0], 11); D.generate_collocation_points(100)
        fx = lambda y: -tf.math.sin(tf.constant(math.pi) * y) * tf.math.sin(tf.constant(math.pi))
        bcs = [FunctionDirichletBC(D, fun=[fx], var="x", target="upper", func_inputs=["y"], n_values=10), dirichletBC(D, val=0.0, var="x", target="upper"),
               FunctionDirichletBC(D, fun=[fx], var="y", target="upper", func_inputs=["x"], n_values=10), dirichletBC(D, val=0.0, var="y", target="lower")]
        m = CollocationSolverND(verbose=False); m.compile([2, 16, 16, 1], f_poisson, D, bcs)
        colloc("3 Poisson [2,16,16,1] N_f=100 Adam", m, 100, "poisson", 500)
        # config 4: CH discovery
        x = np.linspace(-1, 1, 256); t = np.linspace(0, 1, 100); X, T = np.meshgrid(x, t)
        Xs = np.hstack((X.flatten()[:, None], T.flatten()[:, None]))
        u = np.cos(math.pi * Xs[:, 0:1]) * np.exp(-Xs[:, 1:2])
        params = [tf.Variable(1e-4), tf.Variable(1e-4)]
        dm = DiscoveryModel(); dm.compile([2, 128, 128, 128, 128, 1], f_ch, [Xs[:, 0:1], Xs[:, 1:2]], tf.constant(u), params)
        dt, clocks = gpu_timed(lambda k: dm.train_op(k), 100, lambda: dm._engine().sync())
        cpu = None
        if not NO_CPU:
            from oracle import pinn_oracle as po
            import torch
            torch.set_num_threads(os.cpu_count() or 1)
            w = dm.u_model.get_flat().copy()
            ts = []
            for i in range(3):
                t0 = time.perf_counter()
                po.discovery_loss_and_grads([2, 128, 128, 128, 128, 1], "cahn_hilliard", Xs, u.astype(np.float32), w, np.array([1e-4, 1e-4], np.float32))
                if i:
                    ts.append(time.perf_counter() - t0)
            cpu = float(np.mean(ts))
        report("4 CH discovery [2,128x4,1] 25600 points Adam", 25600, dt, clocks, dm._engine().term_kernel(dm._t_res), cpu)
        colloc("5 AC [2,128x4,1] N_f=500000 Adam (examples/AC-dist-new.py's own size; 2M: bench.py)", ac(500000, False), 500000, "allen_cahn", 30)
        mb = ac(50000, True)
        mb.verbose = False
        dt, clocks = gpu_timed(lambda k: mb.fit(tf_iter=k, batch_sz=10000), 20, lambda: mb._engine().sync())
        report("2b AC-SA minibatch (5 batches of 10000 of 50000 per epoch)", 50000, dt, clocks, mb._engine().term_kernel(mb._terms[-1]["tid"]), None,
               {"note": "ms_per_step is one epoch = 5 Adam steps"})
        # predict on the AC example's 512 x 201 grid (models.py:274-284; examples/AC-SA.py:80)
        xs = np.linspace(-1, 1, 512); ts = np.linspace(0, 1, 201); Xg, Tg = np.meshgrid(xs, ts)
        X_star = np.hstack((Xg.flatten()[:, None], Tg.flatten()[:, None]))
        m2.predict(X_star)
        t0 = time.perf_counter()
        for _ in range(20):
            m2.predict(X_star)
        emit(json.dumps({"config": "predict, AC [2,128x4,1], 102912 grid points (host arrays in and out)", "ms_per_call": (time.perf_counter() - t0) * 50}))
