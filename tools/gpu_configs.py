"""Step time of every BASELINE.json config through the public API (developer tool; the bench line is bench.py)."""
import math, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tensordiffeq_b200 as tdq
from tensordiffeq_b200 import tf
from tensordiffeq_b200.boundaries import *
from tensordiffeq_b200.models import CollocationSolverND, DiscoveryModel
from tests.test_jet_host import f_ac, f_burgers, f_poisson, f_ch
from tests.gpu_util import deriv_model


def timed(model, n, label, npts, **kw):
    model.verbose = False
    model.fit(tf_iter=20, **kw)          # warm-up (builds the engine)
    eng = model._engine()
    eng.sync()
    t0 = time.perf_counter()
    model.fit(tf_iter=n, **kw)
    eng.sync()
    dt = (time.perf_counter() - t0) / n
    print("%-34s %8.3f ms/step  %10.2f Mpts/s   loss %.4e" % (label, dt * 1e3, npts / dt / 1e6, model.loss_history[-1]), flush=True)


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
    timed(m, 200, "1 Burgers [2,20x8,1] N_f=10k", 10000)
    t0 = time.perf_counter(); m.fit(tf_iter=0, newton_iter=100); m._engine().sync()
    print("%-34s %8.3f ms/iter (100 L-BFGS iterations, loss %.4e)" % ("1 Burgers L-BFGS", (time.perf_counter() - t0) * 10, m.lbfgs_history[-1]), flush=True)
    m2 = ac(50000, True)
    timed(m2, 100, "2 AC-SA [2,128x4,1] N_f=50k", 50000)
    t0 = time.perf_counter(); m2.fit(tf_iter=0, newton_iter=200); m2._engine().sync()
    print("%-34s %8.3f ms/iter (200 L-BFGS iterations, P = 50 049, loss %.4e)" % ("2 AC-SA L-BFGS", (time.perf_counter() - t0) * 5, m2.lbfgs_history[-1]), flush=True)
    # config 3: Poisson
    D = DomainND(["x", "y"]); D.add("x", [0, 1.0], 11); D.add("y", [0, 1.0], 11); D.generate_collocation_points(100)
    fx = lambda y: -tf.math.sin(tf.constant(math.pi) * y) * tf.math.sin(tf.constant(math.pi))
    bcs = [FunctionDirichletBC(D, fun=[fx], var="x", target="upper", func_inputs=["y"], n_values=10), dirichletBC(D, val=0.0, var="x", target="upper"),
           FunctionDirichletBC(D, fun=[fx], var="y", target="upper", func_inputs=["x"], n_values=10), dirichletBC(D, val=0.0, var="y", target="lower")]
    m = CollocationSolverND(verbose=False); m.compile([2, 16, 16, 1], f_poisson, D, bcs)
    timed(m, 500, "3 Poisson [2,16,16,1] N_f=100", 100)
    # config 4: CH discovery
    x = np.linspace(-1, 1, 256); t = np.linspace(0, 1, 100); X, T = np.meshgrid(x, t)
    Xs = np.hstack((X.flatten()[:, None], T.flatten()[:, None]))
    u = np.cos(math.pi * Xs[:, 0:1]) * np.exp(-Xs[:, 1:2])
    params = [tf.Variable(1e-4), tf.Variable(1e-4)]
    dm = DiscoveryModel(); dm.compile([2, 128, 128, 128, 128, 1], f_ch, [Xs[:, 0:1], Xs[:, 1:2]], tf.constant(u), params)
    dm.train_op(20); dm._engine().sync(); t0 = time.perf_counter(); dm.train_op(100); dm._engine().sync(); dt = (time.perf_counter() - t0) / 100
    print("%-34s %8.3f ms/step  %10.2f Mpts/s" % ("4 CH discovery [2,128x4,1] 25.6k", dt * 1e3, 25600 / dt / 1e6), flush=True)
    timed(ac(500000, False), 30, "5 AC [2,128x4,1] N_f=500k", 500000)
    timed(ac(50000, True), 20, "2b AC-SA minibatch 10k of 50k", 50000, batch_sz=10000)
