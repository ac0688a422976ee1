#!/usr/bin/env python
"""Generate tests/golden/*.npz by executing the UNMODIFIED reference package
(/root/reference/tensordiffeq: models.py, fit.py, optimizers.py, boundaries.py,
domains.py, utils.py) on top of oracle/ref_shims (a torch-backed stand-in for
the TensorFlow API, because TensorFlow cannot be installed in this image).

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
The .npz files it writes are committed; nothing on the GPU box runs this.

Every fixture stores the INPUTS handed to the reference (flat weights in the
utils.py:7-29 layout, collocation points, BC point sets / targets, lambdas) and
the OUTPUTS the reference produced (loss, gradients, post-step state, ...).
"""
import math
import os
import sys

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle", "ref_shims"))
sys.path.insert(0, "/root/reference")

import numpy as np  # noqa: E402
import torch  # noqa: E402
import tensorflow as tf  # noqa: E402  (the shim)
import tensordiffeq as tdq  # noqa: E402  (the reference)
from tensordiffeq.boundaries import IC, dirichletBC, periodicBC, FunctionDirichletBC, DomainND  # noqa: E402
from tensordiffeq.models import CollocationSolverND, DiscoveryModel  # noqa: E402
from tensordiffeq.fit import train_op_inner  # noqa: E402
from tensordiffeq.optimizers import eager_lbfgs, Struct  # noqa: E402
from tensordiffeq.utils import get_weights, set_weights, constant  # noqa: E402

torch.set_num_threads(8)


def glorot_flat(layer_sizes, seed):
    """Flat parameter vector, reference layout (utils.py:19-29): per layer the
    [in,out] kernel row-major, then the bias."""
    rng = np.random.RandomState(seed)
    out = []
    for i in range(1, len(layer_sizes)):
        fi, fo = layer_sizes[i - 1], layer_sizes[i]
        std = math.sqrt(2.0 / (fi + fo)) / 0.87962566103423978
        w = rng.standard_normal((fi, fo))
        bad = np.abs(w) > 2
        while bad.any():
            w[bad] = rng.standard_normal(int(bad.sum()))
            bad = np.abs(w) > 2
        out.append((w * std).astype(np.float32).ravel())
        out.append((0.1 * rng.standard_normal(fo)).astype(np.float32))  # non-zero biases exercise db
    return np.concatenate(out)


def load_flat(model, w):
    set_weights(model.u_model, tf.convert_to_tensor(w), model.sizes_w, model.sizes_b)


def flat_grads(grads, n_net):
    return np.concatenate([g.numpy().ravel() for g in grads[:n_net]])


def bc_dump(bcs, out):
    """Record what the reference built for each BC (point sets are partly random:
    SURVEY Q3), so the oracle / CUDA path receive identical inputs."""
    kinds = []
    for i, bc in enumerate(bcs):
        if bc.isPeriodic:
            kinds.append("periodic")
            out["bc%d_upper" % i] = np.asarray(bc.upper, np.float64)  # [n_var, d, n, 1]
            out["bc%d_lower" % i] = np.asarray(bc.lower, np.float64)
        elif bc.isInit:
            kinds.append("ic")
            out["bc%d_input" % i] = np.asarray(bc.input, np.float64)
            out["bc%d_val" % i] = bc.val.numpy()
        elif bc.isNeumann:
            kinds.append("neumann")
        else:
            kinds.append("dirichlet")
            out["bc%d_input" % i] = np.asarray(bc.input, np.float64)
            v = bc.val
            out["bc%d_val" % i] = v.numpy() if hasattr(v, "numpy") else np.asarray(v, np.float64)
    out["bc_kinds"] = np.array(kinds)


def ac_problem(N_x, N_t, N_f, seed):
    np.random.seed(seed)
    Domain = DomainND(["x", "t"], time_var="t")
    Domain.add("x", [-1.0, 1.0], N_x)
    Domain.add("t", [0.0, 1.0], N_t)
    Domain.generate_collocation_points(N_f)

    def func_ic(x):
        return x ** 2 * np.cos(math.pi * x)

    def deriv_model(u_model, x, t):  # examples/AC-SA.py:21-27
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)[0]
        return u, u_x

    init = IC(Domain, [func_ic], var=[["x"]])
    x_periodic = periodicBC(Domain, ["x"], [deriv_model])

    def f_model(u_model, x, t):  # examples/AC-SA.py:36-44
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)
        u_xx = tf.gradients(u_x, x)
        u_t = tf.gradients(u, t)
        c1 = tdq.utils.constant(.0001)
        c2 = tdq.utils.constant(5.0)
        f_u = u_t - c1 * u_xx + c2 * u * u * u - c2 * u
        return f_u

    return Domain, [init, x_periodic], f_model


def run_collocation(name, model, n_adam, n_lbfgs, out, X_star=None, snapshots=()):
    """Drive the reference exactly as fit.py:17-51 does, but keep every step's loss (and the flat weight
    vector after the steps listed in `snapshots`)."""
    n_net = len(model.u_model.trainable_variables)
    model.batch_sz = model.X_f_len[0]
    model.n_batches = 1
    # --- one loss/grad evaluation (models.py:195-199 via fit.py:92-96)
    model.variables = model.u_model.trainable_variables
    lam = model.lambdas if (model.isAdaptive and model.lambdas) else []
    model.variables.extend(lam)
    loss, grads = model.grad()
    out["loss0"] = np.float32(loss.numpy())
    out["grad0"] = flat_grads(grads, n_net)
    for i, g in enumerate(grads[n_net:]):
        out["lamgrad0_%d" % i] = g.numpy()
    # residual field at the collocation points (models.py:157)
    f0 = model.f_model(model.u_model, *model.X_f_in)
    out["f0"] = f0.numpy().reshape(-1)
    if X_star is not None:
        u_s, f_s = model.predict(X_star)  # models.py:274-284
        out["X_star"] = X_star
        out["u_star"] = u_s
        out["f_star"] = f_s
    # --- Adam (fit.py:23,28-31)
    model.grad = tf.function(model.grad)
    train_op_fn = train_op_inner(model)
    hist = []
    for it in range(n_adam):
        hist.append(np.float32(train_op_fn(model).numpy()))
        if it + 1 in snapshots:
            out["w_after_%d" % (it + 1)] = get_weights(model.u_model).numpy()
    out["adam_loss_hist"] = np.array(hist, np.float32)
    out["w_after_adam"] = get_weights(model.u_model).numpy()
    for i, l in enumerate(lam):
        out["lam_after_adam_%d" % i] = l.numpy()
    # --- eager L-BFGS (fit.py:43-51)
    if n_lbfgs:
        loss_and_flat_grad = model.get_loss_and_flat_grad()
        x, f_hist, n_eval = eager_lbfgs(loss_and_flat_grad, get_weights(model.u_model), Struct(),
                                        maxIter=n_lbfgs, learningRate=0.8)
        out["lbfgs_f_hist"] = np.array([np.float32(f.numpy()) for f in f_hist], np.float32)
        out["w_after_lbfgs"] = x.numpy()
        out["lbfgs_n_eval"] = np.int32(n_eval)
    print("%-14s loss0=%.6e  adam->%.6e%s" % (
        name, out["loss0"], hist[-1] if hist else float("nan"),
        "  lbfgs->%.6e" % out["lbfgs_f_hist"][-1] if n_lbfgs else ""))


def case_ac_sa(layer_sizes, N_x, N_t, N_f, adaptive, n_adam, n_lbfgs, seed, name, g=None, snapshots=()):
    Domain, BCs, f_model = ac_problem(N_x, N_t, N_f, seed)
    torch.manual_seed(seed)
    out = {"layer_sizes": np.array(layer_sizes, np.int32)}
    if adaptive:
        dict_adaptive = {"residual": [True], "BCs": [True, False]}
        lam_res = tf.random.uniform([N_f, 1])
        lam_ic = 100 * tf.random.uniform([N_x, 1])
        init_weights = {"residual": [lam_res], "BCs": [lam_ic, None]}
        out["lam0_0"] = lam_res.numpy()
        out["lam0_1"] = lam_ic.numpy()
    else:
        dict_adaptive = init_weights = None
    model = CollocationSolverND(verbose=False)
    model.compile(layer_sizes, f_model, Domain, BCs, isAdaptive=adaptive,
                  dict_adaptive=dict_adaptive, init_weights=init_weights, g=g)
    w0 = glorot_flat(layer_sizes, seed)
    load_flat(model, w0)
    out["w0"] = w0
    out["X_f"] = np.asarray(Domain.X_f, np.float64)
    bc_dump(BCs, out)
    xs = np.linspace(-1, 1, 17)
    ts = np.linspace(0, 1, 9)
    Xg, Tg = np.meshgrid(xs, ts)
    X_star = np.hstack((Xg.flatten()[:, None], Tg.flatten()[:, None]))
    run_collocation(name, model, n_adam, n_lbfgs, out, X_star, snapshots)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


def case_ac_two_res(name, adaptive_flags, seed):
    """A callback that returns a TUPLE of residuals (models.py:160-188), with per-residual adaptivity flags.
    Both residuals live on the same collocation points; with two adaptive residuals the reference weights
    both with lambdas_map['residual'][0] (models.py:169)."""
    layer_sizes = [2, 24, 24, 24, 1]
    N_x, N_t, N_f = 64, 30, 1200
    Domain, BCs, f_ac = ac_problem(N_x, N_t, N_f, seed)
    torch.manual_seed(seed)

    def f_model(u_model, x, t):
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)[0]
        u_xx = tf.gradients(u_x, x)[0]
        u_t = tf.gradients(u, t)[0]
        u_tt = tf.gradients(u_t, t)[0]
        f_1 = u_t - tdq.utils.constant(.0001) * u_xx + tdq.utils.constant(5.0) * u * u * u - tdq.utils.constant(5.0) * u
        f_2 = u_tt - u * u_x + tdq.utils.constant(0.25) * x
        return f_1, f_2

    dict_adaptive = {"residual": list(adaptive_flags), "BCs": [True, False]}
    lam_res = tf.random.uniform([N_f, 1])
    lam_ic = 100 * tf.random.uniform([N_x, 1])
    init_weights = {"residual": [lam_res], "BCs": [lam_ic, None]}
    out = {"layer_sizes": np.array(layer_sizes, np.int32), "lam0_0": lam_res.numpy(), "lam0_1": lam_ic.numpy(),
           "res_adaptive": np.array(adaptive_flags, np.int32)}
    model = CollocationSolverND(verbose=False)
    model.compile(layer_sizes, f_model, Domain, BCs, isAdaptive=True, dict_adaptive=dict_adaptive, init_weights=init_weights)
    w0 = glorot_flat(layer_sizes, seed)
    load_flat(model, w0)
    out["w0"] = w0
    out["X_f"] = np.asarray(Domain.X_f, np.float64)
    bc_dump(BCs, out)
    n_net = len(model.u_model.trainable_variables)
    model.batch_sz = model.X_f_len[0]
    model.n_batches = 1
    model.variables = model.u_model.trainable_variables
    model.variables.extend(model.lambdas)
    loss, grads = model.grad()
    out["loss0"] = np.float32(loss.numpy())
    out["grad0"] = flat_grads(grads, n_net)
    for i, g in enumerate(grads[n_net:]):
        out["lamgrad0_%d" % i] = g.numpy()
    fs = model.f_model(model.u_model, *model.X_f_in)
    out["f0"] = fs[0].numpy().reshape(-1)
    out["f0_b"] = fs[1].numpy().reshape(-1)
    model.grad = tf.function(model.grad)
    train_op_fn = train_op_inner(model)
    hist = [np.float32(train_op_fn(model).numpy()) for _ in range(8)]
    out["adam_loss_hist"] = np.array(hist, np.float32)
    out["w_after_adam"] = get_weights(model.u_model).numpy()
    for i, l in enumerate(model.lambdas):
        out["lam_after_adam_%d" % i] = l.numpy()
    print("%-14s loss0=%.6e  adam->%.6e" % (name, out["loss0"], hist[-1]))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


def case_burgers(name, seed=3, n_adam=10, n_lbfgs=8, snapshots=()):
    np.random.seed(seed)
    layer_sizes = [2, 20, 20, 20, 20, 20, 20, 20, 20, 1]
    Domain = DomainND(["x", "t"], time_var="t")
    Domain.add("x", [-1.0, 1.0], 64)
    Domain.add("t", [0.0, 1.0], 25)
    Domain.generate_collocation_points(1500)

    def func_ic(x):
        return -np.sin(x * math.pi)

    init = IC(Domain, [func_ic], var=[["x"]])
    upper_x = dirichletBC(Domain, val=0.0, var="x", target="upper")
    lower_x = dirichletBC(Domain, val=0.0, var="x", target="lower")
    BCs = [init, upper_x, lower_x]

    def f_model(u_model, x, t):  # examples/burgers-new.py:26-32
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)
        u_xx = tf.gradients(u_x, x)
        u_t = tf.gradients(u, t)
        f_u = u_t + u * u_x - (0.01 / tf.constant(math.pi)) * u_xx
        return f_u

    model = CollocationSolverND(verbose=False)
    model.compile(layer_sizes, f_model, Domain, BCs)
    out = {"layer_sizes": np.array(layer_sizes, np.int32)}
    w0 = glorot_flat(layer_sizes, seed)
    load_flat(model, w0)
    out["w0"] = w0
    out["X_f"] = np.asarray(Domain.X_f, np.float64)
    bc_dump(BCs, out)
    xs = np.linspace(-1, 1, 13)
    ts = np.linspace(0, 1, 7)
    Xg, Tg = np.meshgrid(xs, ts)
    X_star = np.hstack((Xg.flatten()[:, None], Tg.flatten()[:, None]))
    run_collocation(name, model, n_adam, n_lbfgs, out, X_star, snapshots)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


def case_poisson(name, seed=5):
    from tensorflow.math import sin
    np.random.seed(seed)
    Domain = DomainND(["x", "y"])
    Domain.add("x", [0, 1.0], 11)
    Domain.add("y", [0, 1.0], 11)
    Domain.generate_collocation_points(100)

    def f_model(u_model, x, y):  # examples/steady-state-poisson.py:19-37
        u = u_model(tf.concat([x, y], 1))
        u_x = tf.gradients(u, x)[0]
        u_y = tf.gradients(u, y)[0]
        u_xx = tf.gradients(u_x, x)[0]
        u_yy = tf.gradients(u_y, y)[0]
        a1 = constant(1.0)
        a2 = constant(1.0)
        pi = constant(math.pi)
        forcing = - sin(a1 * pi * x) * sin(a2 * pi * y)
        f_u = u_xx + u_yy - forcing
        return f_u

    def func_upper_x(y):
        return -sin(constant(math.pi) * y) * sin(constant(math.pi))

    def func_upper_y(x):
        return -sin(constant(math.pi) * x) * sin(constant(math.pi))

    lower_x = dirichletBC(Domain, val=0.0, var="x", target="upper")
    upper_x = FunctionDirichletBC(Domain, fun=[func_upper_x], var="x", target="upper", func_inputs=["y"], n_values=10)
    upper_y = FunctionDirichletBC(Domain, fun=[func_upper_y], var="y", target="upper", func_inputs=["x"], n_values=10)
    lower_y = dirichletBC(Domain, val=0.0, var="y", target="lower")
    BCs = [upper_x, lower_x, upper_y, lower_y]
    layer_sizes = [2, 16, 16, 1]
    model = CollocationSolverND(verbose=False)
    model.compile(layer_sizes, f_model, Domain, BCs)
    model.tf_optimizer = tf.keras.optimizers.Adam(lr=.005)  # examples/steady-state-poisson.py:59
    out = {"layer_sizes": np.array(layer_sizes, np.int32)}
    w0 = glorot_flat(layer_sizes, seed)
    load_flat(model, w0)
    out["w0"] = w0
    out["X_f"] = np.asarray(Domain.X_f, np.float64)
    bc_dump(BCs, out)
    xs = np.linspace(0, 1, 6)
    Xg, Yg = np.meshgrid(xs, xs)
    X_star = np.hstack((Xg.flatten()[:, None], Yg.flatten()[:, None]))
    run_collocation(name, model, 12, 0, out, X_star)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


def case_mixed(name, layer_sizes, N_f, seed):
    """A residual with the mixed partial u_xy = tf.gradients(tf.gradients(u, x), y) (as examples/testing.py:23-31 takes
    it) on the unit square with two Dirichlet edges."""
    from tensorflow.math import sin
    np.random.seed(seed)
    Domain = DomainND(["x", "y"])
    Domain.add("x", [0, 1.0], 11)
    Domain.add("y", [0, 1.0], 11)
    Domain.generate_collocation_points(N_f)

    def f_model(u_model, x, y):
        u = u_model(tf.concat([x, y], 1))
        u_x = tf.gradients(u, x)[0]
        u_y = tf.gradients(u, y)[0]
        u_xx = tf.gradients(u_x, x)[0]
        u_yy = tf.gradients(u_y, y)[0]
        u_xy = tf.gradients(u_x, y)[0]
        return u_xx + constant(0.5) * u_xy * u + u_yy + u_xy * u_xy - sin(constant(math.pi) * x) * sin(constant(math.pi) * y)

    BCs = [dirichletBC(Domain, val=0.0, var="x", target="upper"), dirichletBC(Domain, val=0.0, var="y", target="lower")]
    model = CollocationSolverND(verbose=False)
    model.compile(layer_sizes, f_model, Domain, BCs)
    out = {"layer_sizes": np.array(layer_sizes, np.int32)}
    w0 = glorot_flat(layer_sizes, seed)
    load_flat(model, w0)
    out["w0"] = w0
    out["X_f"] = np.asarray(Domain.X_f, np.float64)
    bc_dump(BCs, out)
    xs = np.linspace(0, 1, 6)
    Xg, Yg = np.meshgrid(xs, xs)
    X_star = np.hstack((Xg.flatten()[:, None], Yg.flatten()[:, None]))
    run_collocation(name, model, 6, 0, out, X_star)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


def case_ch_discovery(name, seed=7, with_col_weights=False, layer_sizes=(2, 32, 32, 32, 1), nx=24, nt=10, n_adam=8):
    rng = np.random.RandomState(seed)
    torch.manual_seed(seed)
    layer_sizes = list(layer_sizes)
    params = [tf.Variable(0.0001, dtype=tf.float32), tf.Variable(0.0001, dtype=tf.float32)]

    def f_model(u_model, var, x, t):  # examples/CH-discovery.py:20-35
        g1 = var[0]
        g2 = var[1]
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)
        u_t = tf.gradients(u, t)
        u_xx = tf.gradients(u_x, x)
        tmp = g1 * (u ** 3 - u) - g2 * u_xx
        tmp_x = tf.gradients(tmp, x)[0]
        tmp_xx = tf.gradients(tmp_x, x)[0]
        f_u = u_t - tmp_xx
        return f_u

    x = np.linspace(-1, 1, nx)
    t = np.linspace(0, 1, nt)
    Xg, Tg = np.meshgrid(x, t)
    X_star = np.hstack((Xg.flatten()[:, None], Tg.flatten()[:, None]))
    u_np = (np.cos(math.pi * X_star[:, 0:1]) * np.exp(-X_star[:, 1:2]) + 0.05 * rng.standard_normal((nx * nt, 1)))
    u_star = tdq.utils.tensor(u_np)
    # the example passes float64 numpy columns (CH-discovery.py:51-55): a float32 net then makes
    # tf.gradients(u, x) float64; the cast at models.py:321 brings f back to float32.
    X = [X_star[:, 0:1], X_star[:, 1:2]]
    X = [tf.cast(c, tf.float32) for c in X]
    out = {"layer_sizes": np.array(layer_sizes, np.int32), "X": X_star, "u": u_np.astype(np.float32)}
    col_weights = None
    if with_col_weights:
        col_weights = tf.Variable(tf.random.uniform([nx * nt, 1]))
        out["col_weights0"] = col_weights.numpy()
    model = DiscoveryModel()
    model.compile(layer_sizes, f_model, X, u_star, params, col_weights=col_weights)
    w0 = glorot_flat(layer_sizes, seed)
    from tensordiffeq.utils import get_sizes
    sw, sb = get_sizes(layer_sizes)
    set_weights(model.u_model, tf.convert_to_tensor(w0), sw, sb)
    out["w0"] = w0
    out["vars0"] = np.array([p.numpy() for p in params], np.float32)
    # one loss/grad evaluation, same variable order as train_op (models.py:332-344)
    n_net = len(model.u_model.trainable_variables)
    model.variables = model.u_model.trainable_variables
    if with_col_weights:
        model.variables.extend([col_weights])
    model.variables.extend(model.vars)
    loss, grads = model.grad()
    out["loss0"] = np.float32(loss.numpy())
    out["grad0"] = flat_grads(grads, n_net)
    out["vargrad0"] = np.array([g.numpy() for g in grads[-2:]], np.float32)
    if with_col_weights:
        out["colgrad0"] = grads[n_net].numpy()
    hist = []
    for _ in range(n_adam):
        hist.append(np.float32(model.train_op().numpy()))
    out["adam_loss_hist"] = np.array(hist, np.float32)
    out["w_after_adam"] = get_weights(model.u_model).numpy()
    out["vars_after_adam"] = np.array([p.numpy() for p in params], np.float32)
    if with_col_weights:
        out["col_weights_after_adam"] = col_weights.numpy()
    print("%-14s loss0=%.6e adam->%.6e vars=%s" % (name, out["loss0"], hist[-1], out["vars_after_adam"]))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


AC128 = [2, 128, 128, 128, 128, 1]
CASES = {
    "ac_sa_small": lambda: case_ac_sa([2, 21, 21, 21, 21, 1], 100, 50, 2000, True, 10, 8, 11, "ac_sa_small"),
    "ac_sa_g": lambda: case_ac_sa([2, 21, 21, 21, 21, 1], 64, 30, 1000, True, 6, 0, 12, "ac_sa_g", g=lambda lam: lam ** 2 + 1.0),
    "ac_sa_128": lambda: case_ac_sa(AC128, 512, 201, 1024, True, 3, 0, 13, "ac_sa_128"),
    "ac_dist_128": lambda: case_ac_sa(AC128, 512, 201, 1000, False, 3, 3, 14, "ac_dist_128"),
    "burgers": lambda: case_burgers("burgers"),
    "poisson": lambda: case_poisson("poisson"),
    "ch_discovery": lambda: case_ch_discovery("ch_discovery"),
    "ch_discovery_sa": lambda: case_ch_discovery("ch_discovery_sa", seed=8, with_col_weights=True),
    # BASELINE config 4 at its own net shape (examples/CH-discovery.py:63), with and without col_weights
    "ch_discovery_128": lambda: case_ch_discovery("ch_discovery_128", seed=9, layer_sizes=AC128, nx=32, nt=12, n_adam=6),
    "ch_discovery_128_sa": lambda: case_ch_discovery("ch_discovery_128_sa", seed=10, with_col_weights=True, layer_sizes=AC128, nx=32, nt=12, n_adam=6),
    # f_model returning a tuple of two residuals: one adaptive, then both (sharing lambda 0 as the reference does)
    "ac_two_res": lambda: case_ac_two_res("ac_two_res", (True, False), 21),
    "ac_two_res_both": lambda: case_ac_two_res("ac_two_res_both", (True, True), 22),
    # a residual that reads the mixed partial u_xy (small net: FMA kernel; 128-wide: 16-point tensor-core kernel)
    "mixed_xy": lambda: case_mixed("mixed_xy", [2, 24, 24, 1], 150, 31),
    "mixed_xy_128": lambda: case_mixed("mixed_xy_128", [2, 128, 128, 128, 1], 200, 32),
    # 1000 Adam steps (north_star's trajectory band): same problems as ac_sa_small / burgers / ac_sa_128 with the
    # whole loss history and the weights after 100 and 1000 steps
    "ac_sa_small_long": lambda: case_ac_sa([2, 21, 21, 21, 21, 1], 100, 50, 2000, True, 1000, 0, 11, "ac_sa_small_long", snapshots=(100,)),
    "burgers_long": lambda: case_burgers("burgers_long", n_adam=1000, n_lbfgs=0, snapshots=(100,)),
    "ac_sa_128_long": lambda: case_ac_sa(AC128, 512, 201, 1024, True, 1000, 0, 13, "ac_sa_128_long", snapshots=(100,)),
    "ac_dist_128_long": lambda: case_ac_sa(AC128, 512, 201, 1000, False, 1000, 0, 14, "ac_dist_128_long", snapshots=(100,)),
}

if __name__ == "__main__":
    for case in (sys.argv[1:] or list(CASES)):
        CASES[case]()
