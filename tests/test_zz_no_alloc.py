"""GPU: tdq_loss_grad_static -- the entry point for hosts that cannot allocate device memory (a
plug-in inside another program's memory pool, a loop that must stay out of the driver's
allocator).  It is driven here exactly as such a host would drive it: every input and output is a
device buffer that already exists before the call (torch tensors stand in for the host program's
own pool), and the call itself must get by on the scratch array compiled into the library.

The result is checked against the ordinary handle path (tdq_loss_grad) on the same inputs, at the
tolerance the other GPU tests use (rel 1e-4), on two consecutive calls with different parameters
and point sets -- the second call finds the first call's leftovers in the shared scratch.

This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test.  It also passes on its own in a fresh process (the first call does the
per-device set-up)."""
import ctypes as C

import numpy as np
import pytest

from tensordiffeq_b200 import _lib
from tensordiffeq_b200 import trace as T
from tensordiffeq_b200.models import _trace, _integrand
from tensordiffeq_b200.synthetic import centred_lhs, glorot_normal_flat
from tests.test_jet_host import f_ac

pytestmark = pytest.mark.gpu
TOL = 1e-4


def rel(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def handle_path(ls, ops, slots, orders, w, X, lam):
    eng = _lib.Engine(ls)
    t = eng.add_term(2, orders, ops, slots["out"], -1, 0)
    eng.set_points(t, X, len(X))
    eng.set_lambda(t, lam)
    eng.set_params(w)
    loss, g, _ = eng.loss_grad()
    lg = eng.get_lambda_grad(t)
    eng.close()
    return loss, g, lg


def static_path(lib, torch, ls, ops, slots, orders, w, X, lam):
    dev = torch.device("cuda:0")
    P = w.size
    # buffers the "host program" owns before the call
    params_dev = torch.from_numpy(w).to(dev)
    x_cols_dev = torch.from_numpy(np.ascontiguousarray(X.T)).to(dev)        # [d][n]
    lam_dev = torch.from_numpy(lam).to(dev)
    gvec_dev = torch.full((P + 16,), float("nan"), dtype=torch.float32, device=dev)
    lam_grad_dev = torch.full((len(X),), float("nan"), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    spec, keep = _lib.make_spec(2, orders, ops, slots["out"], -1, 0)
    lsa = (C.c_int32 * len(ls))(*ls)
    err = C.create_string_buffer(512)
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.tdq_loss_grad_static(lsa, len(ls), C.byref(spec), params_dev.data_ptr(), None, 0, x_cols_dev.data_ptr(),
                                  None, lam_dev.data_ptr(), len(X), len(X), gvec_dev.data_ptr(), lam_grad_dev.data_ptr(),
                                  None, err, 512)
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    g = gvec_dev.cpu().numpy()
    return float(g[P + 8]), g[:P].copy(), lam_grad_dev.cpu().numpy(), g[P:], free_before - free_after


@pytest.mark.parametrize("ls", [[2, 128, 128, 128, 128, 1], [2, 20, 20, 20, 1]])
def test_static_entry_matches_handle_path_twice(ls):
    torch = pytest.importorskip("torch")
    lib = _lib.load()
    fe = T.as_sym(_trace(f_ac, 2))
    ops, slots, orders = T.lower({"out": _integrand(fe, False, "mse")}, 2)
    grew = []
    for call, (n, seed) in enumerate([(5003, 1), (7001, 2)]):     # ragged sizes, different data per call
        X = centred_lhs(n, [[-1, 1], [0, 1]], seed=100 + seed).astype(np.float32)
        w = glorot_normal_flat(ls, seed=seed) + np.float32(0.01 * seed)
        lam = np.random.RandomState(seed).uniform(0.2, 1.5, n).astype(np.float32)
        loss_h, g_h, lg_h = handle_path(ls, ops, slots, orders, w, X, lam)
        loss_s, g_s, lg_s, tail, used = static_path(lib, torch, ls, ops, slots, orders, w, X, lam)
        grew.append(used)
        assert abs(loss_s - loss_h) <= TOL * abs(loss_h), (call, loss_s, loss_h)
        assert rel(g_s, g_h) < TOL, call
        assert rel(lg_s, lg_h) < TOL, call
        assert np.all(tail[:8] == 0) and np.all(tail[9:] == 0), "output tail must be cleared"
    # Free device memory around the calls.  The very first use on a device may show the one-time cost of
    # the library's device image being brought in (CUDA loads modules lazily; the image carries the
    # 32 MiB scratch array) -- that is the documented standing cost, not an allocation by the call.
    # A call made after that must leave free memory untouched.
    assert grew[0] <= (32 << 20) + (8 << 20), grew
    assert grew[1] <= (2 << 20), grew          # at most the granularity cudaMemGetInfo reports in: no allocation


def test_static_entry_paired_term_and_coefficients():
    """The two argument shapes the first test leaves out, each against the handle path: a paired (periodic) term,
    whose loss couples consecutive points, and a program that reads trainable PDE coefficients from vars_dev (the
    Cahn-Hilliard residual; d loss / d coefficient comes back in the output's tail).  A program that reads a
    coefficient while vars_dev is NULL must be refused instead of reading something else."""
    torch = pytest.importorskip("torch")
    from tensordiffeq_b200 import tf
    from tests.test_jet_host import f_ch
    lib = _lib.load()
    dev = torch.device("cuda:0")
    err = C.create_string_buffer(512)

    def run_static(ls, spec, w, X, n_mean, vars_np=None, pass_vars=True):
        P = w.size
        params_dev = torch.from_numpy(w).to(dev)
        x_cols_dev = torch.from_numpy(np.ascontiguousarray(X.T)).to(dev)
        vars_dev = torch.from_numpy(vars_np).to(dev) if vars_np is not None else None
        gvec_dev = torch.full((P + 16,), float("nan"), dtype=torch.float32, device=dev)
        lsa = (C.c_int32 * len(ls))(*ls)
        rc = lib.tdq_loss_grad_static(lsa, len(ls), C.byref(spec), params_dev.data_ptr(),
                                      vars_dev.data_ptr() if (vars_dev is not None and pass_vars) else None,
                                      0 if vars_np is None else len(vars_np), x_cols_dev.data_ptr(), None, None, len(X), n_mean,
                                      gvec_dev.data_ptr(), None, None, err, 512)
        torch.cuda.synchronize()
        return rc, gvec_dev.cpu().numpy()

    # ---- paired term: mean over couples of (u(upper) - u(lower))^2
    ls = [2, 128, 128, 128, 128, 1]
    ops, slots, orders = T.lower({"out": T.jet((0, 0))}, 2)
    rs = np.random.RandomState(5)
    t_lv = rs.uniform(0, 1, 301).astype(np.float32)
    X = np.empty((602, 2), np.float32)
    X[0::2, 0], X[1::2, 0] = 1.0, -1.0
    X[0::2, 1] = X[1::2, 1] = t_lv
    w = glorot_normal_flat(ls, seed=3) + np.float32(0.02)
    eng = _lib.Engine(ls)
    t = eng.add_term(2, orders, ops, slots["out"], -1, 1)
    eng.set_points(t, X, len(X) // 2)
    eng.set_params(w)
    loss_h, g_h, _ = eng.loss_grad()
    eng.close()
    spec, keep = _lib.make_spec(2, orders, ops, slots["out"], -1, 1)
    rc, g = run_static(ls, spec, w, X, len(X) // 2)
    assert rc == 0, err.value.decode()
    P = w.size
    assert abs(float(g[P + 8]) - loss_h) <= TOL * abs(loss_h)
    assert rel(g[:P], g_h) < TOL

    # ---- trainable coefficients
    ls = [2, 32, 32, 32, 1]
    vs = [tf.Variable(0.3, dtype=tf.float32), tf.Variable(-0.2, dtype=tf.float32)]
    for i, v in enumerate(vs):
        v._sym = T.var(i)
    fe = T.as_sym(_trace(f_ch, 2, extra_first=vs))
    for v in vs:
        v._sym = None
    ops, slots, orders = T.lower({"out": _integrand(fe)}, 2)
    X = centred_lhs(1203, [[-1, 1], [0, 1]], seed=77).astype(np.float32)
    w = glorot_normal_flat(ls, seed=4) + np.float32(0.03)
    vars_np = np.array([0.3, -0.2], np.float32)
    eng = _lib.Engine(ls)
    t = eng.add_term(2, orders, ops, slots["out"], -1, 0)
    eng.set_points(t, X, len(X))
    eng.set_vars(vars_np)
    eng.set_params(w)
    loss_h, g_h, vg_h = eng.loss_grad(2)
    eng.close()
    spec, keep = _lib.make_spec(2, orders, ops, slots["out"], -1, 0)
    rc, g = run_static(ls, spec, w, X, len(X), vars_np)
    assert rc == 0, err.value.decode()
    P = w.size
    assert abs(float(g[P + 8]) - loss_h) <= 2 * TOL * abs(loss_h)
    assert rel(g[:P], g_h) < 2 * TOL                 # 4th-order jets: the looser bound the discovery tests use
    assert rel(g[P:P + 2], vg_h) < 2 * TOL
    rc, _ = run_static(ls, spec, w, X, len(X), vars_np, pass_vars=False)
    assert rc != 0 and b"vars_dev" in err.value


def test_static_entry_limits():
    """Shapes beyond what the built-in scratch was sized for (hidden width > 128) come back as an
    error text, not a crash; the largest shape it does serve (12 tanh layers of 128, 4th-order jets)
    still runs, on however many CTAs fit into the array."""
    torch = pytest.importorskip("torch")
    lib = _lib.load()
    ls = [2] + [128] * 12 + [1]
    d = 2

    def f4(u_model, x, t):
        from tensordiffeq_b200 import tf
        u = u_model(tf.concat([x, t], 1))
        g = u
        for _ in range(4):
            g = tf.gradients(g, x)[0]
        return g + tf.gradients(u, t)[0]
    fe = T.as_sym(_trace(f4, d))
    ops, slots, orders = T.lower({"out": _integrand(fe)}, d)
    spec, keep = _lib.make_spec(d, orders, ops, slots["out"], -1, 0)
    P = sum(ls[i - 1] * ls[i] + ls[i] for i in range(1, len(ls)))
    dev = torch.device("cuda:0")
    params = torch.zeros(P, device=dev)
    X = torch.zeros(2 * 64, device=dev)
    out = torch.zeros(P + 16, device=dev)
    err = C.create_string_buffer(512)
    lsa = (C.c_int32 * len(ls))(*ls)
    too_wide = (C.c_int32 * 3)(2, 300, 1)
    rc = lib.tdq_loss_grad_static(too_wide, 3, C.byref(spec), params.data_ptr(), None, 0, X.data_ptr(), None, None, 64, 64,
                                  out.data_ptr(), None, None, err, 512)
    assert rc != 0 and b"width" in err.value
    rc = lib.tdq_loss_grad_static(lsa, len(ls), C.byref(spec), params.data_ptr(), None, 0, X.data_ptr(), None, None, 64, 64,
                                  out.data_ptr(), None, None, err, 512)
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
