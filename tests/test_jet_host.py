"""CPU: the device math headers (csrc/jet_math.cuh, csrc/vm.cuh), compiled for the host by
tests/harness/jet_host.cu, and the residual tracer (tensordiffeq_b200/trace.py, tf.py), against
nested torch autograd on the oracle's restatement of the reference callbacks."""
import math
import os
import struct
import subprocess

import numpy as np
import pytest
import torch

from oracle import pinn_oracle as po
from tensordiffeq_b200 import _lib, tf
from tensordiffeq_b200 import trace as T
from tensordiffeq_b200.models import _trace, _integrand

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HARNESS = os.path.join(ROOT, "build", "jet_host")


@pytest.fixture(scope="module")
def harness():
    src = os.path.join(ROOT, "tests", "harness", "jet_host.cu")
    deps = [src] + [os.path.join(ROOT, "tensordiffeq_b200", "csrc", f) for f in ("jet_math.cuh", "f32x2.cuh", "vm.cuh", "vm_poly.h", "vm_poly.cuh")]
    if not os.path.exists(HARNESS) or os.path.getmtime(HARNESS) < max(os.path.getmtime(f) for f in deps):
        os.makedirs(os.path.dirname(HARNESS), exist_ok=True)
        subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-o", HARNESS, src])
    return HARNESS


def run_harness(exe, cfg, ls, w, X, target, lam, vars_, ops, out_slot, C):
    n, d = X.shape
    buf = struct.pack("5i", cfg[0], cfg[1], cfg[2], d, len(ls)) + struct.pack("%di" % len(ls), *ls)
    buf += struct.pack("3i", n, len(ops), out_slot)
    buf += np.asarray(w, np.float32).tobytes() + np.ascontiguousarray(X, np.float32).tobytes()
    buf += np.asarray(target, np.float32).tobytes() + np.asarray(lam, np.float32).tobytes()
    v8 = np.zeros(8, np.float32)
    v8[:len(vars_)] = vars_
    buf += v8.tobytes()
    for (op, a, b, imm) in ops:
        buf += struct.pack("3if", _lib.OP[op], a, b, imm)
    out = subprocess.run([exe], input=buf, stdout=subprocess.PIPE, check=True).stdout
    P = po.n_params(ls)
    arr = np.frombuffer(out, np.float32)
    o = 0
    jets = arr[o:o + n * C].reshape(n, C); o += n * C
    outv = arr[o:o + n]; o += n
    grad = arr[o:o + P]; o += P
    lamb = arr[o:o + n]; o += n
    varb = arr[o:o + 8]
    return jets, outv, grad, lamb, varb


# the user callbacks, written against the tf-like namespace exactly as in the reference's examples
def f_ac(u_model, x, t):
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)
    u_xx = tf.gradients(u_x, x)
    u_t = tf.gradients(u, t)
    c1 = tf.constant(.0001)
    c2 = tf.constant(5.0)
    return u_t - c1 * u_xx + c2 * u * u * u - c2 * u


def f_burgers(u_model, x, t):
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)
    u_xx = tf.gradients(u_x, x)
    u_t = tf.gradients(u, t)
    return u_t + u * u_x - (0.01 / tf.constant(math.pi)) * u_xx


def f_burgers_tape(u_model, x, t):   # test/Burgers2test.py:58-70
    with tf.GradientTape(persistent=True) as tape:
        tape.watch(x)
        tape.watch(t)
        u = u_model(tf.concat([x, t], 1))
        u_x = tape.gradient(u, x)
    u_xx = tape.gradient(u_x, x)
    u_t = tape.gradient(u, t)
    return u_t + u * u_x - 0.01 / tf.constant(math.pi) * u_xx


def f_poisson(u_model, x, y):
    u = u_model(tf.concat([x, y], 1))
    u_x = tf.gradients(u, x)[0]
    u_y = tf.gradients(u, y)[0]
    u_xx = tf.gradients(u_x, x)[0]
    u_yy = tf.gradients(u_y, y)[0]
    a1 = tf.constant(1.0)
    pi = tf.constant(math.pi)
    forcing = - tf.math.sin(a1 * pi * x) * tf.math.sin(a1 * pi * y)
    return u_xx + u_yy - forcing


def f_ch(u_model, var, x, t):
    g1, g2 = var[0], var[1]
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)
    u_t = tf.gradients(u, t)
    u_xx = tf.gradients(u_x, x)
    tmp = g1 * (u ** 3 - u) - g2 * u_xx
    tmp_x = tf.gradients(tmp, x)[0]
    tmp_xx = tf.gradients(tmp_x, x)[0]
    return u_t - tmp_xx


def f_special(u_model, x, t):
    """not a reference example: exercises divide, sin, exp and a negative integer power in the program"""
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)[0]
    u_t = tf.gradients(u, t)[0]
    return tf.math.sin(u) / (2.0 + u_x * u_x) + tf.math.exp(-u) * (1.5 + x) ** -2 + 0.1 * u_t


def _f_special_torch(net, x, t):
    u = net(torch.cat([x, t], 1))
    u_x = po.gradients(u, x)
    u_t = po.gradients(u, t)
    return torch.sin(u) / (2.0 + u_x * u_x) + torch.exp(-u) * (1.5 + x) ** -2 + 0.1 * u_t


CASES = [
    ("special", f_special, [2, 12, 12, 1], (1, 1, 0), [(-1, 1), (0, 1)]),
    ("allen_cahn", f_ac, [2, 24, 24, 24, 1], (2, 1, 0), [(-1, 1), (0, 1)]),
    ("burgers", f_burgers, [2, 20, 20, 20, 20, 1], (2, 1, 0), [(-1, 1), (0, 1)]),
    ("burgers", f_burgers_tape, [2, 20, 20, 1], (2, 1, 0), [(-1, 1), (0, 1)]),
    ("poisson", f_poisson, [2, 16, 16, 1], (2, 2, 0), [(0, 1), (0, 1)]),
    ("cahn_hilliard", f_ch, [2, 32, 32, 32, 1], (4, 1, 0), [(-1, 1), (0, 1)]),
]


@pytest.mark.parametrize("compiled", [False, True, "packed"], ids=["interpreted", "compiled", "compiled-packed"])
@pytest.mark.parametrize("name,fn,ls,cfg,lim", CASES)
def test_traced_residual_and_jets(harness, name, fn, ls, cfg, lim, compiled, monkeypatch):
    """`compiled`: the program goes through tc_compile (csrc/vm_poly.h), the form the tensor-core kernel evaluates;
    `packed`: and the jet formulas through their two-points-at-a-time instantiation (csrc/f32x2.cuh)"""
    if compiled == "packed":
        monkeypatch.setenv("TDQ_HARNESS_F2", "1")
    else:
        monkeypatch.delenv("TDQ_HARNESS_F2", raising=False)
    if compiled:
        monkeypatch.setenv("TDQ_HARNESS_POLY", "1")
    else:
        monkeypatch.delenv("TDQ_HARNESS_POLY", raising=False)
    rng = np.random.RandomState(1)
    n, d = 37, 2
    X = np.stack([rng.uniform(lo, hi, n) for lo, hi in lim], 1).astype(np.float32)
    w = po.glorot_normal_flat(ls, seed=2, bias_std=0.1)
    lam = rng.uniform(0.5, 2.0, n).astype(np.float32)
    vars_ = np.array([0.7, 0.05], np.float32) if name == "cahn_hilliard" else np.zeros(0, np.float32)
    # ---- trace through the product's tracer
    if name == "cahn_hilliard":
        vs = [tf.Variable(float(v)) for v in vars_]
        for i, v in enumerate(vs):
            v._sym = T.var(i)
        fe = T.as_sym(_trace(fn, d, extra_first=vs))
    else:
        fe = T.as_sym(_trace(fn, d))
    ops, slots, orders = T.lower({"out": _integrand(fe, False, "mse"), "f": fe}, d)
    assert tuple(orders) + (0,) == cfg
    C = 1 + sum(cfg)
    jets, outv, grad, lamb, varb = run_harness(harness, cfg, ls, w, X, np.zeros(n), lam, vars_, ops, slots["out"], C)
    # ---- oracle: nested autograd, float64 for the truth
    wt = torch.tensor(w, dtype=torch.float64, requires_grad=True)
    net = po.UModel(wt, ls)
    cols = [torch.tensor(X[:, j:j + 1], dtype=torch.float64, requires_grad=True) for j in range(d)]
    vt = [torch.tensor(float(v), dtype=torch.float64, requires_grad=True) for v in vars_]
    lt = torch.tensor(lam.reshape(-1, 1), dtype=torch.float64, requires_grad=True)
    if name == "special":
        f = _f_special_torch(net, *cols)
    else:
        f = po.RESIDUALS[name](net, vt, *cols) if name == "cahn_hilliard" else po.RESIDUALS[name](net, *cols)
    ell = (lt * f) ** 2
    gs = torch.autograd.grad(ell.sum(), [wt, lt] + vt)
    assert np.allclose(outv, ell.detach().numpy().ravel(), rtol=2e-4, atol=1e-6)
    gref = gs[0].numpy()
    assert np.linalg.norm(grad - gref) / np.linalg.norm(gref) < 2e-4
    assert np.allclose(lamb, gs[1].numpy().ravel(), rtol=2e-4, atol=1e-6)
    for i in range(len(vt)):
        assert abs(varb[i] - float(gs[2 + i])) <= 2e-4 * abs(float(gs[2 + i])) + 1e-6
    # ---- jets: channel 0 is u, then pure derivatives per dimension
    u = net(torch.cat(cols, 1))
    assert np.allclose(jets[:, 0], u.detach().numpy().ravel(), rtol=1e-5, atol=1e-6)
    c = 1
    for dim in range(d):
        g = u
        for _o in range(cfg[dim]):
            g = po.gradients(g, cols[dim])
            ref = g.detach().numpy().ravel()
            assert np.allclose(jets[:, c], ref, rtol=5e-4, atol=1e-4 * max(1.0, np.abs(ref).max())), (dim, _o)
            c += 1


def test_program_matches_numpy_eval():
    d = 2
    fe = T.as_sym(_trace(f_ac, d))
    ops, slots, orders = T.lower({"f": fe}, d)
    rng = np.random.RandomState(0)
    jets = rng.standard_normal((1 + sum(orders), 11))
    coords = rng.standard_normal((2, 11))
    sl = T.eval_program(ops, jets, coords)
    u, ux, uxx, ut = jets[0], jets[1], jets[2], jets[3]
    ref = ut - np.float32(1e-4) * uxx + 5 * u ** 3 - 5 * u
    assert np.allclose(sl[slots["f"]], ref, rtol=1e-6)
    assert fe.listy   # the callback mixed tf.gradients lists into the arithmetic (quirk Q1)


def test_mixed_partials_beyond_u_xy_rejected():
    """u_xy of a two-dimensional problem is carried (tests/test_host_cpu.py::test_mixed_partial_lowering); mixed partials
    in three dimensions or of higher order are not"""
    def f(u_model, x, y, z):
        u = u_model(tf.concat([x, y, z], 1))
        u_x = tf.gradients(u, x)[0]
        return tf.gradients(u_x, y)[0]
    with pytest.raises(T.MixedPartialError):
        _trace(f, 3)
