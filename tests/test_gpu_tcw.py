"""GPU: the tensor-core kernel (tcgen05 3xTF32 for every hidden-layer GEMM of 128-wide nets, tc_kernel.cuh) against
the FP32 FMA kernel (developer switch TDQ_DBG=24, read when the engine is created: 16 = no tensor-core kernel,
8 = no tensor-core weight gradient inside the FMA kernel) and against the oracle, layer by layer.
Tolerance: rel 1e-4 on every gradient block (north_star's fp32 parity bar); measured differences are ~2e-6."""
import os

import numpy as np
import pytest

from oracle import pinn_oracle as po
from tests.test_gpu_properties import ac_engine, points, LS

pytestmark = pytest.mark.gpu


def _grad(X, lam, dbg):
    old = os.environ.get("TDQ_DBG")
    os.environ["TDQ_DBG"] = str(dbg)
    try:
        eng, t = ac_engine(X, lam=lam)
    finally:
        if old is None:
            del os.environ["TDQ_DBG"]
        else:
            os.environ["TDQ_DBG"] = old
    L, G, _ = eng.loss_grad()
    f = eng.get_residual(t)
    eng.close()
    return L, G, f


def _layers(G):
    return po.unflatten(G, LS)


@pytest.mark.parametrize("n", [32, 1000, 200003])
def test_tensor_core_kernel_matches_fma_kernel(n):
    X = points(n)
    lam = np.random.RandomState(5).uniform(0, 1, n).astype(np.float32)
    L1, G1, f1 = _grad(X, lam, 0)
    L0, G0, f0 = _grad(X, lam, 24)
    assert abs(L1 - L0) <= 1e-5 * abs(L0)
    assert np.linalg.norm(f1 - f0) <= 1e-5 * np.linalg.norm(f0)
    for i, ((W1, b1), (W0, b0)) in enumerate(zip(_layers(G1), _layers(G0))):
        eW = np.linalg.norm(W1.astype(np.float64) - W0) / np.linalg.norm(W0)
        eb = np.linalg.norm(b1.astype(np.float64) - b0) / np.linalg.norm(b0)
        emax = np.abs(W1.astype(np.float64) - W0).max() / np.abs(W0).max()
        print("n=%d layer %d: dW rel(norm) %.2e  max-abs/max %.2e  db rel %.2e" % (n, i + 1, eW, emax, eb))
        assert eW < 1e-4 and emax < 1e-4 and eb < 1e-4


def test_tensor_core_kernel_matches_oracle():
    n = 1536
    X = points(n, seed=7)
    lam = np.random.RandomState(6).uniform(0, 1, n).astype(np.float32)
    L1, G1, _ = _grad(X, lam, 0)
    prob = po.Problem(LS, "allen_cahn", X.astype(np.float64), [])
    r = po.loss_and_grads(prob, po.glorot_normal_flat(LS, seed=0, bias_std=0.05), lam.reshape(-1, 1))
    assert abs(L1 - r["loss"]) <= 1e-4 * abs(r["loss"])
    for i, ((W1, b1), (W0, b0)) in enumerate(zip(_layers(G1), _layers(r["grad"]))):
        e = np.linalg.norm(W1.astype(np.float64) - W0) / np.linalg.norm(W0)
        print("layer %d: dW vs oracle %.2e" % (i + 1, e))
        assert e < 1e-4
