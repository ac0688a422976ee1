"""GPU: size-independent properties at BASELINE.json's full sizes (AC [2,128x4,1]), where the oracle
is too slow: run-to-run bit reproducibility, shard additivity (the multi-GPU decomposition on one
device), ragged tails, empty shards, minibatch ranges."""
import numpy as np
import pytest

from oracle import pinn_oracle as po
from tensordiffeq_b200 import _lib
from tensordiffeq_b200 import trace as T
from tensordiffeq_b200.models import _trace, _integrand
from tests.test_jet_host import f_ac

pytestmark = pytest.mark.gpu
LS = [2, 128, 128, 128, 128, 1]


def ac_engine(X, n_global=None, lam=None):
    eng = _lib.Engine(LS)
    fe = T.as_sym(_trace(f_ac, 2))
    ops, slots, orders = T.lower({"out": _integrand(fe, False, "mse" if lam is not None else None), "f": fe}, 2)
    t = eng.add_term(2, orders, ops, slots["out"], slots["f"], 0)
    eng.set_points(t, X, n_global or len(X))
    if lam is not None and len(X):
        eng.set_lambda(t, lam)
    eng.set_params(po.glorot_normal_flat(LS, seed=0, bias_std=0.05))
    return eng, t


def points(n, seed=1234):
    return po.centred_lhs(n, [[-1, 1], [0, 1]], seed).astype(np.float32)


def test_bit_reproducible_50k():
    X = points(50000)
    eng, _ = ac_engine(X)
    l1, g1, _ = eng.loss_grad()
    l2, g2, _ = eng.loss_grad()
    assert l1 == l2 and np.array_equal(g1, g2)
    eng.close()


def test_shard_additivity_and_tail_200k():
    n = 200003                       # not a multiple of the 32-point tile
    X = points(n)
    lam = np.random.RandomState(3).uniform(0, 1, n).astype(np.float32)
    eng, _ = ac_engine(X, lam=lam)
    L, G, _ = eng.loss_grad()
    eng.close()
    cuts = [0, 70001, 70001, 150000, n]      # includes an empty shard
    Ls, Gs = 0.0, np.zeros_like(G, dtype=np.float64)
    for a, b in zip(cuts[:-1], cuts[1:]):
        e, _ = ac_engine(X[a:b], n_global=n, lam=lam[a:b])
        l, g, _ = e.loss_grad()
        Ls += l
        Gs += g
        e.close()
    assert abs(Ls - L) <= 1e-5 * abs(L)
    assert np.linalg.norm(Gs - G) / np.linalg.norm(G) < 1e-5


def test_full_size_2m_runs_and_matches_subsample_scaling():
    """N_f = 2M (config 5): finite loss / gradient, and the mean over the full LHS set agrees with the
    mean over a 1/8 strided subsample to sampling accuracy (the loss is a Monte-Carlo mean)."""
    n = 2_000_000
    X = points(n)
    eng, _ = ac_engine(X)
    L, G, _ = eng.loss_grad()
    assert np.isfinite(L) and np.all(np.isfinite(G))
    e2, _ = ac_engine(np.ascontiguousarray(X[::8]))
    L2, G2, _ = e2.loss_grad()
    assert abs(L - L2) < 0.05 * abs(L)
    eng.close(); e2.close()


def test_minibatch_range_equals_subset():
    X = points(4096)
    eng, t = ac_engine(X)
    eng.set_range(t, 1024, 512, 512)
    l1, g1, _ = eng.loss_grad()
    e2, _ = ac_engine(np.ascontiguousarray(X[1024:1536]))
    l2, g2, _ = e2.loss_grad()
    assert abs(l1 - l2) <= 1e-6 * abs(l2)
    assert np.linalg.norm(g1 - g2) / np.linalg.norm(g2) < 1e-6
    eng.close(); e2.close()


def test_errors_are_loud():
    with pytest.raises(_lib.TdqError):
        _lib.Engine([2, 300, 1])              # hidden width beyond the compiled kernels
    with pytest.raises(_lib.TdqError):
        _lib.Engine([2, 20, 3])               # vector-valued network
    eng = _lib.Engine([2, 20, 1])
    with pytest.raises(_lib.TdqError):
        eng.set_params(np.zeros(5, np.float32))
    eng.close()
