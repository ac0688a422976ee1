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


def _multi_term_engine(X, ic_X, ic_val, per):
    """residual + IC + periodic terms on one handle: evaluated by ONE launch (the BC tiles join the residual's tile queue)"""
    eng = _lib.Engine(LS)
    fe = T.as_sym(_trace(f_ac, 2))
    ops, slots, orders = T.lower({"out": _integrand(fe), "f": fe}, 2)
    t_res = eng.add_term(2, orders, ops, slots["out"], slots["f"], 0)
    u = T.jet((0, 0))
    ops, slots, orders = T.lower({"out": _integrand(u, True)}, 2)
    t_ic = eng.add_term(2, orders, ops, slots["out"], -1, 0)
    ops, slots, orders = T.lower({"out": u}, 2)
    t_per = eng.add_term(2, orders, ops, slots["out"], -1, 1)
    eng.set_points(t_res, X, len(X))
    eng.set_points(t_ic, ic_X, len(ic_X))
    eng.set_target(t_ic, ic_val)
    eng.set_points(t_per, per, len(per) // 2)
    eng.set_params(po.glorot_normal_flat(LS, seed=0, bias_std=0.05))
    return eng, (t_res, t_ic, t_per)


def test_merged_launch_equals_sum_of_single_term_launches():
    """one launch for residual + IC + periodic (api.cu:build_plan) against three handles holding one term each"""
    rs = np.random.RandomState(4)
    X = points(30011)
    x_ic = np.linspace(-1, 1, 300)
    ic_X = np.stack([x_ic, np.zeros_like(x_ic)], 1).astype(np.float32)
    ic_val = (x_ic ** 2 * np.cos(np.pi * x_ic)).astype(np.float32)
    t_lv = rs.uniform(0, 1, 77).astype(np.float32)
    per = np.empty((154, 2), np.float32)
    per[0::2, 0], per[1::2, 0] = 1.0, -1.0
    per[0::2, 1] = per[1::2, 1] = t_lv
    eng, ts = _multi_term_engine(X, ic_X, ic_val, per)
    launches0 = eng.info()["launches"]
    L, G, _ = eng.loss_grad()
    assert eng.info()["launches"] - launches0 == 2          # the jet kernel for all three terms + the slab reduction
    assert len({eng.term_kernel(t) for t in ts}) == 1
    eng.close()
    Ls, Gs = 0.0, np.zeros_like(G, dtype=np.float64)
    for keep in range(3):
        e, ts = _multi_term_engine(X if keep == 0 else X[:0], ic_X if keep == 1 else ic_X[:0], ic_val if keep == 1 else ic_val[:0],
                                   per if keep == 2 else per[:0])
        l, g, _ = e.loss_grad()
        Ls += l
        Gs += g
        e.close()
    assert abs(Ls - L) <= 1e-5 * abs(L)
    assert np.linalg.norm(Gs - G) / np.linalg.norm(G) < 1e-5


def test_16_point_kernel_bit_reproducible():
    """BASELINE config 4's shape (6 jet channels, 25 600 points) twice on the 16-point tensor-core kernel"""
    from tests.test_gpu_tc_shapes import f_x4
    eng = _lib.Engine(LS)
    fe = T.as_sym(_trace(f_x4, 2))
    ops, slots, orders = T.lower({"out": _integrand(fe), "f": fe}, 2)
    t = eng.add_term(2, orders, ops, slots["out"], slots["f"], 0)
    eng.set_points(t, points(25600), 25600)
    eng.set_params(po.glorot_normal_flat(LS, seed=0, bias_std=0.05))
    assert eng.term_kernel(t).startswith("fused_jet_tc16_kernel")
    l1, g1, _ = eng.loss_grad()
    l2, g2, _ = eng.loss_grad()
    assert l1 == l2 and np.array_equal(g1, g2)
    eng.close()


def test_epochs_call_equals_per_batch_calls():
    """tdq_adam_epochs (the whole minibatch schedule in one call) against tdq_set_term_range + tdq_adam_steps per batch"""
    X = points(4000)
    lam = np.random.RandomState(5).uniform(0.5, 1.5, 4000).astype(np.float32)
    w_cfg = _lib.Engine.adam_cfg(0.005, 0.99, 0.999, 1e-7, 1)
    l_cfg = _lib.Engine.adam_cfg(0.005, 0.99, 0.999, 1e-7, -1)
    e1, t1 = ac_engine(X, lam=lam)
    h1 = e1.adam_epochs(2, [t1], 1000, 3, w=w_cfg, lam=l_cfg)          # 3 batches of 1000: the last 1000 points are never visited
    e2, t2 = ac_engine(X, lam=lam)
    h2 = []
    for _ in range(2):
        for b in range(3):
            e2.set_range(t2, b * 1000, 1000, 1000)
            h = e2.adam_steps(1, w=w_cfg, lam=l_cfg)
        h2.append(h[0])
    e2.set_range(t2, 0, -1, 1)
    assert np.array_equal(h1, np.array(h2, np.float32))
    assert np.array_equal(e1.get_params(), e2.get_params())
    assert np.array_equal(e1.get_lambda(t1), e2.get_lambda(t2))
    with pytest.raises(_lib.TdqError):
        e1.adam_epochs(1, [t1], 3000, 2, w=w_cfg)                       # 2 x 3000 > 4000 stored points
    e1.close(); e2.close()


def test_graph_mode_lbfgs_converges_and_reports_status():
    """tdq_lbfgs_hz on a tiny net: monotone decrease, status 0 (max_iterations) after a short budget"""
    ls = [2, 8, 1]
    eng = _lib.Engine(ls)
    fe = T.as_sym(_trace(f_ac, 2))
    ops, slots, orders = T.lower({"out": _integrand(fe)}, 2)
    t = eng.add_term(2, orders, ops, slots["out"], -1, 0)
    eng.set_points(t, points(500), 500)
    eng.set_params(po.glorot_normal_flat(ls, seed=1, bias_std=0.1))
    hist, n_eval, status = eng.lbfgs_hz(15)
    assert status == 0 and len(hist) == 16 and n_eval >= 16
    assert np.all(np.diff(hist) <= 0) and hist[-1] < hist[0]
    eng.close()


def test_prefetched_point_set_takes_effect_at_the_swap():
    """tdq_prefetch_term_points / tdq_swap_term_points: the staged set changes nothing until the swap, then equals a
    fresh upload of the same points bit for bit; two swaps in a row alternate the two buffers"""
    X0, X1 = points(30011, 1), points(30011, 2)
    eng, t = ac_engine(X0)
    l0, g0, _ = eng.loss_grad()
    ref, _ = ac_engine(X1)
    l1, g1, _ = ref.loss_grad()
    ref.close()
    assert l0 != l1
    eng.prefetch_points(t, X1)
    la, ga, _ = eng.loss_grad()                       # still the first set
    assert la == l0 and np.array_equal(ga, g0)
    eng.swap_points(t)
    lb, gb, _ = eng.loss_grad()
    assert lb == l1 and np.array_equal(gb, g1)
    eng.prefetch_points(t, X0)                        # lands in the buffer the first set was in
    eng.swap_points(t)
    lc, gc, _ = eng.loss_grad()
    assert lc == l0 and np.array_equal(gc, g0)
    with pytest.raises(_lib.TdqError):
        eng.swap_points(t)                            # nothing staged
    with pytest.raises(_lib.TdqError):
        eng.prefetch_points(t, X1[:100])              # a different count needs tdq_set_term_points
    eng.close()
