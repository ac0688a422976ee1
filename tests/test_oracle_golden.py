"""CPU: the oracle restatement (oracle/pinn_oracle.py) against fixtures produced by the
reference's own Python running on the TF stand-in (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle import pinn_oracle as po
from tests import golden_util as gu

COLLOC = ["ac_sa_small", "ac_sa_g", "ac_sa_128", "ac_dist_128", "burgers", "poisson", "mixed_xy", "mixed_xy_128"]   # mixed_xy*: a residual reading u_xy


def rel(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


@pytest.mark.parametrize("name", COLLOC)
def test_loss_grad_residual(name):
    z = gu.load(name)
    prob = gu.problem_of(name, z)
    lam_res, lam_bcs = gu.lambdas_of(z)
    r = po.loss_and_grads(prob, z["w0"], lam_res, lam_bcs)
    # same fp32 op order on the same torch build: tight tolerances
    assert abs(float(r["loss"]) - float(z["loss0"])) <= 2e-6 * abs(float(z["loss0"]))
    assert rel(r["grad"], z["grad0"]) < 1e-5
    assert rel(r["f"], z["f0"]) < 1e-5
    if lam_res is not None:
        assert rel(r["lam_res_grad"], z["lamgrad0_0"]) < 1e-5
        assert rel(r["lam_bc_grads"][0], z["lamgrad0_1"]) < 1e-5


@pytest.mark.parametrize("name", COLLOC)
def test_predict(name):
    z = gu.load(name)
    prob = gu.problem_of(name, z)
    u, f = po.predict(prob, z["w0"], z["X_star"])
    assert rel(u, z["u_star"]) < 1e-6
    assert rel(f, z["f_star"]) < 1e-5


@pytest.mark.parametrize("name", ["ac_sa_small", "ac_sa_g", "burgers", "poisson", "mixed_xy"])
def test_adam_trajectory(name):
    z = gu.load(name)
    prob = gu.problem_of(name, z)
    lam_res, lam_bcs = gu.lambdas_of(z)
    n = len(z["adam_loss_hist"])
    opt_w = po.KerasAdam(beta_1=0.9) if name == "poisson" else None  # steady-state-poisson.py:59
    r = po.adam_train(prob, z["w0"], n, lam_res, lam_bcs, opt_w=opt_w)
    assert np.allclose(r["loss_hist"], z["adam_loss_hist"], rtol=2e-4)
    assert rel(r["w"], z["w_after_adam"]) < 1e-4
    if lam_res is not None:
        assert rel(r["lam_res"], z["lam_after_adam_0"]) < 1e-5
        assert rel(r["lam_bcs"][0], z["lam_after_adam_1"]) < 1e-5


@pytest.mark.parametrize("name", ["ac_two_res", "ac_two_res_both"])
def test_residual_tuple(name):
    """f_model returning two residuals, one or both self-adaptive (models.py:160-188)"""
    z = gu.load(name)
    prob = gu.problem_of(name, z)
    lam_res, lam_bcs = gu.lambdas_of(z)
    r = po.loss_and_grads(prob, z["w0"], lam_res, lam_bcs)
    assert abs(float(r["loss"]) - float(z["loss0"])) <= 2e-6 * abs(float(z["loss0"]))
    assert rel(r["grad"], z["grad0"]) < 1e-5
    assert rel(r["f"], z["f0"]) < 1e-5
    assert rel(r["lam_res_grad"], z["lamgrad0_0"]) < 1e-5
    assert rel(r["lam_bc_grads"][0], z["lamgrad0_1"]) < 1e-5
    t = po.adam_train(prob, z["w0"], len(z["adam_loss_hist"]), lam_res, lam_bcs)
    assert np.allclose(t["loss_hist"], z["adam_loss_hist"], rtol=2e-4)
    assert rel(t["lam_res"], z["lam_after_adam_0"]) < 1e-5


def test_adam_long_first_300_steps():
    """the 1000-step fixture (tests/golden/ac_sa_small_long.npz); the CPU suite follows its first 300 steps, the GPU
    suite all 1000 (tests/test_gpu_adam_long.py); tools/adam_band.py cpu runs the oracle over the whole trajectory"""
    z = gu.load("ac_sa_small_long")
    prob = gu.problem_of("ac_sa_small_long", z)
    lam_res, lam_bcs = gu.lambdas_of(z)
    opt_w, opt_lam = po.KerasAdam(), po.KerasAdam()
    r = po.adam_train(prob, z["w0"], 100, lam_res, lam_bcs, opt_w, opt_lam)
    assert rel(r["w"], z["w_after_100"]) < 1e-5
    r2 = po.adam_train(prob, r["w"], 200, r["lam_res"], r["lam_bcs"], opt_w, opt_lam)
    hist = np.concatenate([r["loss_hist"], r2["loss_hist"]])
    assert np.allclose(hist, z["adam_loss_hist"][:300], rtol=1e-3)


def test_lbfgs_burgers():
    z = gu.load("burgers")
    prob = gu.problem_of("burgers", z)
    n = len(z["lbfgs_f_hist"]) - 1
    x, f_hist, n_eval = po.lbfgs_train(prob, z["w_after_adam"], n)
    assert n_eval == int(z["lbfgs_n_eval"])
    assert np.allclose(np.array(f_hist), z["lbfgs_f_hist"], rtol=2e-3)
    assert rel(x, z["w_after_lbfgs"]) < 2e-3


def test_lbfgs_first_steps_ac():
    # the reference's fixed-step L-BFGS diverges on this SA state; compare the first iterations only
    z = gu.load("ac_sa_small")
    prob = gu.problem_of("ac_sa_small", z)
    x, f_hist, _ = po.lbfgs_train(prob, z["w_after_adam"], 3, z["lam_after_adam_0"], {0: z["lam_after_adam_1"]})
    assert np.allclose(np.array(f_hist)[:3], z["lbfgs_f_hist"][:3], rtol=5e-3)


@pytest.mark.parametrize("name,cw", [("ch_discovery", False), ("ch_discovery_sa", True),
                                     ("ch_discovery_128", False), ("ch_discovery_128_sa", True)])   # _128: BASELINE config 4's own net shape
def test_discovery(name, cw):
    z = gu.load(name)
    ls = [int(v) for v in z["layer_sizes"]]
    col = z["col_weights0"] if cw else None
    r = po.discovery_loss_and_grads(ls, "cahn_hilliard", z["X"], z["u"], z["w0"], z["vars0"], col)
    assert abs(float(r["loss"]) - float(z["loss0"])) <= 1e-5 * abs(float(z["loss0"]))
    assert rel(r["grad"], z["grad0"]) < 1e-4
    assert rel(r["var_grads"], z["vargrad0"]) < 1e-4
    if cw:
        assert rel(r["col_grad"], z["colgrad0"]) < 1e-4
    t = po.discovery_train(ls, "cahn_hilliard", z["X"], z["u"], z["w0"], z["vars0"], len(z["adam_loss_hist"]), col)
    assert np.allclose(t["loss_hist"], z["adam_loss_hist"], rtol=1e-3)
    assert rel(t["w"], z["w_after_adam"]) < 1e-3
    assert np.allclose(t["vars"], z["vars_after_adam"], rtol=2e-2, atol=1e-5)
    assert t["w"][-1] == z["w0"][-1]  # quirk Q6: output bias never updated
    if cw:
        assert rel(t["col_weights"], z["col_weights_after_adam"]) < 1e-4
