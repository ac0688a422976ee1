"""GPU: the CUDA path, called through the C-ABI (ctypes -> libtdq_b200.so) behind the public host API,
against the oracle and against the golden fixtures written by the reference's own code.

Tolerance: north_star asks for loss / residual / parameter gradients within rel 1e-4 in fp32; the
checks below use 1e-4 on norms (and say so where an individual check is looser and why)."""
import numpy as np
import pytest

from oracle import pinn_oracle as po
from tests import golden_util as gu
from tests.gpu_util import solver_from_golden, rel

pytestmark = pytest.mark.gpu

COLLOC = ["ac_sa_small", "ac_sa_g", "ac_sa_128", "ac_dist_128", "burgers", "poisson", "mixed_xy", "mixed_xy_128"]   # mixed_xy*: a residual reading u_xy
TOL = 1e-4


@pytest.mark.parametrize("name", COLLOC)
def test_predict_matches_reference(name):
    model, z = solver_from_golden(name)
    u, f = model.predict(z["X_star"])
    assert u.shape == z["u_star"].shape
    assert f.shape == z["f_star"].shape          # [1,M,1] where the callback mixed lists in (Q1)
    assert rel(u, z["u_star"]) < TOL, "u"
    assert rel(f, z["f_star"]) < TOL, "f"


@pytest.mark.parametrize("name", COLLOC)
def test_loss_grad_matches_reference(name):
    model, z = solver_from_golden(name)
    loss, grads = model.grad()
    n_net = 2 * (len(z["layer_sizes"]) - 1)
    flat = np.concatenate([g.ravel() for g in grads[:n_net]])
    assert abs(float(loss) - float(z["loss0"])) <= TOL * abs(float(z["loss0"])), (float(loss), float(z["loss0"]))
    assert rel(flat, z["grad0"]) < TOL, "weight gradient"
    f = model._engine().get_residual(model._terms[-1]["tid"])
    assert rel(f, z["f0"]) < TOL, "residual"
    if "lamgrad0_0" in z:
        assert rel(grads[n_net], z["lamgrad0_0"]) < TOL, "d loss / d lambda_residual"
        assert rel(grads[n_net + 1], z["lamgrad0_1"]) < TOL, "d loss / d lambda_IC"


@pytest.mark.parametrize("name", COLLOC)
def test_loss_grad_matches_oracle_per_layer(name):
    """same inputs through the oracle directly; reports the worst layer so a kernel fault localises"""
    model, z = solver_from_golden(name)
    prob = gu.problem_of(name, z)
    lam_res, lam_bcs = gu.lambdas_of(z)
    r = po.loss_and_grads(prob, z["w0"], lam_res, lam_bcs)
    loss, grads = model.grad()
    assert abs(float(loss) - float(r["loss"])) <= TOL * abs(float(r["loss"]))
    ref_layers = po.unflatten(r["grad"], [int(v) for v in z["layer_sizes"]])
    worst = 0.0
    for i, (W, b) in enumerate(ref_layers):
        eW, eb = rel(grads[2 * i], W), rel(grads[2 * i + 1], b)
        worst = max(worst, eW, eb)
        assert eW < 5 * TOL and eb < 5 * TOL, "layer %d: dW rel %.2e db rel %.2e" % (i + 1, eW, eb)
    assert worst < 5 * TOL


@pytest.mark.parametrize("name", ["ac_sa_small", "ac_sa_g", "burgers", "poisson", "ac_sa_128", "mixed_xy", "mixed_xy_128"])
def test_adam_trajectory(name):
    model, z = solver_from_golden(name)
    if name == "poisson":
        from tensordiffeq_b200 import tf
        model.tf_optimizer = tf.keras.optimizers.Adam(lr=.005)       # steady-state-poisson.py:59
    n = len(z["adam_loss_hist"])
    model.verbose = False
    model.fit(tf_iter=n)
    hist = np.array(model.loss_history[:n], np.float32)
    # band: each step's loss within 1e-3 of the reference trajectory (errors compound through Adam's
    # sqrt(v) normalisation: an fp32-rounding-sized gradient change moves early steps by O(lr))
    assert np.allclose(hist, z["adam_loss_hist"], rtol=1e-3), (hist, z["adam_loss_hist"])
    assert rel(model.u_model.get_flat(), z["w_after_adam"]) < 1e-3
    if "lam_after_adam_0" in z:
        assert rel(model.lambdas[0].numpy(), z["lam_after_adam_0"]) < 1e-4
        assert rel(model.lambdas[1].numpy(), z["lam_after_adam_1"]) < 1e-4


@pytest.mark.parametrize("name", ["ac_two_res", "ac_two_res_both"])
def test_residual_tuple(name):
    """A callback returning a tuple of two residuals (models.py:160-188), the first self-adaptive, then both (the
    reference weights every adaptive residual with lambda vector 0): loss, gradients w.r.t. weights and lambdas,
    first residual field and 8 Adam steps against the reference's own numbers."""
    model, z = solver_from_golden(name)
    loss, grads = model.grad()
    n_net = 2 * (len(z["layer_sizes"]) - 1)
    flat = np.concatenate([g.ravel() for g in grads[:n_net]])
    assert abs(float(loss) - float(z["loss0"])) <= TOL * abs(float(z["loss0"])), (float(loss), float(z["loss0"]))
    assert rel(flat, z["grad0"]) < TOL
    assert rel(grads[n_net], z["lamgrad0_0"]) < TOL
    assert rel(grads[n_net + 1], z["lamgrad0_1"]) < TOL
    assert rel(model._engine().get_residual(model._terms[-1]["tid"]), z["f0"]) < TOL
    model.verbose = False
    model.fit(tf_iter=len(z["adam_loss_hist"]))
    assert np.allclose(model.loss_history, z["adam_loss_hist"], rtol=1e-3)
    assert rel(model.u_model.get_flat(), z["w_after_adam"]) < 1e-3
    assert rel(model.lambdas[0].numpy(), z["lam_after_adam_0"]) < 1e-4


def test_lbfgs_burgers():
    model, z = solver_from_golden("burgers")
    model.u_model.set_flat(z["w_after_adam"])
    model.verbose = False
    n = len(z["lbfgs_f_hist"]) - 1
    model.fit(tf_iter=0, newton_iter=n)
    assert model.lbfgs_n_eval == int(z["lbfgs_n_eval"])
    assert np.allclose(np.array(model.lbfgs_history), z["lbfgs_f_hist"], rtol=5e-3), model.lbfgs_history
    assert rel(model.u_model.get_flat(), z["w_after_lbfgs"]) < 5e-3


def test_lbfgs_ac_first_steps():
    model, z = solver_from_golden("ac_dist_128")
    model.u_model.set_flat(z["w_after_adam"])
    model.verbose = False
    model.fit(tf_iter=0, newton_iter=3)
    assert np.allclose(np.array(model.lbfgs_history)[:3], z["lbfgs_f_hist"][:3], rtol=5e-3), model.lbfgs_history


@pytest.mark.parametrize("name,cw,fma", [("ch_discovery", False, False), ("ch_discovery_sa", True, False),
                                         ("ch_discovery_128", False, False), ("ch_discovery_128_sa", True, False),
                                         ("ch_discovery_128", False, True), ("ch_discovery_128_sa", True, True)])
def test_discovery(name, cw, fma, monkeypatch):
    """_128: BASELINE config 4's own net shape, on the tensor-core kernel and (fma) on fused_jet_kernel<4,1,0,128,2,true>"""
    if fma:
        monkeypatch.setenv("TDQ_DBG", "24")
    else:
        monkeypatch.delenv("TDQ_DBG", raising=False)
    from tensordiffeq_b200 import tf
    from tensordiffeq_b200.models import DiscoveryModel
    from tests.test_jet_host import f_ch
    z = gu.load(name)
    ls = [int(v) for v in z["layer_sizes"]]
    params = [tf.Variable(float(v), dtype=tf.float32) for v in z["vars0"]]
    X = [z["X"][:, 0:1], z["X"][:, 1:2]]
    col = tf.Variable(z["col_weights0"]) if cw else None
    model = DiscoveryModel()
    model.compile(ls, f_ch, X, tf.constant(z["u"]), params, col_weights=col)
    model.u_model.set_flat(z["w0"])
    loss, grads = model.grad()
    kernel = model._engine().term_kernel(model._t_res)
    # the 128-wide net (BASELINE config 4's own shape) runs the 16-point-tile tensor-core kernel, the 32-wide one the FMA kernel
    assert kernel.startswith("fused_jet_tc16_kernel<4,1,0" if (ls[1] == 128 and not fma) else "fused_jet_kernel<4,1,0"), kernel
    n_net = 2 * (len(ls) - 1)
    flat = np.concatenate([g.ravel() for g in grads[:n_net]])
    assert abs(float(loss) - float(z["loss0"])) <= TOL * abs(float(z["loss0"]))
    assert rel(flat, z["grad0"]) < 2 * TOL          # 4th-order jets: a little more fp32 rounding
    assert rel(np.array(grads[-2:]), z["vargrad0"]) < 2 * TOL
    if cw:
        assert rel(grads[n_net], z["colgrad0"]) < 2 * TOL
    hist = model.train_op(len(z["adam_loss_hist"]))
    model._sync_host()
    assert np.allclose(hist, z["adam_loss_hist"], rtol=2e-3)
    w = model.u_model.get_flat()
    assert rel(w, z["w_after_adam"]) < 2e-3
    assert w[-1] == z["w0"][-1]                      # quirk Q6: the output bias never moves
    assert np.allclose([float(p.numpy()) for p in params], z["vars_after_adam"], rtol=5e-2, atol=1e-5)
    if cw:
        assert rel(col.numpy(), z["col_weights_after_adam"]) < 1e-4


def test_minibatch_epoch_matches_oracle():
    """fit(batch_sz=...) (models.py:148-155,172-174; fit.py:86-104): one frozen permutation, consecutive slices,
    BC terms re-evaluated for every batch, Adam on weights and ascent on the batch's lambdas per batch."""
    name = "ac_sa_small"
    model, z = solver_from_golden(name)
    n_f, bs = len(z["X_f"]), 500
    np.random.seed(123)
    perm = np.random.choice(n_f, size=n_f, replace=False)
    np.random.seed(123)                       # fit() draws the same permutation
    model.verbose = False
    model.fit(tf_iter=2, batch_sz=bs)
    assert np.array_equal(model.batch_indx_map, perm)
    # oracle: same schedule
    lam_res, lam_bcs = gu.lambdas_of(z)
    w = z["w0"].copy()
    lam_res = lam_res.copy()
    lam_ic = lam_bcs[0].copy()
    opt_w, opt_l = po.KerasAdam(), po.KerasAdam()
    losses = []
    for _epoch in range(2):
        for b in range(n_f // bs):
            idx = perm[b * bs:(b + 1) * bs]
            prob = po.Problem([int(v) for v in z["layer_sizes"]], gu.RESIDUAL_OF[name], z["X_f"][idx], gu.terms_of(z))
            r = po.loss_and_grads(prob, w, lam_res[idx], {0: lam_ic})
            opt_w.begin_step()
            w = opt_w.apply("w", w, r["grad"])
            opt_l.begin_step()
            g_full = np.zeros_like(lam_res)
            g_full[idx] = r["lam_res_grad"]
            # keras applies Adam to the whole lambda variable: entries outside the batch see a zero gradient
            lam_res = opt_l.apply("res", lam_res, -g_full)
            lam_ic = opt_l.apply("ic", lam_ic, -r["lam_bc_grads"][0])
            loss = float(r["loss"])
        losses.append(loss)
    assert np.allclose(model.loss_history[:2], losses, rtol=1e-3), (model.loss_history, losses)
    assert rel(model.u_model.get_flat(), w) < 1e-3
    assert rel(model.lambdas[1].numpy(), lam_ic) < 1e-4
    # residual lambdas: the reference's dense Adam also moves entries outside the current batch through their
    # decaying moments; the engine reproduces that with a dense update after each batch
    assert rel(model.lambdas[0].numpy(), lam_res) < 1e-4
