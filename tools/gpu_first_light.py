"""Stage-by-stage bring-up report for a fresh GPU box (developer tool, not a test):
forward-only kernel -> residual -> loss/grad per layer -> Adam, each against the oracle."""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import pinn_oracle as po  # noqa: E402  (developer tool: the oracle is the checker)
from tests import golden_util as gu  # noqa: E402
from tests.gpu_util import solver_from_golden, rel  # noqa: E402


def stage(name):
    def deco(fn):
        def run():
            t0 = time.time()
            try:
                fn()
                print("[ok  ] %-40s %.2fs" % (name, time.time() - t0), flush=True)
            except Exception:
                print("[FAIL] %-40s" % name, flush=True)
                traceback.print_exc()
                sys.stdout.flush()
        return run
    return deco


def per_layer(grads, ref_flat, ls):
    ref = po.unflatten(ref_flat, ls)
    for i, (W, b) in enumerate(ref):
        print("      layer %d: dW rel %.2e  db rel %.2e   |dW| %.3e" % (i + 1, rel(grads[2 * i], W), rel(grads[2 * i + 1], b), np.linalg.norm(W)))


def check_case(name):
    @stage("case " + name)
    def run():
        model, z = solver_from_golden(name)
        ls = [int(v) for v in z["layer_sizes"]]
        u, f = model.predict(z["X_star"])
        print("   predict: u rel %.2e  f rel %.2e" % (rel(u, z["u_star"]), rel(f, z["f_star"])))
        loss, grads = model.grad()
        n_net = 2 * (len(ls) - 1)
        flat = np.concatenate([g.ravel() for g in grads[:n_net]])
        print("   loss %.6e ref %.6e rel %.2e ; grad rel %.2e" % (loss, z["loss0"], abs(loss - z["loss0"]) / abs(z["loss0"]), rel(flat, z["grad0"])))
        per_layer(grads, z["grad0"], ls)
        fres = model._engine().get_residual(model._terms[-1]["tid"])
        print("   residual rel %.2e" % rel(fres, z["f0"]))
        if "lamgrad0_0" in z:
            print("   lam grads rel %.2e %.2e" % (rel(grads[n_net], z["lamgrad0_0"]), rel(grads[n_net + 1], z["lamgrad0_1"])))
        # term by term against the oracle
        prob = gu.problem_of(name, z)
        n = len(z["adam_loss_hist"])
        model.verbose = False
        if name == "poisson":
            from tensordiffeq_b200 import tf
            model.tf_optimizer = tf.keras.optimizers.Adam(lr=.005)
        model.fit(tf_iter=n)
        print("   adam hist ours %s" % np.array(model.loss_history[:n]))
        print("   adam hist ref  %s" % z["adam_loss_hist"])
        print("   w after adam rel %.2e" % rel(model.u_model.get_flat(), z["w_after_adam"]))
        print("   info", model._engine().info())
    return run


@stage("engine + ffma peak")
def s0():
    from tensordiffeq_b200 import _lib
    eng = _lib.Engine([2, 20, 20, 1])
    print("   info", eng.info(), "launches", eng.info()["launches"])
    eng.close()


if __name__ == "__main__":
    s0()
    for nm in (sys.argv[1:] or ["poisson", "burgers", "ac_sa_small", "ac_dist_128", "ac_sa_128", "ac_sa_g"]):
        check_case(nm)()
