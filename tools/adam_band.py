"""1000 Adam steps of the golden *_long cases on the CPU oracle (float32 and float64) or on the GPU (FP32 FMA
kernel and, for 128-wide nets, the tensor-core kernel), against the reference's trajectory: prints the band
statistics (tests/adam_band.py).  python tools/adam_band.py cpu|gpu [case ...]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import golden_util as gu
from tests.adam_band import band_stats

CASES = ["ac_sa_small_long", "burgers_long", "ac_sa_128_long", "ac_dist_128_long"]


def run_cpu(name, dtype):
    import torch
    from oracle import pinn_oracle as po
    z = gu.load(name)
    prob = gu.problem_of(name, z)
    prob.dtype = dtype
    prob.cols = [c.to(dtype) for c in prob.cols]
    lam_res, lam_bcs = gu.lambdas_of(z)
    opt_w, opt_lam = po.KerasAdam(), po.KerasAdam()
    r100 = po.adam_train(prob, z["w0"], 100, lam_res, lam_bcs, opt_w, opt_lam)
    r = po.adam_train(prob, r100["w"], len(z["adam_loss_hist"]) - 100, r100["lam_res"], r100["lam_bcs"], opt_w, opt_lam)
    return band_stats(np.concatenate([r100["loss_hist"], r["loss_hist"]]), r100["w"], r["w"], z)


def run_gpu(name, dbg):
    from tests.gpu_util import solver_from_golden
    if dbg:
        os.environ["TDQ_DBG"] = str(dbg)
    else:
        os.environ.pop("TDQ_DBG", None)
    model, z = solver_from_golden(name)
    model.verbose = False
    n = len(z["adam_loss_hist"])
    model.fit(tf_iter=100)
    w100 = model.u_model.get_flat().copy()
    model.fit(tf_iter=n - 100)
    kern = model._engine().term_kernel(model._terms[-1]["tid"])
    st = band_stats(model.loss_history[:n], w100, model.u_model.get_flat(), z)
    st["kernel"] = kern
    return st


if __name__ == "__main__":
    mode = sys.argv[1]
    for name in (sys.argv[2:] or CASES):
        if mode == "cpu":
            import torch
            for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
                print(name, "oracle", tag, json.dumps(run_cpu(name, dt)), flush=True)
        else:
            print(name, "fma", json.dumps(run_gpu(name, 24)), flush=True)
            if "128" in name:
                print(name, "tc", json.dumps(run_gpu(name, 0)), flush=True)
