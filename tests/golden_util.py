"""Load tests/golden/*.npz (written by tests/golden/make_golden.py from the reference's own
code) into oracle.Problem objects and neutral dicts the CUDA-path tests reuse."""
import os

import numpy as np

from oracle import pinn_oracle as po

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

RESIDUAL_OF = {"ac_sa_small": "allen_cahn", "ac_sa_g": "allen_cahn", "ac_sa_128": "allen_cahn",
               "ac_dist_128": "allen_cahn", "burgers": "burgers", "poisson": "poisson",
               "ac_sa_small_long": "allen_cahn", "burgers_long": "burgers", "ac_sa_128_long": "allen_cahn",
               "ac_dist_128_long": "allen_cahn", "ac_two_res": "ac_two_res", "ac_two_res_both": "ac_two_res",
               "mixed_xy": "mixed_xy", "mixed_xy_128": "mixed_xy"}


def f_mixed_oracle(u_model, x, y):
    """tests/golden/make_golden.py:case_mixed on torch autograd"""
    import math
    import torch
    u = u_model(torch.cat([x, y], 1))
    u_x = po.gradients(u, x)
    u_y = po.gradients(u, y)
    u_xx = po.gradients(u_x, x)
    u_yy = po.gradients(u_y, y)
    u_xy = po.gradients(u_x, y)
    pi = float(np.float32(math.pi))
    return u_xx + 0.5 * u_xy * u + u_yy + u_xy * u_xy - torch.sin(pi * x) * torch.sin(pi * y)


def f_two_res_oracle(u_model, x, t):
    """the tuple-returning callback of tests/golden/make_golden.py:case_ac_two_res, on torch autograd"""
    import torch
    u = u_model(torch.cat([x, t], 1))
    u_x = po.gradients(u, x)
    u_xx = po.gradients(u_x, x)
    u_t = po.gradients(u, t)
    u_tt = po.gradients(u_t, t)
    f_1 = u_t - 0.0001 * u_xx + 5.0 * u * u * u - 5.0 * u
    f_2 = u_tt - u * u_x + 0.25 * x
    return f_1, f_2


def g_of(name):
    if name == "ac_sa_g":
        return lambda lam: lam ** 2 + 1.0
    return None


def load(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))


def terms_of(z):
    terms = []
    for i, kind in enumerate(z["bc_kinds"]):
        kind = str(kind)
        if kind == "periodic":
            terms.append(po.Term("periodic", upper=z["bc%d_upper" % i], lower=z["bc%d_lower" % i], var=["x"]))
        else:
            terms.append(po.Term(kind, X=z["bc%d_input" % i], val=np.asarray(z["bc%d_val" % i], np.float32)))
    return terms


def problem_of(name, z):
    res = RESIDUAL_OF[name]
    flags = [bool(v) for v in z["res_adaptive"]] if "res_adaptive" in z else None
    fm = {"ac_two_res": f_two_res_oracle, "mixed_xy": f_mixed_oracle}.get(res, res)
    return po.Problem([int(v) for v in z["layer_sizes"]], fm, z["X_f"],
                      terms_of(z), g=g_of(name), res_adaptive=flags)


def lambdas_of(z):
    """(lam_res, {bc index: lam}) as the golden case used them (lam0_0 residual, lam0_1 IC = BC 0)."""
    lam_res = z["lam0_0"] if "lam0_0" in z else None
    lam_bcs = {0: z["lam0_1"]} if "lam0_1" in z else {}
    return lam_res, lam_bcs
