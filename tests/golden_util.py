"""Load tests/golden/*.npz (written by tests/golden/make_golden.py from the reference's own
code) into oracle.Problem objects and neutral dicts the CUDA-path tests reuse."""
import os

import numpy as np

from oracle import pinn_oracle as po

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

RESIDUAL_OF = {"ac_sa_small": "allen_cahn", "ac_sa_g": "allen_cahn", "ac_sa_128": "allen_cahn",
               "ac_dist_128": "allen_cahn", "burgers": "burgers", "poisson": "poisson"}


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
    return po.Problem([int(v) for v in z["layer_sizes"]], RESIDUAL_OF[name], z["X_f"], terms_of(z), g=g_of(name))


def lambdas_of(z):
    """(lam_res, {bc index: lam}) as the golden case used them (lam0_0 residual, lam0_1 IC = BC 0)."""
    lam_res = z["lam0_0"] if "lam0_0" in z else None
    lam_bcs = {0: z["lam0_1"]} if "lam0_1" in z else {}
    return lam_res, lam_bcs
