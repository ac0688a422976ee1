"""Builds a tensordiffeq_b200 solver for a golden case through the PUBLIC host API (DomainND, IC,
periodicBC, CollocationSolverND.compile), then swaps in the exact point sets / weights / lambdas the
reference run used (they are partly random in the reference: SURVEY Q3)."""
import math

import numpy as np

import tensordiffeq_b200 as tdq
from tensordiffeq_b200 import tf
from tensordiffeq_b200.boundaries import DomainND, IC, dirichletBC, periodicBC, FunctionDirichletBC
from tensordiffeq_b200.models import CollocationSolverND
from tests import golden_util as gu
from tests.test_jet_host import f_ac, f_burgers, f_poisson, f_ch  # the reference-style callbacks



def f_two_res(u_model, x, t):
    """tests/golden/make_golden.py:case_ac_two_res -- a callback returning a tuple of residuals (models.py:160-161)"""
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)[0]
    u_xx = tf.gradients(u_x, x)[0]
    u_t = tf.gradients(u, t)[0]
    u_tt = tf.gradients(u_t, t)[0]
    f_1 = u_t - tf.constant(.0001) * u_xx + tf.constant(5.0) * u * u * u - tf.constant(5.0) * u
    f_2 = u_tt - u * u_x + tf.constant(0.25) * x
    return f_1, f_2


def f_mixed(u_model, x, y):
    """tests/golden/make_golden.py:case_mixed -- reads the mixed partial u_xy"""
    u = u_model(tf.concat([x, y], 1))
    u_x = tf.gradients(u, x)[0]
    u_y = tf.gradients(u, y)[0]
    u_xx = tf.gradients(u_x, x)[0]
    u_yy = tf.gradients(u_y, y)[0]
    u_xy = tf.gradients(u_x, y)[0]
    return u_xx + tf.constant(0.5) * u_xy * u + u_yy + u_xy * u_xy - tf.math.sin(tf.constant(math.pi) * x) * tf.math.sin(tf.constant(math.pi) * y)


F_MODEL = {"allen_cahn": f_ac, "burgers": f_burgers, "poisson": f_poisson, "ac_two_res": f_two_res, "mixed_xy": f_mixed}


def deriv_model(u_model, x, t):
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)[0]
    return u, u_x


def solver_from_golden(name, z=None, dist=False, batch_info=None):
    z = z if z is not None else gu.load(name)
    res = gu.RESIDUAL_OF[name]
    ls = [int(v) for v in z["layer_sizes"]]
    kinds = [str(k) for k in z["bc_kinds"]]
    if res in ("poisson", "mixed_xy"):
        D = DomainND(["x", "y"])
        D.add("x", [0, 1.0], 11)
        D.add("y", [0, 1.0], 11)
    else:
        D = DomainND(["x", "t"], time_var="t")
        nx = len(z["bc0_input"])
        D.add("x", [-1.0, 1.0], nx)
        D.add("t", [0.0, 1.0], 7)
    D.X_f = np.asarray(z["X_f"], np.float64)
    bcs = []
    for i, kind in enumerate(kinds):
        if kind == "periodic":
            bc = periodicBC(D, ["x"], [deriv_model])
            bc.upper = [np.asarray(a) for a in z["bc%d_upper" % i]]
            bc.lower = [np.asarray(a) for a in z["bc%d_lower" % i]]
        elif kind == "ic":
            bc = IC(D, [lambda x: 0 * x], var=[["x"]])
            bc.input = np.asarray(z["bc%d_input" % i])
            bc.val = tf.constant(z["bc%d_val" % i])
        else:
            bc = dirichletBC(D, val=0.0, var=D.vars[0], target="upper")
            bc.input = np.asarray(z["bc%d_input" % i])
            v = np.asarray(z["bc%d_val" % i], np.float32)
            bc.val = tf.constant(v) if v.ndim else float(v)
        bcs.append(bc)
    lam_res, lam_bcs = gu.lambdas_of(z)
    model = CollocationSolverND(verbose=False)
    if lam_res is not None:
        flags = [bool(v) for v in z["res_adaptive"]] if "res_adaptive" in z else [True]
        dict_adaptive = {"residual": flags, "BCs": [True] + [False] * (len(bcs) - 1)}
        init_weights = {"residual": [tf.constant(lam_res)], "BCs": [tf.constant(lam_bcs[0])] + [None] * (len(bcs) - 1)}
        model.compile(ls, F_MODEL[res], D, bcs, isAdaptive=True, dict_adaptive=dict_adaptive,
                      init_weights=init_weights, g=gu.g_of(name), dist=dist)
    else:
        model.compile(ls, F_MODEL[res], D, bcs, dist=dist)
    model.u_model.set_flat(z["w0"])
    return model, z


def rel(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
