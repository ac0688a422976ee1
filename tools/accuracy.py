#!/usr/bin/env python
"""End-to-end solution quality: the reference's two headline examples trained through the public API with the
examples' own budget (10 k Adam + 10 k eager L-BFGS, examples/burgers-new.py:40-41, examples/AC-SA.py:65) and compared
with the solution fields the reference ships (tests/golden/{burgers,ac}_solution.npz, taken from
examples/burgers_shock.mat and examples/AC.mat) by tdq.helpers.find_L2_error, exactly as the example scripts do
(burgers-new.py:48-68, AC-SA.py:68-89).  Also timed.  python tools/accuracy.py [burgers] [ac_sa] [--adam N --newton N]"""
import argparse
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tensordiffeq_b200 as tdq
from tensordiffeq_b200 import tf
from tensordiffeq_b200.boundaries import DomainND, IC, dirichletBC, periodicBC
from tensordiffeq_b200.models import CollocationSolverND


def report(name, model, Domain, sol, t0, t1, t2):
    x = Domain.domaindict[0]['xlinspace']
    t = Domain.domaindict[1]["tlinspace"]
    X, T = np.meshgrid(x, t)
    X_star = np.hstack((X.flatten()[:, None], T.flatten()[:, None]))
    u_star = sol["u"].T.flatten()[:, None]
    u_pred, f_pred = model.predict(X_star)
    err = tdq.helpers.find_L2_error(u_pred, u_star)
    lb = getattr(model, "lbfgs_history", [])
    print("%-8s Error u: %e   (Adam %d steps %.1f s, final loss %.3e; L-BFGS %d evaluations %.1f s, final loss %s; "
          "mean |f| on the grid %.3e)" % (name, err, len(model.loss_history), t1 - t0, model.loss_history[-1] if model.loss_history else float("nan"),
                                          len(lb), t2 - t1, "%.3e" % lb[-1] if lb else "-", float(np.abs(f_pred).mean())), flush=True)
    return err


def burgers(n_adam, n_newton):
    sol = np.load(os.path.join(ROOT, "tests", "golden", "burgers_solution.npz"))
    Domain = DomainND(["x", "t"], time_var='t')
    Domain.add("x", [-1.0, 1.0], 256)
    Domain.add("t", [0.0, 1.0], 100)
    Domain.generate_collocation_points(10000)
    init = IC(Domain, [lambda x: -np.sin(x * math.pi)], var=[['x']])
    BCs = [init, dirichletBC(Domain, val=0.0, var='x', target="upper"), dirichletBC(Domain, val=0.0, var='x', target="lower")]

    def f_model(u_model, x, t):
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)
        u_xx = tf.gradients(u_x, x)
        u_t = tf.gradients(u, t)
        return u_t + u * u_x - (0.01 / tf.constant(math.pi)) * u_xx

    model = CollocationSolverND(verbose=False)
    model.compile([2, 20, 20, 20, 20, 20, 20, 20, 20, 1], f_model, Domain, BCs)
    t0 = time.time()
    model.fit(tf_iter=n_adam)
    model._engine().sync()
    t1 = time.time()
    model.fit(tf_iter=0, newton_iter=n_newton)
    model._engine().sync()
    return report("burgers", model, Domain, sol, t0, t1, time.time())


def ac_sa(n_adam, n_newton):
    sol = np.load(os.path.join(ROOT, "tests", "golden", "ac_solution.npz"))
    Domain = DomainND(["x", "t"], time_var='t')
    Domain.add("x", [-1.0, 1.0], 512)
    Domain.add("t", [0.0, 1.0], 201)
    N_f = 50000
    Domain.generate_collocation_points(N_f)

    def deriv_model(u_model, x, t):
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)[0]
        return u, u_x

    BCs = [IC(Domain, [lambda x: x ** 2 * np.cos(math.pi * x)], var=[['x']]), periodicBC(Domain, ['x'], [deriv_model])]

    def f_model(u_model, x, t):
        u = u_model(tf.concat([x, t], 1))
        u_x = tf.gradients(u, x)
        u_xx = tf.gradients(u_x, x)
        u_t = tf.gradients(u, t)
        c1 = tdq.utils.constant(.0001)
        c2 = tdq.utils.constant(5.0)
        return u_t - c1 * u_xx + c2 * u * u * u - c2 * u

    dict_adaptive = {"residual": [True], "BCs": [True, False]}
    init_weights = {"residual": [tf.random.uniform([N_f, 1])], "BCs": [100 * tf.random.uniform([512, 1]), None]}
    model = CollocationSolverND(verbose=False)
    model.compile([2, 128, 128, 128, 128, 1], f_model, Domain, BCs, isAdaptive=True, dict_adaptive=dict_adaptive,
                  init_weights=init_weights)
    t0 = time.time()
    model.fit(tf_iter=n_adam)
    model._engine().sync()
    t1 = time.time()
    model.fit(tf_iter=0, newton_iter=n_newton)
    model._engine().sync()
    return report("ac_sa", model, Domain, sol, t0, t1, time.time())


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("cases", nargs="*", default=["burgers", "ac_sa"])
    ap.add_argument("--adam", type=int, default=10000)
    ap.add_argument("--newton", type=int, default=10000)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    np.random.seed(a.seed)
    for c in a.cases:
        {"burgers": burgers, "ac_sa": ac_sa}[c](a.adam, a.newton)
