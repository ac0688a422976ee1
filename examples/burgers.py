"""Burgers' equation, the reference's examples/burgers-new.py against tensordiffeq_b200 (the reference solution
file burgers_shock.mat is not shipped here; the script reports the training loss and the residual norm)."""
import math
import os

import numpy as np

import tensordiffeq_b200 as tdq
from tensordiffeq_b200.boundaries import *
from tensordiffeq_b200.models import CollocationSolverND

Domain = DomainND(["x", "t"], time_var='t')
Domain.add("x", [-1.0, 1.0], 256)
Domain.add("t", [0.0, 1.0], 100)
N_f = 10000
Domain.generate_collocation_points(N_f)


def func_ic(x):
    return -np.sin(x * math.pi)


init = IC(Domain, [func_ic], var=[['x']])
upper_x = dirichletBC(Domain, val=0.0, var='x', target="upper")
lower_x = dirichletBC(Domain, val=0.0, var='x', target="lower")
BCs = [init, upper_x, lower_x]


def f_model(u_model, x, t):
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)
    u_xx = tf.gradients(u_x, x)
    u_t = tf.gradients(u, t)
    f_u = u_t + u * u_x - (0.01 / tf.constant(math.pi)) * u_xx
    return f_u


layer_sizes = [2, 20, 20, 20, 20, 20, 20, 20, 20, 1]
model = CollocationSolverND()
model.compile(layer_sizes, f_model, Domain, BCs)
model.fit(tf_iter=int(os.environ.get("TDQ_ADAM", "10000")), newton_iter=int(os.environ.get("TDQ_NEWTON", "10000")))

x = Domain.domaindict[0]['xlinspace']
t = Domain.domaindict[1]["tlinspace"]
X, T = np.meshgrid(x, t)
X_star = np.hstack((X.flatten()[:, None], T.flatten()[:, None]))
u_pred, f_u_pred = model.predict(X_star)
print("final Adam loss %.4e, L-BFGS loss %.4e, |f|_2/sqrt(n) on the grid %.4e, u range [%.3f, %.3f]" % (
    model.loss_history[-1], model.lbfgs_history[-1], np.linalg.norm(f_u_pred) / math.sqrt(f_u_pred.size), u_pred.min(), u_pred.max()))
