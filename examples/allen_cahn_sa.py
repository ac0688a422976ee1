"""Self-adaptive Allen-Cahn, the reference's examples/AC-SA.py against tensordiffeq_b200."""
import math
import os

import numpy as np

import tensordiffeq_b200 as tdq
from tensordiffeq_b200.models import CollocationSolverND
from tensordiffeq_b200.boundaries import *

Domain = DomainND(["x", "t"], time_var='t')
Domain.add("x", [-1.0, 1.0], 512)
Domain.add("t", [0.0, 1.0], 201)
N_f = 50000
Domain.generate_collocation_points(N_f)


def func_ic(x):
    return x ** 2 * np.cos(math.pi * x)


def deriv_model(u_model, x, t):
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)[0]
    return u, u_x


init = IC(Domain, [func_ic], var=[['x']])
x_periodic = periodicBC(Domain, ['x'], [deriv_model])
BCs = [init, x_periodic]


def f_model(u_model, x, t):
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)
    u_xx = tf.gradients(u_x, x)
    u_t = tf.gradients(u, t)
    c1 = tdq.utils.constant(.0001)
    c2 = tdq.utils.constant(5.0)
    f_u = u_t - c1 * u_xx + c2 * u * u * u - c2 * u
    return f_u


dict_adaptive = {"residual": [True], "BCs": [True, False]}
init_weights = {"residual": [tf.random.uniform([N_f, 1])], "BCs": [100 * tf.random.uniform([512, 1]), None]}
model = CollocationSolverND()
model.compile([2, 128, 128, 128, 128, 1], f_model, Domain, BCs, isAdaptive=True, dict_adaptive=dict_adaptive,
              init_weights=init_weights)
model.fit(tf_iter=int(os.environ.get("TDQ_ADAM", "10000")), newton_iter=int(os.environ.get("TDQ_NEWTON", "10000")))
print("final Adam loss %.4e; lambda_residual range [%.3f, %.3f]" % (
    model.loss_history[-1], model.lambdas[0].numpy().min(), model.lambdas[0].numpy().max()))
