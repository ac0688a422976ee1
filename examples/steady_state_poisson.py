"""2-D steady-state Poisson with an analytic solution -- the tensordiffeq example
(examples/steady-state.py in the reference) written against tensordiffeq_b200: only the import lines
change, the problem definition, f_model and the compile / fit / predict calls are the reference's."""
import math

import numpy as np

import tensordiffeq_b200 as tdq
from tensordiffeq_b200.boundaries import *          # DomainND, dirichletBC, np, tf
from tensordiffeq_b200.models import CollocationSolverND
from tensordiffeq_b200.utils import constant

sin = tf.math.sin

Domain = DomainND(["x", "y"])
Domain.add("x", [-1.0, 1.0], 101)
Domain.add("y", [-1.0, 1.0], 101)
N_f = 10000
Domain.generate_collocation_points(N_f)


def f_model(u_model, x, y):
    u = u_model(tf.concat([x, y], 1))
    u_x = tf.gradients(u, x)[0]
    u_y = tf.gradients(u, y)[0]
    u_xx = tf.gradients(u_x, x)[0]
    u_yy = tf.gradients(u_y, y)[0]
    a1, a2, ksq, pi = constant(1.0), constant(4.0), constant(1.0), constant(math.pi)
    forcing = - (a1 * pi) ** 2 * sin(a1 * pi * x) * sin(a2 * pi * y) - \
        (a2 * pi) ** 2 * sin(a1 * pi * x) * sin(a2 * pi * y) + ksq * sin(a1 * pi * x) * sin(a2 * pi * y)
    return u_xx + u_yy + ksq * u - forcing


BCs = [dirichletBC(Domain, val=0.0, var=v, target=t) for v in ("x", "y") for t in ("upper", "lower")]
model = CollocationSolverND()
model.compile([2, 50, 50, 50, 50, 1], f_model, Domain, BCs)
model.fit(tf_iter=int(__import__("os").environ.get("TDQ_ADAM", "10000")), newton_iter=int(__import__("os").environ.get("TDQ_NEWTON", "10000")))

x = Domain.domaindict[0]["xlinspace"]
y = Domain.domaindict[1]["ylinspace"]
X, Y = np.meshgrid(x, y)
X_star = np.hstack((X.flatten()[:, None], Y.flatten()[:, None]))
u_star = (np.sin(math.pi * X) * np.sin(4 * math.pi * Y)).flatten()[:, None]
u_pred, f_u_pred = model.predict(X_star)
print("Error u: %e" % tdq.helpers.find_L2_error(u_pred, u_star))
