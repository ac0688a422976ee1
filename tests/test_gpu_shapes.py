"""GPU: kernel instantiations the BASELINE configs do not reach -- 3 input dimensions with second derivatives in
each (C = 7 channels, 2-point tiles), a single input dimension, padded widths 64 and 128 with odd layer sizes,
a Neumann boundary term, first-order-only residuals -- each against the oracle (nested autograd) at rel 1e-4."""
import math

import numpy as np
import pytest
import torch

from oracle import pinn_oracle as po
from tensordiffeq_b200 import tf
from tensordiffeq_b200.boundaries import DomainND, IC, dirichletBC, FunctionNeumannBC
from tensordiffeq_b200.models import CollocationSolverND

pytestmark = pytest.mark.gpu
TOL = 1e-4


def rel(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def flat_of(grads, n_net):
    return np.concatenate([g.ravel() for g in grads[:n_net]])


def check(model, prob, w0, n_layers):
    model.u_model.set_flat(w0)
    loss, grads = model.grad()
    r = po.loss_and_grads(prob, w0)
    assert abs(float(loss) - float(r["loss"])) <= TOL * abs(float(r["loss"])), (float(loss), float(r["loss"]))
    assert rel(flat_of(grads, 2 * (n_layers - 1)), r["grad"]) < TOL
    f = model._engine().get_residual(model._terms[-1]["tid"])
    assert rel(f, r["f"]) < TOL


@pytest.mark.parametrize("ls", [[3, 40, 40, 1], [3, 100, 90, 80, 1]])
def test_poisson_3d(ls):
    D = DomainND(["x", "y", "z"])
    for v in "xyz":
        D.add(v, [0.0, 1.0], 5)
    D.generate_collocation_points(777)

    def f_model(u_model, x, y, z):
        u = u_model(tf.concat([x, y, z], 1))
        u_xx = tf.gradients(tf.gradients(u, x)[0], x)[0]
        u_yy = tf.gradients(tf.gradients(u, y)[0], y)[0]
        u_zz = tf.gradients(tf.gradients(u, z)[0], z)[0]
        return u_xx + u_yy + u_zz + tf.math.sin(x) * tf.math.cos(y) * z

    def f_oracle(u_model, x, y, z):
        u = u_model(torch.cat([x, y, z], 1))
        g = po.gradients
        return g(g(u, x), x) + g(g(u, y), y) + g(g(u, z), z) + torch.sin(x) * torch.cos(y) * z

    bcs = [dirichletBC(D, val=0.0, var="x", target="upper"), dirichletBC(D, val=0.5, var="z", target="lower")]
    model = CollocationSolverND(verbose=False)
    model.compile(ls, f_model, D, bcs)
    terms = [po.Term("dirichlet", X=b.input, val=np.float32(b.val)) for b in bcs]
    prob = po.Problem(ls, f_oracle, D.X_f, terms)
    check(model, prob, po.glorot_normal_flat(ls, seed=4, bias_std=0.1), len(ls))


def test_ode_1d():
    ls = [1, 33, 33, 1]
    D = DomainND(["x"])
    D.add("x", [0.0, 2.0], 9)
    D.generate_collocation_points(300)

    def f_model(u_model, x):
        u = u_model(tf.concat([x], 1))
        u_x = tf.gradients(u, x)[0]
        u_xx = tf.gradients(u_x, x)[0]
        return u_xx + 2.0 * u_x + u * u - tf.math.exp(-x)

    def f_oracle(u_model, x):
        u = u_model(torch.cat([x], 1))
        u_x = po.gradients(u, x)
        return po.gradients(u_x, x) + 2.0 * u_x + u * u - torch.exp(-x)

    model = CollocationSolverND(verbose=False)
    model.compile(ls, f_model, D, [])
    prob = po.Problem(ls, f_oracle, D.X_f, [])
    check(model, prob, po.glorot_normal_flat(ls, seed=5, bias_std=0.2), len(ls))


def test_first_order_transport_width_64():
    ls = [2, 50, 50, 50, 50, 1]              # examples/steady-state.py width
    D = DomainND(["x", "t"], time_var="t")
    D.add("x", [-1.0, 1.0], 16)
    D.add("t", [0.0, 1.0], 8)
    D.generate_collocation_points(1234)

    def f_model(u_model, x, t):
        u = u_model(tf.concat([x, t], 1))
        return tf.gradients(u, t)[0] + 0.7 * tf.gradients(u, x)[0]

    def f_oracle(u_model, x, t):
        u = u_model(torch.cat([x, t], 1))
        return po.gradients(u, t) + 0.7 * po.gradients(u, x)

    init = IC(D, [lambda x: np.sin(math.pi * x)], var=[["x"]])
    model = CollocationSolverND(verbose=False)
    model.compile(ls, f_model, D, [init])
    prob = po.Problem(ls, f_oracle, D.X_f, [po.Term("ic", X=init.input, val=np.asarray(init.val, np.float32))])
    check(model, prob, po.glorot_normal_flat(ls, seed=6, bias_std=0.1), len(ls))


def test_neumann_term():
    """FunctionNeumannBC (boundaries.py:103-156, models.py:120-130): MSE(val, d u / d x at the boundary)"""
    ls = [2, 24, 24, 1]
    np.random.seed(3)
    D = DomainND(["x", "t"], time_var="t")
    D.add("x", [0.0, 1.0], 12)
    D.add("t", [0.0, 1.0], 7)
    D.generate_collocation_points(200)

    def deriv(u_model, x, t):
        u = u_model(tf.concat([x, t], 1))
        return (tf.gradients(u, x)[0],)

    def f_model(u_model, x, t):
        u = u_model(tf.concat([x, t], 1))
        return tf.gradients(u, t)[0] - 0.1 * tf.gradients(tf.gradients(u, x)[0], x)[0]

    nbc = FunctionNeumannBC(D, fun=[lambda t: 0.3 * t], var=["x"], target="upper", deriv_model=[deriv], func_inputs=[["t"]])
    model = CollocationSolverND(verbose=False)
    model.compile(ls, f_model, D, [nbc])
    w0 = po.glorot_normal_flat(ls, seed=8, bias_std=0.1)
    model.u_model.set_flat(w0)
    loss, grads = model.grad()
    # oracle by hand: residual MSE + MSE(val, u_x(boundary points))
    w = torch.tensor(w0, requires_grad=True)
    net = po.UModel(w, ls)
    cols = [torch.tensor(D.X_f[:, j:j + 1], dtype=torch.float32, requires_grad=True) for j in range(2)]
    u = net(torch.cat(cols, 1))
    f = po.gradients(u, cols[1]) - 0.1 * po.gradients(po.gradients(u, cols[0]), cols[0])
    Xb = np.asarray(nbc.input[0], np.float64)[:, :, 0].T
    bcols = [torch.tensor(Xb[:, j:j + 1], dtype=torch.float32, requires_grad=True) for j in range(2)]
    ub = net(torch.cat(bcols, 1))
    ux = po.gradients(ub, bcols[0])
    val = torch.tensor(np.asarray(nbc.val, np.float32))
    ref = torch.mean(f ** 2) + torch.mean((val - ux) ** 2)
    (g,) = torch.autograd.grad(ref, w)
    assert abs(float(loss) - float(ref)) <= TOL * abs(float(ref))
    assert rel(flat_of(grads, 6), g.numpy()) < TOL
