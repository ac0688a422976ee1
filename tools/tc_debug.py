"""Developer check: tensor-core kernel vs the FP32 FMA kernel (TDQ_DBG=24) on the Allen-Cahn 128x4 net."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tensordiffeq_b200 import _lib
from tensordiffeq_b200 import trace as T
from tensordiffeq_b200.models import _trace, _integrand
from tensordiffeq_b200 import tf


def f_ac(u_model, x, t):
    u = u_model(tf.concat([x, t], 1))
    u_x = tf.gradients(u, x)[0]
    u_xx = tf.gradients(u_x, x)[0]
    u_t = tf.gradients(u, t)[0]
    return u_t - 0.0001 * u_xx + 5.0 * u ** 3 - 5.0 * u


LS = [2, 128, 128, 128, 128, 1]


def run(n, dbg):
    os.environ["TDQ_DBG"] = str(dbg)
    rs = np.random.RandomState(0)
    X = np.stack([rs.uniform(-1, 1, n), rs.uniform(0, 1, n)], 1).astype(np.float32)
    eng = _lib.Engine(LS)
    fe = T.as_sym(_trace(f_ac, 2))
    ops, slots, orders = T.lower({"out": _integrand(fe, False, None), "f": fe}, 2)
    t = eng.add_term(2, orders, ops, slots["out"], slots["f"], 0)
    eng.set_points(t, X, n)
    rs = np.random.RandomState(1)
    P = sum((a + 1) * b for a, b in zip(LS[:-1], LS[1:]))
    eng.set_params((rs.randn(P) * 0.1).astype(np.float32))
    L, G, _ = eng.loss_grad()
    f = eng.get_residual(t)
    eng.close()
    return L, G, f


def layers(G):
    out, o = [], 0
    for a, b in zip(LS[:-1], LS[1:]):
        out.append((G[o:o + a * b].reshape(a, b), G[o + a * b:o + a * b + b]))
        o += a * b + b
    return out


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    L1, G1, f1 = run(n, 0)
    L0, G0, f0 = run(n, 24)
    print("loss tc %.8g  fma %.8g" % (L1, L0))
    print("residual rel %.3e" % (np.linalg.norm(f1 - f0) / np.linalg.norm(f0)), f1[:4], f0[:4])
    for i, ((W1, b1), (W0, b0)) in enumerate(zip(layers(G1), layers(G0))):
        print("layer %d  dW rel %.3e (|tc| %.3e |fma| %.3e)   db rel %.3e (|tc| %.3e |fma| %.3e)" % (
            i + 1, np.linalg.norm(W1 - W0) / np.linalg.norm(W0), np.linalg.norm(W1), np.linalg.norm(W0),
            np.linalg.norm(b1 - b0) / np.linalg.norm(b0), np.linalg.norm(b1), np.linalg.norm(b0)))
