"""CPU: the restatement of the graph-mode Newton phase (oracle/hz_lbfgs.py: L-BFGS, 10 pairs, Hager-Zhang line
search with tfp's documented defaults; fit.py:65-80) against the defining properties of the method.  tfp itself is
not available, so there are no reference numbers to compare with (the oracle's header says "parity unpinned")."""
import numpy as np

from oracle import hz_lbfgs as hz
from oracle import pinn_oracle as po
from tests import golden_util as gu

F = np.float32


def test_quadratic_exact_minimiser_and_wolfe_steps():
    rs = np.random.RandomState(0)
    n = 40
    Q = rs.standard_normal((n, n))
    A = (Q @ Q.T / n + np.eye(n) * 0.5).astype(np.float64)
    b = rs.standard_normal(n)

    def op(x):
        x64 = x.astype(np.float64)
        return F(0.5 * x64 @ A @ x64 - b @ x64), (A @ x64 - b).astype(F)
    r = hz.lbfgs_hz(op, np.zeros(n, F), 200, tolerance=1e-5)
    assert r["converged"] and not r["failed"]
    assert np.allclose(r["x"], np.linalg.solve(A, b), atol=2e-4)
    assert np.all(np.diff(r["f_hist"]) <= 1e-6)


def test_every_accepted_step_satisfies_wolfe():
    def op(x):                                   # Rosenbrock in 6 dimensions
        x = x.astype(np.float64)
        f = np.sum(100 * (x[1:] - x[:-1] ** 2) ** 2 + (1 - x[:-1]) ** 2)
        g = np.zeros_like(x)
        g[:-1] += -400 * x[:-1] * (x[1:] - x[:-1] ** 2) - 2 * (1 - x[:-1])
        g[1:] += 200 * (x[1:] - x[:-1] ** 2)
        return F(f), g.astype(F)
    x = np.full(6, -1.2, F)
    f, g = op(x)
    f_start = f
    S, Y = [], []
    for _ in range(40):
        d = hz.two_loop(g, S, Y)
        ls = hz.LineSearch(op, x, d, f, g)
        acc = hz.hager_zhang(ls)
        assert acc is not None and ls.wolfe(acc) and ls.n_eval <= hz.MAX_LS + 2
        c, f_new, _, g_new = acc
        S.append(c * d)
        Y.append(g_new - g)
        S, Y = S[-hz.HISTORY:], Y[-hz.HISTORY:]
        x, f, g = x + c * d, f_new, g_new
    assert f < 0.05 * f_start


def test_burgers_golden_start_point_decreases():
    z = gu.load("burgers")
    prob = gu.problem_of("burgers", z)

    def op(x):
        r = po.loss_and_grads(prob, x)
        return F(r["loss"]), r["grad"]
    r = hz.lbfgs_hz(op, z["w_after_adam"], 12)
    assert not r["failed"] and len(r["f_hist"]) == 13
    assert np.all(np.diff(r["f_hist"]) < 0)
    assert r["f_hist"][-1] < 0.7 * r["f_hist"][0]
