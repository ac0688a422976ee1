"""CPU restatement of the graph-mode Newton phase -- TEST INFRASTRUCTURE, NOT PRODUCT.

The reference's `fit(newton_eager=False)` hands the loss to
`tfp.optimizer.lbfgs_minimize(value_and_gradients_function, initial_position, max_iterations=newton_iter,
tolerance=1e-20)` (tensordiffeq/fit.py:65-80, optimizers.py:118-202).  tensorflow-probability (pinned 0.14.1 in
the reference's Pipfile.lock) is a third-party dependency that is NOT vendored under /root/reference and cannot
be installed here, so this file restates the PUBLISHED algorithm that call documents:

* L-BFGS with `num_correction_pairs = 10`, initial inverse Hessian gamma_k I with
  gamma_k = s_k.y_k / y_k.y_k (Nocedal & Wright, Algorithm 7.4/7.5), first direction -g;
* the Hager-Zhang line search (W. Hager, H. Zhang, "Algorithm 851: CG_DESCENT", ACM TOMS 32 (2006), the
  bracket / secant^2 / update / bisect routines L0-L3, B0-B3, S1-S4, U0-U3) with tfp's documented defaults:
  initial step 1, delta = 0.1 (sufficient decrease), sigma = 0.9 (curvature), epsilon = 1e-6 (approximate
  Wolfe threshold, f_lim = f(0) + epsilon |f(0)|), gamma = 0.66 (shrinkage), rho = 5 (expansion), theta = 0.5
  (bisection), at most 50 line-search iterations, a non-finite trial step shrunk by 0.1;
* stop: max |g_i| <= tolerance (1e-20 as the reference passes it), or the line search fails to find a step.

PARITY UNPINNED: nothing here was compared with outputs of tfp itself (it is absent); iteration-for-iteration
agreement with tfp 0.14.1 is therefore not claimed.  What IS pinned: the CUDA path (tdq_lbfgs_hz) against this
file (tests/test_gpu_hz.py), and this file against the defining properties of the method
(tests/test_hz_oracle.py: the accepted steps satisfy the (approximate) Wolfe conditions, exact minimiser of a
convex quadratic, monotone decrease on the Burgers golden start point).
"""
import numpy as np

F = np.float32

DELTA, SIGMA, EPSILON, GAMMA, RHO, THETA = 0.1, 0.9, 1e-6, 0.66, 5.0, 0.5
MAX_LS = 50
HISTORY = 10


class LineSearch:
    """phi(c) = f(x + c d) along one direction; counts evaluations, remembers every trial"""

    def __init__(self, opfunc, x, d, f0, g0):
        self.opfunc, self.x, self.d = opfunc, x, d
        self.f0 = F(f0)
        self.df0 = F(np.sum(g0 * d))
        self.f_lim = F(self.f0 + F(EPSILON) * abs(self.f0))
        self.n_eval = 0
        self.best = None      # (c, f, df, g) of the accepted step

    def phi(self, c):
        c = F(c)
        f, g = self.opfunc(self.x + c * self.d)
        self.n_eval += 1
        return c, F(f), F(np.sum(g * self.d)), g

    def wolfe(self, p):
        c, f, df, _ = p
        exact = f <= self.f0 + F(DELTA) * c * self.df0 and df >= F(SIGMA) * self.df0
        approx = f <= self.f_lim and F(2 * DELTA - 1) * self.df0 >= df >= F(SIGMA) * self.df0
        return bool(exact or approx)


def _u3(ls, a, c, budget):
    """U3 entered from the bracketing phase: f'(c) < 0 and f(c) > f_lim on [a = 0, c]"""
    ah, bh = a, c
    while budget[0] > 0:
        budget[0] -= 1
        d = ls.phi((1 - THETA) * ah[0] + THETA * bh[0])
        if ls.wolfe(d):
            return ah, bh, d
        if d[2] >= 0:
            return ah, d, None
        if d[1] <= ls.f_lim:
            ah = d
        else:
            bh = d
    return ah, bh, None


def _update(ls, a, b, c, budget):
    """U0-U3: shrink [a, b] with the trial point c (a, b, c are (c, f, df, g) tuples).  Returns (a, b, accepted)"""
    if not (a[0] < c[0] < b[0]):
        return a, b, None                                   # U0
    if ls.wolfe(c):
        return a, b, c
    if c[2] >= 0:
        return a, c, None                                   # U1
    if c[1] <= ls.f_lim:
        return c, b, None                                   # U2
    return _u3(ls, a, c, budget)                            # U3


def _secant(a, b):
    den = b[2] - a[2]
    if den == 0:
        return F(0.5) * (a[0] + b[0])
    return (a[0] * b[2] - b[0] * a[2]) / den


def hager_zhang(ls, initial_step=1.0):
    """returns the accepted trial (c, f, df, g) or None when the search fails"""
    budget = [MAX_LS]
    c = ls.phi(initial_step)
    while not (np.isfinite(c[1]) and np.isfinite(c[2])) and budget[0] > 0:
        budget[0] -= 1
        c = ls.phi(c[0] * F(0.1))
    if ls.wolfe(c):
        return c
    zero = (F(0), ls.f0, ls.df0, None)
    # ---- bracket B0-B3
    last_ok = zero                                           # largest c_i so far with f(c_i) <= f_lim
    a = b = None
    while budget[0] > 0:
        if c[2] >= 0:                                        # B1
            a, b = last_ok, c
            break
        if c[1] > ls.f_lim:                                  # B2: U3 on [0, c]
            a, b, acc = _u3(ls, zero, c, budget)
            if acc is not None:
                return acc
            break
        last_ok = c                                          # B3
        budget[0] -= 1
        c = ls.phi(c[0] * F(RHO))
        if ls.wolfe(c):
            return c
    if a is None:
        return None
    # ---- L1-L3
    while budget[0] > 0:
        budget[0] -= 1
        width = b[0] - a[0]
        # secant^2 (S1-S4)
        cs = ls.phi(_secant(a, b)) if a[0] < _secant(a, b) < b[0] else None
        A, B = a, b
        if cs is not None:
            A, B, acc = _update(ls, a, b, cs, budget)
            if acc is not None:
                return acc
            c2 = None
            if cs is B:
                c2 = _secant(b, B)
            elif cs is A:
                c2 = _secant(a, A)
            if c2 is not None and A[0] < c2 < B[0]:
                p2 = ls.phi(c2)
                A, B, acc = _update(ls, A, B, p2, budget)
                if acc is not None:
                    return acc
        a, b = A, B
        if b[0] - a[0] > F(GAMMA) * width:                   # L2: not enough shrinkage -> bisect
            m = ls.phi(F(0.5) * (a[0] + b[0]))
            a, b, acc = _update(ls, a, b, m, budget)
            if acc is not None:
                return acc
        if b[0] - a[0] <= 0:
            break
    return None


def two_loop(g, S, Y):
    """-H g with H from the stored pairs, H0 = gamma I (Nocedal & Wright Alg. 7.4)"""
    if not S:
        return -g
    q = g.copy()
    al = []
    for s, y in zip(reversed(S), reversed(Y)):
        rho = F(1) / F(np.sum(y * s))
        a = rho * F(np.sum(s * q))
        al.append(a)
        q = q - a * y
    gamma = F(np.sum(S[-1] * Y[-1])) / F(np.sum(Y[-1] * Y[-1]))
    r = gamma * q
    for (s, y), a in zip(zip(S, Y), reversed(al)):
        rho = F(1) / F(np.sum(y * s))
        be = rho * F(np.sum(y * r))
        r = r + (a - be) * s
    return -r


def lbfgs_hz(opfunc, x0, max_iter, tolerance=1e-20, history=HISTORY):
    """opfunc(x) -> (f, g), float32 numpy.  Returns dict(x, f_hist [one entry per accepted step, first = f(x0)],
    steps, n_eval, converged, failed)."""
    x = np.array(x0, F)
    f, g = opfunc(x)
    f = F(f)
    n_eval = 1
    f_hist, steps = [f], []
    S, Y = [], []
    converged = bool(np.max(np.abs(g)) <= tolerance)
    failed = False
    it = 0
    while it < max_iter and not converged and not failed:
        it += 1
        d = two_loop(g, S, Y)
        ls = LineSearch(opfunc, x, d, f, g)
        acc = hager_zhang(ls)
        n_eval += ls.n_eval
        if acc is None:
            failed = True
            break
        c, f_new, _, g_new = acc
        s = c * d
        y = g_new - g
        x = x + s
        if F(np.sum(y * s)) > 0:                 # (an approximate-Wolfe step may lack curvature: the pair is skipped)
            if len(S) == history:
                S.pop(0)
                Y.pop(0)
            S.append(s)
            Y.append(y)
        f, g = f_new, g_new
        f_hist.append(f)
        steps.append(float(c))
        converged = bool(np.max(np.abs(g)) <= tolerance)
    return {"x": x, "f_hist": np.array(f_hist, F), "steps": steps, "n_eval": n_eval, "converged": converged, "failed": failed}
