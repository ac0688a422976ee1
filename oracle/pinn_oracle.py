"""CPU oracle for the tensordiffeq hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

A plain restatement, in PyTorch CPU autograd, of what the reference computes
per training step (loss, residual, parameter / lambda gradients, the Keras-Adam
update, the eager L-BFGS iteration, DiscoveryModel).  Every function cites the
reference file:line it follows (paths relative to /root/reference).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module; tensordiffeq_b200/ never does.

PARITY PINNING.  TensorFlow cannot be installed in the build image, and the
reference's own tests hold no numeric assertion, so this oracle
is NOT pinned against outputs of real TensorFlow ("parity unpinned" in that
sense).  It IS pinned against the reference's own Python -- models.py, fit.py,
optimizers.py, boundaries.py executed unmodified on oracle/ref_shims (a torch
stand-in for the TF API): tests/golden/make_golden.py wrote tests/golden/*.npz
that way and tests/test_oracle_golden.py checks this module against them.

Derivatives are taken the way the reference takes them -- nested reverse-mode
autodiff (tf.gradients -> torch.autograd.grad(create_graph=True)) -- and NOT
with the Taylor-jet formulas the CUDA kernels use, so agreement between the two
is a real check.
"""
import math

import numpy as np
import torch

F32 = torch.float32


# --------------------------------------------------------------------------
# parameter layout + network   (tensordiffeq/utils.py:7-35, networks.py:10-20)
# --------------------------------------------------------------------------
def get_sizes(layer_sizes):
    """utils.py:32-35."""
    sizes_w = [layer_sizes[i] * layer_sizes[i - 1] for i in range(1, len(layer_sizes))]
    sizes_b = list(layer_sizes[1:])
    return sizes_w, sizes_b


def n_params(layer_sizes):
    sw, sb = get_sizes(layer_sizes)
    return sum(sw) + sum(sb)


def unflatten(w, layer_sizes):
    """utils.py:7-16: per layer, kernel [in,out] row-major then bias."""
    out, off = [], 0
    for i in range(1, len(layer_sizes)):
        fi, fo = layer_sizes[i - 1], layer_sizes[i]
        W = w[off:off + fi * fo].reshape(fi, fo)
        off += fi * fo
        b = w[off:off + fo]
        off += fo
        out.append((W, b))
    return out


def glorot_normal_flat(layer_sizes, seed=0, bias_std=0.0):
    """networks.py:13-19 kernel_initializer="glorot_normal" (keras: truncated
    normal, stddev sqrt(2/(fan_in+fan_out))/.8796..., cut at 2 sigma), bias 0."""
    rng = np.random.RandomState(seed)
    out = []
    for i in range(1, len(layer_sizes)):
        fi, fo = layer_sizes[i - 1], layer_sizes[i]
        std = math.sqrt(2.0 / (fi + fo)) / 0.87962566103423978
        W = rng.standard_normal((fi, fo))
        bad = np.abs(W) > 2
        while bad.any():
            W[bad] = rng.standard_normal(int(bad.sum()))
            bad = np.abs(W) > 2
        out.append((W * std).astype(np.float32).ravel())
        out.append((bias_std * rng.standard_normal(fo)).astype(np.float32))
    return np.concatenate(out)


class UModel:
    """networks.py:10-20: Dense(tanh) x (L-2), Dense(linear).  Inputs are cast to
    the parameter dtype on entry (keras autocast)."""

    def __init__(self, w_flat, layer_sizes):
        self.layer_sizes = list(layer_sizes)
        self.w = w_flat
        self.layers = unflatten(w_flat, layer_sizes)

    def __call__(self, X):
        a = X.to(self.w.dtype)
        n = len(self.layers)
        for i, (W, b) in enumerate(self.layers):
            a = a @ W + b
            if i < n - 1:
                a = torch.tanh(a)
        return a


# --------------------------------------------------------------------------
# loss primitives   (utils.py:38-45)
# --------------------------------------------------------------------------
def MSE(pred, actual, weights=None):
    if weights is not None:
        return torch.mean(torch.square(weights * (pred - actual)))
    return torch.mean(torch.square(pred - actual))


def g_MSE(pred, actual, g_lam):
    return torch.mean(g_lam * torch.square(pred - actual))


def gradients(y, x):
    """tf.gradients(y, x)[0]: d(sum y)/dx, itself differentiable."""
    (g,) = torch.autograd.grad(y.sum(), x, create_graph=True)
    return g


# --------------------------------------------------------------------------
# the residual callbacks of BASELINE.json's configs (user f_model)
# --------------------------------------------------------------------------
def f_allen_cahn(u_model, x, t):
    """examples/AC-SA.py:36-44."""
    u = u_model(torch.cat([x, t], 1))
    u_x = gradients(u, x)
    u_xx = gradients(u_x, x)
    u_t = gradients(u, t)
    return u_t - 0.0001 * u_xx + 5.0 * u * u * u - 5.0 * u


def f_burgers(u_model, x, t):
    """examples/burgers-new.py:26-32; 0.01/tf.constant(pi) is evaluated in fp32."""
    u = u_model(torch.cat([x, t], 1))
    u_x = gradients(u, x)
    u_xx = gradients(u_x, x)
    u_t = gradients(u, t)
    nu = float(np.float32(0.01) / np.float32(math.pi))
    return u_t + u * u_x - nu * u_xx


def f_poisson(u_model, x, y):
    """examples/steady-state-poisson.py:19-37."""
    u = u_model(torch.cat([x, y], 1))
    u_x = gradients(u, x)
    u_y = gradients(u, y)
    u_xx = gradients(u_x, x)
    u_yy = gradients(u_y, y)
    pi = float(np.float32(math.pi))
    forcing = -torch.sin(pi * x) * torch.sin(pi * y)
    return u_xx + u_yy - forcing


def f_cahn_hilliard(u_model, var, x, t):
    """examples/CH-discovery.py:20-35 (var = [g1, g2] trainable coefficients)."""
    g1, g2 = var[0], var[1]
    u = u_model(torch.cat([x, t], 1))
    u_x = gradients(u, x)
    u_t = gradients(u, t)
    u_xx = gradients(u_x, x)
    tmp = g1 * (u ** 3 - u) - g2 * u_xx
    tmp_x = gradients(tmp, x)
    tmp_xx = gradients(tmp_x, x)
    return u_t - tmp_xx


RESIDUALS = {
    "allen_cahn": f_allen_cahn,
    "burgers": f_burgers,
    "poisson": f_poisson,
    "cahn_hilliard": f_cahn_hilliard,
}


def deriv_u_ux(u_model, *cols):
    """examples/AC-SA.py:21-27 periodic deriv_model: (u, u_x)."""
    u = u_model(torch.cat(list(cols), 1))
    u_x = gradients(u, cols[0])
    return u, u_x


# --------------------------------------------------------------------------
# problem container
# --------------------------------------------------------------------------
class Term:
    """One BC/IC entry of `bcs` as the reference builds it (boundaries.py)."""

    def __init__(self, kind, X=None, val=None, lam=None, upper=None, lower=None,
                 deriv_model=None, var=None):
        self.kind = kind            # "ic" | "dirichlet" | "periodic"
        self.X = X                  # [n,d] float64 (boundaries.py:57-59,177-186)
        self.val = val              # [n,1] float32 or python scalar (dirichletBC.val)
        self.lam = lam              # [n,1] float32 or None (adaptive BC, models.py:114-116,133-135)
        self.upper = upper          # periodic: [n_var, d, n, 1] (boundaries.py:227-246)
        self.lower = lower
        self.deriv_model = deriv_model or [deriv_u_ux]
        self.var = var or ["x"]


class Problem:
    def __init__(self, layer_sizes, f_model, X_f, terms, lam_res=None, g=None, dtype=F32, res_adaptive=None):
        self.layer_sizes = list(layer_sizes)
        self.f_model = RESIDUALS[f_model] if isinstance(f_model, str) else f_model
        self.dtype = dtype
        # models.py:46-49: collocation columns explicitly cast to float32
        self.cols = [torch.tensor(np.asarray(X_f)[:, j:j + 1], dtype=dtype) for j in range(np.shape(X_f)[1])]
        self.terms = terms
        self.lam_res = lam_res
        self.g = g
        # dict_adaptive["residual"]: one flag per residual of a tuple-returning callback (models.py:166-168);
        # None = the single residual is adaptive iff a lambda vector is passed
        self.res_adaptive = res_adaptive


def _t(a, dtype, grad=False):
    t = torch.tensor(np.asarray(a), dtype=dtype)
    return t.requires_grad_(True) if grad else t


def update_loss(prob, w, lam_res=None, lam_bcs=None):
    """models.py:80-192.  w: flat torch parameter vector; lam_res: tensor or None;
    lam_bcs: {term index: tensor}.  Returns (loss_total, f_u_pred)."""
    u_model = UModel(w, prob.layer_sizes)
    loss_bcs = _bc_loss(prob, u_model, lam_bcs)
    loss_r, f_u_pred = _res_loss(prob, u_model, prob.cols, lam_res)
    return loss_r + loss_bcs, f_u_pred                       # models.py:190


def _bc_loss(prob, u_model, lam_bcs=None):
    """models.py:91-142: the sum of the BC / IC terms"""
    dt = prob.dtype
    lam_bcs = lam_bcs or {}
    loss_bcs = torch.zeros((), dtype=dt)
    for i, bc in enumerate(prob.terms):
        if bc.kind == "periodic":
            # models.py:100-111 incl. quirk Q2: iterate the *characters* of the variable name
            loss_bc = torch.zeros((), dtype=dt)
            for vi, dim in enumerate(bc.var):
                for j, lst in enumerate(dim):
                    for k, _tup in enumerate(lst):
                        up = [torch.tensor(c, dtype=torch.float64).requires_grad_(True) for c in bc.upper[vi]]
                        lo = [torch.tensor(c, dtype=torch.float64).requires_grad_(True) for c in bc.lower[vi]]
                        upper = bc.deriv_model[j](u_model, *up)[k]
                        lower = bc.deriv_model[j](u_model, *lo)[k]
                        loss_bc = loss_bc + MSE(upper, lower)
        elif bc.kind in ("ic", "dirichlet"):
            # models.py:113-118, 132-137
            pred = u_model(torch.tensor(np.asarray(bc.X), dtype=torch.float64))
            val = bc.val if isinstance(bc.val, torch.Tensor) else torch.tensor(np.asarray(bc.val), dtype=dt)
            if i in lam_bcs:
                loss_bc = MSE(pred, val, lam_bcs[i])
            else:
                loss_bc = MSE(pred, val)
        else:
            raise Exception("Boundary condition type is not acceptable")  # models.py:140
        loss_bcs = loss_bcs + loss_bc
    return loss_bcs


def _res_loss(prob, u_model, cols, lam_res=None):
    """models.py:144-188: the residual terms on the collocation columns `cols`"""
    dt = prob.dtype
    cols = [c.detach().clone().requires_grad_(True) for c in cols]
    f_u_preds = prob.f_model(u_model, *cols)  # models.py:157
    if not isinstance(f_u_preds, tuple):      # models.py:160-161
        f_u_preds = (f_u_preds,)
    zero = torch.zeros((), dtype=dt)
    loss_r = torch.zeros((), dtype=dt)
    for r, f_u_pred in enumerate(f_u_preds):  # models.py:164-188
        adaptive = lam_res is not None and (prob.res_adaptive is None or bool(prob.res_adaptive[r]))
        if adaptive:
            # every adaptive residual reads lambdas_map['residual'][0] (models.py:169; the increment at :181 is dead)
            if prob.g is not None:
                loss_r = loss_r + g_MSE(f_u_pred, zero, prob.g(lam_res))  # models.py:176-177
            else:
                loss_r = loss_r + MSE(f_u_pred, zero, lam_res)            # models.py:179
        else:
            loss_r = loss_r + MSE(f_u_pred, zero)                         # models.py:184-186
    return loss_r, f_u_preds[0]


def loss_and_grads(prob, w_np, lam_res_np=None, lam_bcs_np=None):
    """models.py:195-199 `grad()`: loss and d loss / d (weights ++ lambdas)."""
    dt = prob.dtype
    w = _t(w_np, dt, True)
    lam_res = _t(lam_res_np, dt, True) if lam_res_np is not None else None
    lam_bcs = {i: _t(v, dt, True) for i, v in (lam_bcs_np or {}).items()}
    loss, f = update_loss(prob, w, lam_res, lam_bcs)
    srcs = [w] + ([lam_res] if lam_res is not None else []) + [lam_bcs[i] for i in sorted(lam_bcs)]
    gs = torch.autograd.grad(loss, srcs, allow_unused=True)
    out = {"loss": loss.detach().numpy(), "grad": gs[0].numpy(), "f": f.detach().numpy().reshape(-1)}
    k = 1
    if lam_res is not None:
        out["lam_res_grad"] = gs[k].numpy()
        k += 1
    out["lam_bc_grads"] = {i: gs[k + n].numpy() for n, i in enumerate(sorted(lam_bcs))}
    return out


def loss_and_grads_chunked(prob, w_np, chunk=100000):
    """loss_and_grads() for a problem without lambdas whose collocation set is too large for one autograd
    graph: the residual mean is accumulated over chunks of points (each weighted n_chunk / N_f), the BC terms
    are added once.  Accumulation in float64; same terms as update_loss (models.py:80-192)."""
    dt = prob.dtype
    n = prob.cols[0].shape[0]
    w = _t(w_np, dt, True)
    loss = float(0.0)
    grad = np.zeros(np.shape(w_np), np.float64)
    lb = _bc_loss(prob, UModel(w, prob.layer_sizes))
    if lb.requires_grad:
        (g,) = torch.autograd.grad(lb, [w])
        grad += g.numpy().astype(np.float64)
    loss += float(lb.detach())
    for s in range(0, n, chunk):
        cols = [c[s:s + chunk] for c in prob.cols]
        lr, _ = _res_loss(prob, UModel(w, prob.layer_sizes), cols)
        (g,) = torch.autograd.grad(lr, [w])
        wgt = cols[0].shape[0] / n
        loss += float(lr.detach()) * wgt
        grad += g.numpy().astype(np.float64) * wgt
    return {"loss": np.float64(loss), "grad": grad}


def predict(prob, w_np, X_star):
    """models.py:274-284: u from the float64 grid (autocast), f from float32 columns."""
    dt = prob.dtype
    u_model = UModel(_t(w_np, dt), prob.layer_sizes)
    u = u_model(torch.tensor(np.asarray(X_star), dtype=torch.float64))
    cols = [torch.tensor(np.asarray(X_star)[:, j:j + 1], dtype=dt).requires_grad_(True)
            for j in range(np.shape(X_star)[1])]
    f = prob.f_model(u_model, *cols)
    return u.detach().numpy(), f.detach().numpy()


# --------------------------------------------------------------------------
# Keras Adam   (models.py:35-36; TF 2.6 ResourceApplyAdam, third-party arithmetic)
# --------------------------------------------------------------------------
class KerasAdam:
    """alpha = lr*sqrt(1-b2^t)/(1-b1^t); m += (g-m)(1-b1); v += (g^2-v)(1-b2);
    x -= alpha*m/(sqrt(v)+eps).  Defaults are the reference's (lr .005, beta_1 .99)."""

    def __init__(self, lr=0.005, beta_1=0.99, beta_2=0.999, epsilon=1e-7):
        self.lr, self.b1, self.b2, self.eps = lr, beta_1, beta_2, epsilon
        self.t = 0
        self.m = {}
        self.v = {}

    def begin_step(self):
        self.t += 1
        f = np.float32
        b1p = f(np.power(f(self.b1), f(self.t)))
        b2p = f(np.power(f(self.b2), f(self.t)))
        self.alpha = f(f(self.lr) * np.sqrt(f(1) - b2p) / (f(1) - b1p))

    def apply(self, key, x, g):
        f = np.float32
        if key not in self.m:
            self.m[key] = np.zeros_like(x)
            self.v[key] = np.zeros_like(x)
        m, v = self.m[key], self.v[key]
        m += (g - m) * (f(1) - f(self.b1))
        v += (g * g - v) * (f(1) - f(self.b2))
        x -= (m * self.alpha) / (np.sqrt(v) + f(self.eps))
        return x


def adam_train(prob, w_np, n_steps, lam_res_np=None, lam_bcs_np=None, opt_w=None, opt_lam=None):
    """fit.py:83-113: per step, Adam descent on the weights and Adam *ascent*
    (negated gradients, fit.py:101,104) on every lambda."""
    opt_w = opt_w or KerasAdam()
    opt_lam = opt_lam or KerasAdam()
    w = np.array(w_np, np.float32)
    lam_res = None if lam_res_np is None else np.array(lam_res_np, np.float32)
    lam_bcs = {i: np.array(v, np.float32) for i, v in (lam_bcs_np or {}).items()}
    hist = []
    for _ in range(n_steps):
        r = loss_and_grads(prob, w, lam_res, lam_bcs)
        hist.append(np.float32(r["loss"]))
        opt_w.begin_step()
        w = opt_w.apply("w", w, r["grad"])
        if lam_res is not None or lam_bcs:
            opt_lam.begin_step()
            if lam_res is not None:
                lam_res = opt_lam.apply("res", lam_res, -r["lam_res_grad"])
            for i in sorted(lam_bcs):
                lam_bcs[i] = opt_lam.apply(("bc", i), lam_bcs[i], -r["lam_bc_grads"][i])
    return {"w": w, "lam_res": lam_res, "lam_bcs": lam_bcs, "loss_hist": np.array(hist, np.float32)}


# --------------------------------------------------------------------------
# eager L-BFGS   (optimizers.py:214-402; called from fit.py:43-51)
# --------------------------------------------------------------------------
def eager_lbfgs(opfunc, x0, max_iter, learning_rate=0.8, n_correction=50):
    """opfunc(x) -> (f, g) on numpy float32.  History 50, no line search, fixed step
    0.8 (first step min(1, 1/|g|_1)), stop tests as written incl. quirk Q8
    (`tf.abs(f, f_old)` tests |f| < tolX).  Naming trap kept: old_dirs holds s,
    old_stps holds y (optimizers.py:285-286)."""
    f32 = np.float32
    max_eval = max_iter * 1.25
    tol_fun = tol_x = 1e-12
    x = np.array(x0, f32)
    f, g = opfunc(x)
    f_hist = [f32(f)]
    n_eval = 1
    if np.sum(np.abs(g)) <= tol_fun:                      # :241-244
        return x, f_hist, n_eval
    n_iter = 0
    old_dirs, old_stps = [], []
    Hdiag = f32(1)
    d = None
    t = None
    g_old = g
    for _epoch in range(max_iter):
        # :252-256 -- evaluated BEFORE the counter moves: on the 2nd pass (state.nIter == 1)
        # t is re-derived from the *new* gradient, and `s = d * t` below then uses that value,
        # not the step length actually taken (quirk Q12); from the 3rd pass on t = learningRate.
        if n_iter == 1:
            t = f32(min(1.0, 1.0 / float(np.sum(np.abs(g)))))
        else:
            t = f32(learning_rate)
        n_iter += 1
        if n_iter == 1:                                   # :263-267
            d = -g
            old_dirs, old_stps = [], []
            Hdiag = f32(1)
        else:
            y = g - g_old                                 # :273-275
            s = d * t
            ys = f32(np.sum(y * s))
            if ys > 1e-10:                                # :277-289
                if len(old_dirs) == n_correction:
                    del old_dirs[0]
                    del old_stps[0]
                old_dirs.append(s)
                old_stps.append(y)
                Hdiag = ys / f32(np.sum(y * y))
            k = len(old_dirs)
            ro = [f32(1) / f32(np.sum(old_stps[i] * old_dirs[i])) for i in range(k)]  # :296-298
            al = [f32(0)] * k
            q = -g
            for i in range(k - 1, -1, -1):                # :304-307
                al[i] = f32(np.sum(old_dirs[i] * q)) * ro[i]
                q = q - al[i] * old_stps[i]
            r = q * Hdiag                                 # :310
            for i in range(k):                            # :311-313
                be_i = f32(np.sum(old_stps[i] * r)) * ro[i]
                r = r + (al[i] - be_i) * old_dirs[i]
            d = r
        g_old = g
        gtd = f32(np.sum(g * d))                          # :325
        if gtd > -tol_x:                                  # :328-330
            break
        if n_iter == 1:                                   # :333-337
            t = f32(min(1.0, 1.0 / float(np.sum(np.abs(g)))))
        else:
            t = f32(learning_rate)
        x = x + t * d                                     # :339
        if n_iter != max_iter:                            # :341-345
            f, g = opfunc(x)
        f_hist.append(f32(f))
        n_eval += 1
        if n_iter == max_iter:
            break
        if n_eval >= max_eval:
            break
        if np.sum(np.abs(g)) <= tol_fun:
            break
        if np.sum(np.abs(d * t)) <= tol_x:
            break
        if abs(f) < tol_x:                                # :377, quirk Q8
            break
    return x, f_hist, n_eval


def lbfgs_train(prob, w_np, max_iter, lam_res_np=None, lam_bcs_np=None):
    """fit.py:43-51 + models.py:260-272: lambdas frozen but applied in the loss."""
    def opfunc(x):
        r = loss_and_grads(prob, x, lam_res_np, lam_bcs_np)
        return np.float32(r["loss"]), r["grad"]
    return eager_lbfgs(opfunc, w_np, max_iter, 0.8)


# --------------------------------------------------------------------------
# DiscoveryModel   (models.py:295-369)
# --------------------------------------------------------------------------
def discovery_loss_and_grads(layer_sizes, f_model, X_np, u_np, w_np, vars_np, col_weights_np=None, dtype=F32):
    """models.py:314-328: MSE(u_model(X), u) + MSE(f, 0) (or g_MSE(f, 0, col_weights^2))."""
    f_model = RESIDUALS[f_model] if isinstance(f_model, str) else f_model
    w = _t(w_np, dtype, True)
    vars_ = [torch.tensor(float(v), dtype=dtype, requires_grad=True) for v in vars_np]
    cw = _t(col_weights_np, dtype, True) if col_weights_np is not None else None
    u_model = UModel(w, layer_sizes)
    cols = [torch.tensor(np.asarray(X_np)[:, j:j + 1], dtype=dtype).requires_grad_(True)
            for j in range(np.shape(X_np)[1])]
    u_pred = u_model(torch.cat(cols, 1))
    f = f_model(u_model, vars_, *cols)
    u_t = torch.tensor(np.asarray(u_np), dtype=dtype)
    zero = torch.zeros((), dtype=dtype)
    if cw is not None:
        loss = MSE(u_pred, u_t) + g_MSE(f, zero, cw ** 2)
    else:
        loss = MSE(u_pred, u_t) + MSE(f, zero)
    srcs = [w] + vars_ + ([cw] if cw is not None else [])
    gs = torch.autograd.grad(loss, srcs)
    out = {"loss": loss.detach().numpy(), "grad": gs[0].numpy(),
           "var_grads": np.array([g.numpy() for g in gs[1:1 + len(vars_)]], np.float32),
           "f": f.detach().numpy().reshape(-1)}
    if cw is not None:
        out["col_grad"] = gs[-1].numpy()
    return out


def discovery_train(layer_sizes, f_model, X_np, u_np, w_np, vars_np, n_steps, col_weights_np=None):
    """models.py:330-350 train_op: three Adams; quirk Q6 -- the slices at :339/:346
    drop the last network variable, so the output bias never moves."""
    opt, opt_vars, opt_cw = KerasAdam(), KerasAdam(), KerasAdam()
    w = np.array(w_np, np.float32)
    vars_ = np.array(vars_np, np.float32)
    cw = None if col_weights_np is None else np.array(col_weights_np, np.float32)
    n_out_bias = layer_sizes[-1]
    hist = []
    for _ in range(n_steps):
        r = discovery_loss_and_grads(layer_sizes, f_model, X_np, u_np, w, vars_, cw)
        hist.append(np.float32(r["loss"]))
        g = r["grad"].copy()
        opt.begin_step()
        head = opt.apply("w", w[:-n_out_bias].copy(), g[:-n_out_bias])
        w = np.concatenate([head, w[-n_out_bias:]])
        if cw is not None:
            opt_cw.begin_step()
            cw = opt_cw.apply("cw", cw, -r["col_grad"])
        opt_vars.begin_step()
        vars_ = opt_vars.apply("vars", vars_, r["var_grads"])
    return {"w": w, "vars": vars_, "col_weights": cw, "loss_hist": np.array(hist, np.float32)}


# --------------------------------------------------------------------------
# synthetic inputs for benchmarks
# --------------------------------------------------------------------------
def centred_lhs(n, limits, seed=1234):
    """sampling.py:258-260,305-311 (criterion 'c' -> pyDOE2 centred LHS) + _scale_to_xlimits
    :238-254: column j = lo + (perm_j + .5)/n * (hi-lo), perm_j seeded 1234+j."""
    limits = np.asarray(limits, np.float64)
    X = np.empty((n, limits.shape[0]), np.float64)
    for j in range(limits.shape[0]):
        perm = np.random.RandomState(seed + j).permutation(n)
        X[:, j] = limits[j, 0] + (perm + 0.5) / n * (limits[j, 1] - limits[j, 0])
    return X
