"""Training drivers with the reference's names (tensordiffeq/fit.py:17-190).  The per-step
arithmetic (loss, gradients, Adam on weights, Adam ascent on lambdas, L-BFGS) is enqueued on
the GPU by libtdq_b200; the host only chunks steps so the progress bar can show a loss every
10 epochs like the reference (fit.py:36-37)."""
import numpy as np
from tqdm.auto import trange

from . import _lib
from .output import print_screen

_CHUNK = 10


def _adam_cfgs(obj):
    E = _lib.Engine
    w = obj.tf_optimizer.cfg(E, 1)
    has_lam = bool(obj.isAdaptive and obj.lambdas)
    lam = obj.tf_optimizer_weights.cfg(E, -1) if has_lam else None    # negated gradients: fit.py:101,104
    return w, lam


def _adam_loop(obj, tf_iter, desc="Adam epoch"):
    eng = obj._engine()
    w_cfg, lam_cfg = _adam_cfgs(obj)
    hist_all = []
    if obj.n_batches > 1:
        # fit.py:86-89 + models.py:148-155: a permutation drawn once (quirk Q7: frozen at trace time),
        # batches are consecutive slices of it, the remainder N_f % batch_sz is never visited
        N_f = int(obj.X_f_len[0])
        if getattr(obj, "_perm", None) is None or len(obj._perm) != N_f:
            obj.batch_indx_map = np.random.choice(N_f, size=N_f, replace=False)
            obj._set_minibatch_order(obj.batch_indx_map)
        res = [t for t in obj._terms if t["kind"] == "residual"]
    with trange(tf_iter) as t:
        done = 0
        while done < tf_iter:
            n = min(_CHUNK, tf_iter - done)
            if obj.n_batches > 1:
                # n epochs x n_batches Adam steps in one library call (no host round trip per batch); each epoch
                # reports its last batch's loss (fit.py:109)
                hist = eng.adam_epochs(n, [dsc["tid"] for dsc in res], obj.batch_sz, obj.n_batches, w=w_cfg, lam=lam_cfg)
            else:
                hist = eng.adam_steps(n, w=w_cfg, lam=lam_cfg)
            hist_all.extend(float(h) for h in hist)
            done += n
            t.update(n)
            t.set_description('%s %i' % (desc, done))
            t.set_postfix(loss=float(hist[0]))
    if obj.n_batches > 1:
        for dsc in res:
            eng.set_range(dsc["tid"], 0, -1, 1)
    obj._pull_lambdas()
    obj.loss_history = getattr(obj, "loss_history", []) + hist_all
    return hist_all


def fit(obj, tf_iter=0, newton_iter=0, newton_eager=True):
    """fit.py:17-61"""
    if obj.verbose:
        print_screen(obj)
    print("Starting Adam training")
    _adam_loop(obj, tf_iter)
    if newton_iter > 0:
        obj.n_batches = 1
        print("Starting L-BFGS training")
        eng = obj._engine()
        if newton_eager:
            print("Executing eager-mode L-BFGS")
            f_hist, n_eval = eng.lbfgs(int(newton_iter), 0.8)      # fit.py:49-51
            obj.lbfgs_history = [float(f) for f in f_hist]
            obj.lbfgs_n_eval = n_eval
        else:
            # fit.py:52-58,65-80: tfp.optimizer.lbfgs_minimize(..., max_iterations=newton_iter, tolerance=1e-20)
            print("Executing graph-mode L-BFGS\n Building graph...")
            f_hist, n_eval, status = eng.lbfgs_hz(int(newton_iter), 1e-20)
            obj.lbfgs_history = [float(f) for f in f_hist]
            obj.lbfgs_n_eval = n_eval
            obj.lbfgs_status = {0: "max_iterations", 1: "converged", 2: "line search failed"}[status]


def fit_dist(obj, tf_iter, newton_iter, batch_sz=None, newton_eager=True):
    """fit.py:116-190.  The reference replicates the full batch on every replica and SUM-reduces
    identical gradients (fit.py:128-153; models.py:148-157 never reads dist_X_f).  Here the
    collocation points and their lambdas are sharded across ranks, each rank evaluates its slice
    scaled by 1/N_f, and one packed NCCL all-reduce yields the single-device gradient."""
    eng = obj._engine()
    if obj.verbose and obj._dist.rank == 0:
        print_screen(obj)
        print("Number of GPU devices: {}".format(obj._dist.world))
        print("starting Adam training")
    _adam_loop(obj, tf_iter)
    if obj._dist.rank == 0:
        print("Starting L-BFGS training")   # fit.py:187-190: printed, never run in dist mode


def train_op_inner(obj):
    """fit.py:83-113: returns a callable doing one epoch of Adam and returning the loss"""
    def apply_grads(obj=obj):
        return np.float32(_adam_loop_quiet(obj, 1)[-1])
    return apply_grads


def _adam_loop_quiet(obj, n):
    eng = obj._engine()
    w_cfg, lam_cfg = _adam_cfgs(obj)
    hist = eng.adam_steps(n, w=w_cfg, lam=lam_cfg)
    obj._pull_lambdas()
    return [float(h) for h in hist]
