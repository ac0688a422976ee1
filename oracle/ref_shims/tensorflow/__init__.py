"""Torch-backed stand-in for the slice of the TensorFlow 2.x API that
tensordiffeq (the reference at /root/reference) touches.

TEST INFRASTRUCTURE ONLY.  TensorFlow is not installable in this image, so
`tests/golden/make_golden.py` puts this directory on sys.path and imports the
UNMODIFIED reference package on top of it.  The reference's own Python
(models.py / fit.py / optimizers.py / boundaries.py) then executes for real;
only the TF primitives are restated here:

  * dense fp32 math, tanh, reductions          -> torch CPU ops
  * tf.gradients / GradientTape.gradient       -> torch.autograd.grad
  * keras Dense / Sequential / glorot_normal   -> see .keras
  * keras.optimizers.Adam (ResourceApplyAdam)  -> see .keras.optimizers

Nothing in the product (tensordiffeq_b200/) imports this.
"""
import numpy as _np
import torch as _torch

_torch.set_default_dtype(_torch.float32)


class DType:
    def __init__(self, name, tdt):
        self.name = name
        self.torch = tdt

    def __repr__(self):
        return "tf." + self.name


float32 = DType("float32", _torch.float32)
float64 = DType("float64", _torch.float64)
int32 = DType("int32", _torch.int32)
int64 = DType("int64", _torch.int64)


def _tdt(dtype):
    if dtype is None:
        return None
    if isinstance(dtype, DType):
        return dtype.torch
    if isinstance(dtype, _torch.dtype):
        return dtype
    return {"float32": _torch.float32, "float64": _torch.float64}[str(dtype)]


def _raw(x, like=None):
    """Anything -> torch.Tensor (lists of tensors are stacked, as
    ops.convert_to_tensor does in TF; this is what produces the [1,N,1]
    residual shape of SURVEY Q1)."""
    if isinstance(x, Tensor):
        return x.t
    if isinstance(x, _torch.Tensor):
        return x
    if isinstance(x, (list, tuple)):
        if len(x) and any(isinstance(e, (Tensor, _torch.Tensor)) for e in x):
            return _torch.stack([_raw(e) for e in x], 0)
        arr = _np.asarray(x)
        if arr.dtype == _np.float64 and like is not None:
            return _torch.as_tensor(arr, dtype=like.dtype)
        return _torch.as_tensor(arr)
    if isinstance(x, _np.ndarray):
        return _torch.as_tensor(x)
    if isinstance(x, (_np.floating, float, int, _np.integer, bool)):
        if like is not None:
            return _torch.tensor(x, dtype=like.dtype)
        if isinstance(x, (float, _np.floating)):
            return _torch.tensor(float(x), dtype=_torch.float32)
        return _torch.tensor(x)
    raise TypeError("tf-shim: cannot convert %r" % type(x))


class Tensor:
    __array_priority__ = 10000
    __array_ufunc__ = None

    def __init__(self, t):
        self.t = t

    # ---- introspection
    @property
    def shape(self):
        return tuple(self.t.shape)

    @property
    def dtype(self):
        return self.t.dtype

    def numpy(self):
        return self.t.detach().cpu().numpy().copy()  # a value, never an alias of variable storage

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a.astype(dtype) if dtype is not None else a

    def __len__(self):
        return self.t.shape[0]

    def __iter__(self):
        for i in range(self.t.shape[0]):
            yield Tensor(self.t[i])

    def __getitem__(self, idx):
        if isinstance(idx, Tensor):
            idx = idx.t
        return Tensor(self.t[idx])

    def __repr__(self):
        return "tf.Tensor(%r)" % (self.t.detach(),)

    def __bool__(self):
        return bool(self.t.item())

    def __float__(self):
        return float(self.t.item())

    def __int__(self):
        return int(self.t.item())

    def __format__(self, spec):
        return format(self.numpy(), spec) if self.t.ndim else format(float(self), spec)

    # ---- arithmetic
    def _bin(self, other, fn, rev=False):
        o = _raw(other, like=self.t)
        a, b = (o, self.t) if rev else (self.t, o)
        if a.is_floating_point() and b.is_floating_point() and a.dtype != b.dtype:
            # python / numpy scalars adopt the tensor dtype; tensors must agree
            if o.ndim == 0 or not isinstance(other, (Tensor, _torch.Tensor)):
                o = o.to(self.t.dtype)
                a, b = (o, self.t) if rev else (self.t, o)
        return Tensor(fn(a, b))

    def __add__(self, o): return self._bin(o, _torch.add)
    def __radd__(self, o): return self._bin(o, _torch.add, True)
    def __iadd__(self, o): return self._bin(o, _torch.add)
    def __sub__(self, o): return self._bin(o, _torch.sub)
    def __rsub__(self, o): return self._bin(o, _torch.sub, True)
    def __mul__(self, o): return self._bin(o, _torch.mul)
    def __rmul__(self, o): return self._bin(o, _torch.mul, True)
    def __truediv__(self, o): return self._bin(o, _torch.div)
    def __rtruediv__(self, o): return self._bin(o, _torch.div, True)
    def __pow__(self, o): return self._bin(o, _torch.pow)
    def __neg__(self): return Tensor(-self.t)
    def __abs__(self): return Tensor(self.t.abs())

    def __lt__(self, o): return bool((self.t < _raw(o, like=self.t)).item())
    def __le__(self, o): return bool((self.t <= _raw(o, like=self.t)).item())
    def __gt__(self, o): return bool((self.t > _raw(o, like=self.t)).item())
    def __ge__(self, o): return bool((self.t >= _raw(o, like=self.t)).item())


class Variable(Tensor):
    def __init__(self, initial_value, trainable=True, dtype=None, validate_shape=True, name=None):
        t = _raw(initial_value).detach().clone()
        if dtype is not None:
            t = t.to(_tdt(dtype))
        elif t.dtype == _torch.float64:
            t = t.to(_torch.float32)
        t.requires_grad_(bool(trainable) and t.is_floating_point())
        super().__init__(t)
        self.trainable = trainable

    def assign(self, value):
        with _torch.no_grad():
            self.t.copy_(_raw(value).to(self.t.dtype).reshape(self.t.shape))
        return self

    def assign_sub(self, value):
        with _torch.no_grad():
            self.t.sub_(_raw(value).to(self.t.dtype))
        return self


def _leaf(t):
    """Constants built from host data behave like tf.function arguments: they
    can be differentiated against (tf.gradients(u, x)), so make them leaves
    that record history."""
    if t.is_floating_point() and not t.requires_grad:
        t = t.detach().clone().requires_grad_(True)
    return t


def cast(x, dtype=None):
    t = _raw(x)
    td = _tdt(dtype)
    if isinstance(x, (Tensor, _torch.Tensor)):
        return Tensor(t.to(td))
    return Tensor(_leaf(t.to(td)))


def convert_to_tensor(x, dtype=None):
    t = _raw(x)
    if dtype is not None:
        t = t.to(_tdt(dtype))
    return Tensor(t)


def constant(val, dtype=None):
    t = _raw(val)
    if dtype is not None:
        t = t.to(_tdt(dtype))
    elif t.dtype == _torch.float64:
        t = t.to(_torch.float32)
    return Tensor(t)


def shape(x):
    return Tensor(_torch.tensor(list(_np.shape(x.numpy() if isinstance(x, Tensor) else x)), dtype=_torch.int32))


def slice(x, begin, size):  # noqa: A001
    t = _raw(x)
    idx = tuple(builtins_slice(b, b + s) for b, s in zip(begin, size))
    return Tensor(t[idx])


import builtins as _builtins
builtins_slice = _builtins.slice


def concat(values, axis):
    ts = [_raw(v) for v in values]
    return Tensor(_torch.cat(ts, dim=axis))


def reshape(x, shp):
    return Tensor(_raw(x).reshape(tuple(int(s) for s in shp)))


def gather(x, idx, axis=0):
    i = _torch.as_tensor(_np.asarray(idx if not isinstance(idx, Tensor) else idx.numpy()), dtype=_torch.long)
    return Tensor(_raw(x).index_select(axis, i))


def reduce_mean(x, axis=None):
    t = _raw(x)
    return Tensor(t.mean() if axis is None else t.mean(dim=axis))


def reduce_sum(x, axis=None):
    t = _raw(x)
    return Tensor(t.sum() if axis is None else t.sum(dim=axis))


def square(x):
    t = _raw(x)
    return Tensor(t * t)


def add(a, b):
    return convert_to_tensor(a) + b


def abs(x, name=None):  # noqa: A001  (reference passes a 2nd positional: SURVEY Q8)
    return Tensor(_raw(x).abs())


def zeros(shp, dtype=float32):
    return Tensor(_torch.zeros(tuple(_np.atleast_1d(shp).tolist()), dtype=_tdt(dtype)))


def timestamp():
    import time
    return Tensor(_torch.tensor(time.time(), dtype=_torch.float64))


def function(f=None, **kw):
    """tf.function: numpy arguments become graph inputs (tensors one can take
    tf.gradients against, dtype preserved)."""
    if f is None:
        return lambda g: function(g)

    def wrapped(*args, **kwargs):
        conv = []
        for a in args:
            if isinstance(a, _np.ndarray) and a.dtype.kind == "f":
                conv.append(Tensor(_leaf(_torch.as_tensor(a))))
            else:
                conv.append(a)
        return f(*conv, **kwargs)

    wrapped.__wrapped__ = f
    return wrapped


def gradients(ys, xs, grad_ys=None):
    """tf.gradients: d(sum ys)/d xs, differentiable again (TF builds the
    backward graph symbolically; create_graph=True is the torch equivalent)."""
    ys_l = ys if isinstance(ys, (list, tuple)) else [ys]
    xs_l = xs if isinstance(xs, (list, tuple)) else [xs]
    total = None
    for y in ys_l:
        s = _raw(y).sum()
        total = s if total is None else total + s
    gs = _torch.autograd.grad(total, [_raw(x) for x in xs_l], create_graph=True, allow_unused=True)
    return [None if g is None else Tensor(g) for g in gs]


class GradientTape:
    def __init__(self, persistent=False, watch_accessed_variables=True):
        self.persistent = persistent

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def watch(self, x):
        t = _raw(x)
        if not t.requires_grad:
            t.requires_grad_(True)

    def gradient(self, target, sources):
        single = not isinstance(sources, (list, tuple))
        src = [sources] if single else list(sources)
        tgt = target if isinstance(target, (list, tuple)) else [target]
        total = None
        for y in tgt:
            s = _raw(y).sum()
            total = s if total is None else total + s
        gs = _torch.autograd.grad(total, [_raw(s) for s in src], create_graph=True,
                                  retain_graph=True, allow_unused=True)
        # the tape result is a value, not a graph the caller differentiates
        # again, unless the tape is persistent (test/Burgers2test.py:58-66)
        out = [None if g is None else Tensor(g if self.persistent else g.detach()) for g in gs]
        return out[0] if single else out


class _Nn:
    @staticmethod
    def tanh(x):
        return Tensor(_torch.tanh(_raw(x)))


nn = _Nn()


class _Random:
    @staticmethod
    def uniform(shp, minval=0.0, maxval=1.0, dtype=float32, seed=None):
        return Tensor(_torch.rand(tuple(int(s) for s in shp), dtype=_tdt(dtype)) * (maxval - minval) + minval)

    @staticmethod
    def set_seed(s):
        _torch.manual_seed(int(s))


random = _Random()


class _Config:
    @staticmethod
    def run_functions_eagerly(flag):
        return None


config = _Config()


class _Distribute:
    class ReduceOp:
        SUM = "SUM"

    class MirroredStrategy:
        def __init__(self, *a, **k):
            raise RuntimeError("tf-shim: tf.distribute is not restated (reference dist mode replicates compute, SURVEY 0)")


distribute = _Distribute()

from . import math  # noqa: E402
from . import keras  # noqa: E402


def dynamic_stitch(*a, **k):
    raise NotImplementedError("tf-shim: graph-mode L-BFGS (tfp) is out of scope")
