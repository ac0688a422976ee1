"""The handful of `tf.*` names the reference's user scripts touch inside f_model /
deriv_model / init-weight code (`from tensordiffeq.boundaries import *` re-exports `tf`
there: boundaries.py:3), mapped onto the tracer.  Not TensorFlow, and not a compute
path: every function here either builds a symbolic expression (during tracing) or does
small host-side numpy work."""
import math as _math

import numpy as np

from . import trace as _T

float32 = np.float32
float64 = np.float64


class Variable:
    """tf.Variable stand-in: a named host value (PDE coefficients, lambda vectors)."""

    def __init__(self, initial_value, trainable=True, dtype=None, validate_shape=True, name=None):
        v = initial_value.numpy() if hasattr(initial_value, "numpy") else initial_value
        self._v = np.array(v, dtype=np.float32)
        self.trainable = trainable
        self._sym = None     # set while a DiscoveryModel traces its residual

    def numpy(self):
        return self._v.copy() if self._v.ndim else np.float32(self._v)

    def assign(self, v):
        self._v = np.array(v.numpy() if hasattr(v, "numpy") else v, dtype=np.float32).reshape(self._v.shape)
        return self

    def sym_value(self):
        if self._sym is not None:
            return self._sym
        if self._v.size == 1:
            return _T.const(float(self._v.reshape(-1)[0]))
        raise TypeError("a vector-valued Variable cannot appear in a traced residual")

    @property
    def shape(self):
        return self._v.shape

    # arithmetic: symbolic while tracing, numpy otherwise
    def _b(self, o, f, r=False):
        if self._sym is not None or isinstance(o, _T.Sym):
            a = _T.as_sym(self)
            return f(o, a) if r else f(a, o)
        ov = o.numpy() if hasattr(o, "numpy") else o
        npf = {_T.add: np.add, _T.sub: np.subtract, _T.mul: np.multiply, _T.div: np.divide}[f]
        return HostTensor(npf(ov, self._v) if r else npf(self._v, ov))

    def __add__(self, o): return self._b(o, _T.add)
    def __radd__(self, o): return self._b(o, _T.add, True)
    def __sub__(self, o): return self._b(o, _T.sub)
    def __rsub__(self, o): return self._b(o, _T.sub, True)
    def __mul__(self, o): return self._b(o, _T.mul)
    def __rmul__(self, o): return self._b(o, _T.mul, True)
    def __truediv__(self, o): return self._b(o, _T.div)
    def __rtruediv__(self, o): return self._b(o, _T.div, True)
    def __neg__(self): return _T.neg(_T.as_sym(self)) if self._sym is not None else HostTensor(-self._v)
    def __pow__(self, n): return _T.power(_T.as_sym(self), n) if self._sym is not None else HostTensor(self._v ** n)


class HostTensor(np.ndarray):
    """numpy array with a .numpy() method (what tf.random.uniform / tf.constant return on the host)."""

    def __new__(cls, a):
        return np.asarray(a, dtype=np.float32 if np.asarray(a).dtype.kind == "f" else None).view(cls)

    def numpy(self):
        return np.asarray(self).copy()

    def sym_value(self):
        if self.size == 1:
            return _T.const(float(np.asarray(self).reshape(-1)[0]))
        raise TypeError("a host array cannot appear in a traced residual")

    # constants meet symbolic values (or tf.gradients lists) inside traced callbacks
    def __add__(self, o): return _T.add(self, o) if _is_sym(o) else np.ndarray.__add__(self, o)
    def __radd__(self, o): return _T.add(o, self) if _is_sym(o) else np.ndarray.__radd__(self, o)
    def __sub__(self, o): return _T.sub(self, o) if _is_sym(o) else np.ndarray.__sub__(self, o)
    def __rsub__(self, o): return _T.sub(o, self) if _is_sym(o) else np.ndarray.__rsub__(self, o)
    def __mul__(self, o): return _T.mul(self, o) if _is_sym(o) else np.ndarray.__mul__(self, o)
    def __rmul__(self, o): return _T.mul(o, self) if _is_sym(o) else np.ndarray.__rmul__(self, o)
    def __truediv__(self, o): return _T.div(self, o) if _is_sym(o) else np.ndarray.__truediv__(self, o)
    def __rtruediv__(self, o): return _T.div(o, self) if _is_sym(o) else np.ndarray.__rtruediv__(self, o)


def _is_sym(x):
    return isinstance(x, _T.Sym) or (isinstance(x, (list, tuple)) and len(x) == 1 and _is_sym(x[0])) or \
        (isinstance(x, Variable) and x._sym is not None)


def constant(val, dtype=None):
    a = np.asarray(val, dtype=np.float32 if dtype in (None, float32, "float32") else dtype)
    return HostTensor(a)


def convert_to_tensor(x, dtype=None):
    if _is_sym(x):
        return _T.as_sym(x)
    return constant(x.numpy() if hasattr(x, "numpy") else x, dtype)


def cast(x, dtype=None):
    if _is_sym(x):
        return _T.as_sym(x)
    return HostTensor(np.asarray(x.numpy() if hasattr(x, "numpy") else x).astype(np.float32 if dtype in (None, float32) else dtype))


class Coords:
    """Result of tf.concat([x, t], 1) on the symbolic input columns."""

    def __init__(self, cols):
        self.cols = cols


def concat(values, axis=1):
    vals = list(values)
    if vals and all(isinstance(v, _T.Sym) and v.kind == "coord" for v in vals):
        return Coords([v.val for v in vals])
    if any(_is_sym(v) for v in vals):
        raise TypeError("tf.concat inside a traced residual must join the raw input columns, in domain order")
    return HostTensor(np.concatenate([np.asarray(v.numpy() if hasattr(v, "numpy") else v) for v in vals], axis=axis))


def gradients(ys, xs, grad_ys=None):
    """tf.gradients(ys, xs): list with d(sum ys)/dx for every x (here: symbolic)."""
    ys_l = ys if isinstance(ys, (list, tuple)) else [ys]
    xs_l = xs if isinstance(xs, (list, tuple)) else [xs]
    out = []
    for x in xs_l:
        x = _T.as_sym(x)
        if x.kind != "coord":
            raise TypeError("tf.gradients inside a traced residual differentiates against an input column")
        tot = None
        for y in ys_l:
            dy = _T.total_derivative(_T.as_sym(y), x.val)
            tot = dy if tot is None else _T.add(tot, dy)
        out.append(tot)
    return out


class GradientTape:
    """tf.GradientTape as test/Burgers2test.py:58-66 uses it."""

    def __init__(self, persistent=False, watch_accessed_variables=True):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def watch(self, x):
        return None

    def gradient(self, target, sources):
        if isinstance(sources, (list, tuple)):
            return gradients(target, list(sources))
        return gradients(target, sources)[0]


def _un(kind, npf):
    def f(x, name=None):
        if _is_sym(x):
            return _T.unary(kind, x)
        return HostTensor(npf(np.asarray(x.numpy() if hasattr(x, "numpy") else x, dtype=np.float32)))
    return f


sin = _un("sin", np.sin)
cos = _un("cos", np.cos)
exp = _un("exp", np.exp)
tanh = _un("tanh", np.tanh)
sqrt = _un("sqrt", np.sqrt)
log = _un("log", np.log)


def square(x):
    return _T.mul(x, x) if _is_sym(x) else HostTensor(np.square(np.asarray(x, np.float32)))


def pow(x, n):  # noqa: A001
    return _T.power(_T.as_sym(x), n) if _is_sym(x) else HostTensor(np.power(np.asarray(x, np.float32), n))


def _bin(tf_name, sym_f, npf):
    def f(a, b, name=None):
        if _is_sym(a) or _is_sym(b):
            return sym_f(a, b)
        return HostTensor(npf(np.asarray(a, np.float32), np.asarray(b, np.float32)))
    f.__name__ = tf_name
    return f


add = _bin("add", _T.add, np.add)
subtract = _bin("subtract", _T.sub, np.subtract)
multiply = _bin("multiply", _T.mul, np.multiply)
divide = _bin("divide", _T.div, np.divide)


class _Math:
    sin = staticmethod(sin); cos = staticmethod(cos); exp = staticmethod(exp); tanh = staticmethod(tanh)
    sqrt = staticmethod(sqrt); log = staticmethod(log); square = staticmethod(square); pow = staticmethod(pow)
    add = staticmethod(add); subtract = staticmethod(subtract); multiply = staticmethod(multiply)
    divide = staticmethod(divide)


math = _Math()


class _Nn:
    tanh = staticmethod(tanh)


nn = _Nn()

_rng = np.random.RandomState()


class _Random:
    @staticmethod
    def uniform(shape, minval=0.0, maxval=1.0, dtype=None, seed=None):
        return HostTensor((_rng.uniform(minval, maxval, size=tuple(int(s) for s in shape))).astype(np.float32))

    @staticmethod
    def set_seed(s):
        global _rng
        _rng = np.random.RandomState(int(s))
        _random_ns_seed(int(s))


def _random_ns_seed(s):
    _Random._seed = s


random = _Random()


class _Config:
    @staticmethod
    def run_functions_eagerly(flag):
        return None


config = _Config()


def function(f=None, **kw):
    return f if f is not None else (lambda g: g)


class _Optimizers:
    @staticmethod
    def Adam(learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, lr=None, **kw):
        from .optimizers import Adam as _Adam
        return _Adam(lr=lr if lr is not None else learning_rate, beta_1=beta_1, beta_2=beta_2, epsilon=epsilon)


class _Keras:
    optimizers = _Optimizers()


keras = _Keras()
pi = _math.pi
