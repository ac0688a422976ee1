"""tf.math subset (see ../__init__.py: test infrastructure only)."""
import torch as _torch
from .. import Tensor, _raw, convert_to_tensor


def subtract(a, b):
    return convert_to_tensor(a) - b


def add(a, b):
    return convert_to_tensor(a) + b


def multiply(a, b):
    return convert_to_tensor(a) * b


def sin(x):
    t = _raw(x)
    if t.dtype == _torch.float64:
        pass
    return Tensor(_torch.sin(t))


def cos(x):
    return Tensor(_torch.cos(_raw(x)))


def exp(x):
    return Tensor(_torch.exp(_raw(x)))


def tanh(x):
    return Tensor(_torch.tanh(_raw(x)))


def square(x):
    t = _raw(x)
    return Tensor(t * t)
