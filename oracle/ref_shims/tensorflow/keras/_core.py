"""Keras subset used by tensordiffeq/networks.py:10-20 and models.py:35-36
(see ../__init__.py: test infrastructure only)."""
import numpy as _np
import torch as _torch
from .. import Tensor, Variable, _raw, float32

_init_rng = _np.random.RandomState(0)


def _glorot_normal(fan_in, fan_out):
    # keras GlorotNormal = VarianceScaling(1.0, "fan_avg", "truncated_normal"):
    # stddev = sqrt(2/(fan_in+fan_out)) / .87962566103423978, resampled outside 2 sigma
    std = _np.sqrt(2.0 / (fan_in + fan_out)) / 0.87962566103423978
    out = _init_rng.standard_normal((fan_in, fan_out))
    bad = _np.abs(out) > 2.0
    while bad.any():
        out[bad] = _init_rng.standard_normal(int(bad.sum()))
        bad = _np.abs(out) > 2.0
    return (out * std).astype(_np.float32)


class _InputLayer:
    def __init__(self, input_shape=None, **k):
        self.input_shape = input_shape


class _Dense:
    def __init__(self, units, activation=None, kernel_initializer="glorot_normal", **k):
        self.units = int(units)
        self.activation = activation
        self.kernel = None
        self.bias = None

    def build(self, fan_in):
        self.kernel = Variable(_glorot_normal(fan_in, self.units), dtype=float32)
        self.bias = Variable(_np.zeros(self.units, _np.float32), dtype=float32)

    def get_weights(self):
        return [self.kernel.numpy().copy(), self.bias.numpy().copy()]

    def set_weights(self, wb):
        k, b = wb
        self.kernel.assign(_np.asarray(k.numpy() if isinstance(k, Tensor) else k, _np.float32))
        self.bias.assign(_np.asarray(b.numpy() if isinstance(b, Tensor) else b, _np.float32))

    def __call__(self, t):
        z = t @ self.kernel.t + self.bias.t
        if self.activation is not None:
            return _raw(self.activation(Tensor(z)))
        return z


class _Layers:
    InputLayer = _InputLayer
    Dense = _Dense
    Input = _InputLayer


layers = _Layers()


class _Activations:
    @staticmethod
    def tanh(x):
        return Tensor(_torch.tanh(_raw(x)))


activations = _Activations()


class Sequential:
    def __init__(self):
        self.layers = []
        self._width = None

    def add(self, layer):
        if isinstance(layer, _InputLayer):
            self._width = int(layer.input_shape[0])
            return
        layer.build(self._width)
        self._width = layer.units
        self.layers.append(layer)

    @property
    def trainable_variables(self):
        out = []
        for l in self.layers:
            out.extend([l.kernel, l.bias])
        return out

    def __call__(self, x):
        # keras autocast: inputs are cast to the layer dtype, float32
        t = _raw(x)
        t = t.to(_torch.float32)
        for l in self.layers:
            t = l(t)
        return Tensor(t)

    def summary(self):
        n = sum(int(_np.prod(v.shape)) for v in self.trainable_variables)
        return "Sequential(%d dense layers, %d params)" % (len(self.layers), n)

    def save(self, path):
        _np.savez(path, *[v.numpy() for v in self.trainable_variables])


class _Models:
    Sequential = Sequential

    @staticmethod
    def load_model(path, compile=False):
        raise NotImplementedError


models = _Models()

