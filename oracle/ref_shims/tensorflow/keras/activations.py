from ._core import _Activations
tanh = _Activations.tanh
