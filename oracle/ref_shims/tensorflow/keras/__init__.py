"""Keras subset (see ../__init__.py: test infrastructure only)."""
from ._core import Sequential, layers as _layers_ns, activations as _act_ns
from . import layers, models, activations, optimizers
