"""Host utilities with the names of tensordiffeq/utils.py."""
import numpy as np

from . import tf as _tf
from .sampling import LHS


def get_sizes(layer_sizes):
    """utils.py:32-35"""
    sizes_w = [layer_sizes[i] * layer_sizes[i - 1] for i in range(1, len(layer_sizes))]
    sizes_b = list(layer_sizes[1:])
    return sizes_w, sizes_b


def set_weights(model, w, sizes_w=None, sizes_b=None):
    """utils.py:7-16: flat vector (per layer: kernel [in,out] row-major, bias) -> network"""
    model.set_flat(np.asarray(w.numpy() if hasattr(w, "numpy") else w, np.float32))


def get_weights(model):
    """utils.py:19-29"""
    return _tf.HostTensor(model.get_flat())


def MSE(pred, actual, weights=None):
    """utils.py:38-41 on host arrays"""
    d = np.asarray(pred, np.float32) - np.asarray(actual, np.float32)
    if weights is not None:
        d = np.asarray(weights, np.float32) * d
    return np.mean(np.square(d))


def g_MSE(pred, actual, g_lam):
    """utils.py:44-45 on host arrays"""
    d = np.asarray(pred, np.float32) - np.asarray(actual, np.float32)
    return np.mean(np.asarray(g_lam, np.float32) * np.square(d))


def constant(val, dtype=None):
    return _tf.constant(val, dtype)


def convertTensor(val, dtype=None):
    return _tf.cast(val, dtype)


def tensor(x, dtype=None):
    return _tf.convert_to_tensor(x, dtype)


def LatinHypercubeSample(N_f, bounds):
    """utils.py:56-58"""
    return LHS(xlimits=bounds)(N_f)


def get_tf_model(model):
    return model


def multimesh(arrs):
    """utils.py:65-84: like np.meshgrid(*arrs, indexing='ij')"""
    return list(np.meshgrid(*[np.asarray(a) for a in arrs], indexing="ij"))


def flatten_and_stack(mesh):
    """utils.py:88-93: [n_points, n_arrays]"""
    return np.stack([np.asarray(m).reshape(-1) for m in mesh], axis=1).astype(np.float64)


def initialize_weights_loss(init_weights):
    """utils.py:99-112: lambda list (residual entries first, then BCs, None skipped) and index map
    with lower-cased keys."""
    lambdas, lambdas_map, counter = [], {}, 0
    for key, values in init_weights.items():
        idx = []
        for value in values:
            if value is not None:
                lambdas.append(_tf.Variable(value, trainable=True, dtype=_tf.float32))
                idx.append(counter)
                counter += 1
        lambdas_map[key.lower()] = idx
    return lambdas, lambdas_map
