"""neural_net (tensordiffeq/networks.py:10-20): Dense(tanh, glorot_normal) x (L-2) + linear
Dense.  The network is a flat fp32 vector in the reference layout; evaluation happens in
the CUDA kernels, this object only holds / initialises / reshapes parameters."""
import math

import numpy as np

_GLOROT_TRUNC = 0.87962566103423978  # keras VarianceScaling truncated-normal correction


class NeuralNet:
    def __init__(self, layer_sizes, seed=None):
        self.layer_sizes = [int(v) for v in layer_sizes]
        rng = np.random.RandomState(seed)
        parts = []
        for i in range(1, len(self.layer_sizes)):
            fi, fo = self.layer_sizes[i - 1], self.layer_sizes[i]
            std = math.sqrt(2.0 / (fi + fo)) / _GLOROT_TRUNC
            W = rng.standard_normal((fi, fo))
            bad = np.abs(W) > 2.0
            while bad.any():
                W[bad] = rng.standard_normal(int(bad.sum()))
                bad = np.abs(W) > 2.0
            parts.append((W * std).astype(np.float32).ravel())
            parts.append(np.zeros(fo, np.float32))
        self._flat = np.concatenate(parts)
        self._engine = None     # set by the solver once parameters live on the device

    # ---- flat view
    def get_flat(self):
        if self._engine is not None:
            self._flat = self._engine.get_params()
        return self._flat.copy()

    def set_flat(self, w):
        w = np.asarray(w, np.float32).ravel()
        if w.size != self._flat.size:
            raise ValueError("expected %d parameters, got %d" % (self._flat.size, w.size))
        self._flat = w.copy()
        if self._engine is not None:
            self._engine.set_params(self._flat)

    # ---- keras-like views
    def get_weights(self):
        out, off, flat = [], 0, self.get_flat()
        for i in range(1, len(self.layer_sizes)):
            fi, fo = self.layer_sizes[i - 1], self.layer_sizes[i]
            out.append(flat[off:off + fi * fo].reshape(fi, fo).copy()); off += fi * fo
            out.append(flat[off:off + fo].copy()); off += fo
        return out

    def set_weights(self, arrays):
        self.set_flat(np.concatenate([np.asarray(a, np.float32).ravel() for a in arrays]))

    @property
    def trainable_variables(self):
        return self.get_weights()

    def count_params(self):
        return int(self._flat.size)

    def summary(self):
        return "tanh MLP %s, %d parameters (fp32, evaluated by libtdq_b200)" % (self.layer_sizes, self._flat.size)

    def save(self, path):
        p = path if str(path).endswith(".npz") else str(path) + ".npz"
        np.savez(p, layer_sizes=np.array(self.layer_sizes, np.int32), flat=self.get_flat())

    @staticmethod
    def load(path):
        p = path if str(path).endswith(".npz") else str(path) + ".npz"
        z = np.load(p)
        net = NeuralNet([int(v) for v in z["layer_sizes"]], seed=0)
        net.set_flat(z["flat"])
        return net


def neural_net(layer_sizes, seed=None):
    return NeuralNet(layer_sizes, seed)
