"""keras.optimizers.Adam as TF 2.6 executes it: optimizer_v2/adam.py
_resource_apply_dense -> ResourceApplyAdam (training_ops.cc ApplyAdam, non-nesterov):
  alpha = lr * sqrt(1 - beta2^t) / (1 - beta1^t)
  m += (g - m) * (1 - beta1);  v += (g*g - v) * (1 - beta2)
  var -= (m * alpha) / (sqrt(v) + epsilon)           epsilon = 1e-7 (keras default)
(test infrastructure only)."""
import numpy as _np
import torch as _torch
from .. import _raw


class Adam:
    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, lr=None, **k):
        self.lr = float(lr if lr is not None else learning_rate)
        self.beta_1 = float(beta_1)
        self.beta_2 = float(beta_2)
        self.epsilon = float(epsilon)
        self.iterations = 0
        self._slots = {}

    def apply_gradients(self, grads_and_vars):
        gv = [(g, v) for g, v in grads_and_vars]
        if not gv or all(g is None for g, _ in gv):
            raise ValueError("No gradients provided for any variable")
        self.iterations += 1
        t = self.iterations
        f32 = _np.float32
        b1, b2 = f32(self.beta_1), f32(self.beta_2)
        b1p = f32(_np.power(b1, f32(t)))
        b2p = f32(_np.power(b2, f32(t)))
        alpha = f32(f32(self.lr) * _np.sqrt(f32(1) - b2p) / (f32(1) - b1p))
        with _torch.no_grad():
            for g, v in gv:
                if g is None:
                    continue
                gt = _raw(g).detach().to(v.t.dtype).reshape(v.t.shape)
                key = id(v)
                if key not in self._slots:
                    self._slots[key] = (_torch.zeros_like(v.t), _torch.zeros_like(v.t))
                m, s = self._slots[key]
                m.add_((gt - m) * float(f32(1) - b1))
                s.add_((gt * gt - s) * float(f32(1) - b2))
                v.t.sub_((m * float(alpha)) / (s.sqrt() + float(f32(self.epsilon))))
