"""Optimizer value objects.  The arithmetic runs on the device (csrc/optim_kernels.cuh):
Adam with TF ResourceApplyAdam semantics, and the eager L-BFGS of
tensordiffeq/optimizers.py:214-402.  These classes only carry hyper-parameters, so user
code can keep assigning `model.tf_optimizer = tf.keras.optimizers.Adam(lr=.005)`
(examples/steady-state-poisson.py:59)."""


class Adam:
    """keras.optimizers.Adam hyper-parameters; the reference's defaults are
    Adam(lr=0.005, beta_1=.99) (models.py:35-36), keras' own are lr 1e-3, beta_1 .9."""

    def __init__(self, lr=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, learning_rate=None):
        self.lr = float(learning_rate if learning_rate is not None else lr)
        self.beta_1 = float(beta_1)
        self.beta_2 = float(beta_2)
        self.epsilon = float(epsilon)

    def cfg(self, engine_cls, sign=1):
        return engine_cls.adam_cfg(self.lr, self.beta_1, self.beta_2, self.epsilon, sign)


class Struct(object):
    """optimizers.py:405-414: Lua-like struct with 0 defaults (kept for API compatibility)."""

    def __getattribute__(self, key):
        if key == "__dict__":
            return object.__getattribute__(self, "__dict__")
        return self.__dict__.get(key, 0)


def eager_lbfgs(model, maxIter=100, learningRate=0.8):
    """optimizers.py:214 entry point re-expressed on a compiled solver: runs the device-side
    iteration and returns (weights, f_hist, n_eval) like the reference."""
    eng = model._engine()
    hist, n_eval = eng.lbfgs(int(maxIter), float(learningRate))
    return eng.get_params(), list(hist), n_eval
