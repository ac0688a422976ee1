"""Placeholder so tensordiffeq/optimizers.py:5 imports; graph-mode L-BFGS is out of scope."""
class _Opt:
    @staticmethod
    def lbfgs_minimize(*a, **k):
        raise NotImplementedError("tfp is not restated")
optimizer = _Opt()
