"""Latin-hypercube sampling of collocation points on the host (one-off set-up work).
Reference: tensordiffeq/sampling.py:256-311 -- class LHS, default criterion "c" (centred),
which delegates to pyDOE2.lhs; scaling per :238-254."""
import numpy as np


class LHS:
    def __init__(self, xlimits, criterion="c", random_state=None):
        self.xlimits = np.asarray(xlimits, dtype=np.float64)
        if criterion not in ("c", "center"):
            raise NotImplementedError("only the centred criterion (the reference default) is available")
        if isinstance(random_state, np.random.RandomState):
            self.random_state = random_state
        elif random_state is None:
            self.random_state = np.random.RandomState()
        else:
            self.random_state = np.random.RandomState(int(random_state))

    def __call__(self, nt):
        nx = self.xlimits.shape[0]
        centres = (np.arange(nt, dtype=np.float64) + 0.5) / nt
        out = np.empty((nt, nx), dtype=np.float64)
        for j in range(nx):
            lo, hi = self.xlimits[j]
            out[:, j] = lo + self.random_state.permutation(centres) * (hi - lo)
        return out
