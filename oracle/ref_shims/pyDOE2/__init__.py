"""pyDOE2 1.3.0 `lhs` restated for the one criterion the reference uses by default
(tensordiffeq/sampling.py:258-260,305-311: criterion 'c' -> pyDOE2 _lhscentered):
cell centres (i+0.5)/n, independently permuted per column by random_state.
Test infrastructure only."""
import numpy as np


def lhs(n, samples=None, criterion=None, iterations=None, random_state=None, correlation_matrix=None):
    if random_state is None:
        random_state = np.random.RandomState()
    elif not isinstance(random_state, np.random.RandomState):
        random_state = np.random.RandomState(random_state)
    if samples is None:
        samples = n
    if criterion not in ("c", "center"):
        raise NotImplementedError("only centred LHS is restated")
    cut = np.linspace(0, 1, samples + 1)
    a, b = cut[:samples], cut[1:samples + 1]
    centre = (a + b) / 2
    H = np.zeros((samples, n))
    for j in range(n):
        H[:, j] = random_state.permutation(centre)
    return H
