"""Synthetic inputs of the shapes BASELINE.json names: centred Latin-hypercube
collocation sets and Glorot-initialised networks, reproducible from a seed."""
import numpy as np

from .networks import NeuralNet


def centred_lhs(n, limits, seed=1234):
    """column j = lo_j + (perm_j + 0.5)/n * (hi_j - lo_j), perm_j = RandomState(seed + j).permutation(n)
    (the reference's default LHS criterion: sampling.py:258-260,305-311)"""
    limits = np.asarray(limits, np.float64)
    X = np.empty((n, limits.shape[0]), np.float64)
    for j in range(limits.shape[0]):
        perm = np.random.RandomState(seed + j).permutation(n)
        X[:, j] = limits[j, 0] + (perm + 0.5) / n * (limits[j, 1] - limits[j, 0])
    return X


def glorot_normal_flat(layer_sizes, seed=0):
    return NeuralNet(layer_sizes, seed=seed).get_flat()
