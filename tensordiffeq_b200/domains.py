"""DomainND (tensordiffeq/domains.py:5-31): same attributes and dictionary keys."""
import numpy as np

from .utils import LatinHypercubeSample


class DomainND:
    def __init__(self, var, time_var=None):
        self.vars = var
        self.domaindict = []
        self.domain_ids = []
        self.time_var = time_var

    def add(self, token, vals, fidel):
        self.domain_ids.append(token)
        self.domaindict.append({
            "identifier": token,
            "range": vals,
            token + "fidelity": fidel,
            token + "linspace": np.linspace(vals[0], vals[1], fidel),
            token + "upper": vals[1],
            token + "lower": vals[0],
        })

    def generate_collocation_points(self, N_f):
        limits = np.array([d["range"] for d in self.domaindict])
        self.X_f = LatinHypercubeSample(N_f, limits)
