"""Boundary / initial condition objects with the reference's names, constructor arguments and
attributes (tensordiffeq/boundaries.py:12-246).  They build point sets and targets on the host
with NumPy exactly once; their loss terms are evaluated by the CUDA kernels.

`from tensordiffeq_b200.boundaries import *` exports np, tf (the tracer's tf-like namespace)
and DomainND, as the reference module does (boundaries.py:1-4)."""
import numpy as np

from . import tf
from .domains import DomainND
from .utils import multimesh, flatten_and_stack, MSE, convertTensor, get_tf_model

__all__ = ["np", "tf", "DomainND", "BC", "dirichletBC", "FunctionDirichletBC", "FunctionNeumannBC", "IC",
           "periodicBC", "get_linspace"]


def get_linspace(dict_):
    return [val for key, val in dict_.items() if "linspace" in key][0]


def _fidelity(dict_):
    return [val for key, val in dict_.items() if "fidelity" in key][0]


class BC(DomainND):
    def __init__(self):
        self.isPeriodic = False
        self.isInit = False
        self.isNeumann = False
        self.isDirichlect = False

    def compile(self):
        self.input = self.create_input()

    def get_dict(self, var):
        return next(item for item in self.domain.domaindict if item["identifier"] == var)

    def _others(self, var):
        return [item for item in self.domain.domaindict if item["identifier"] != var]

    def get_not_dims(self, var):
        self.dicts_ = self._others(var)
        return [get_linspace(d) for d in self.dicts_]

    def _face_mesh(self, var, value):
        """all combinations of the other variables' linspaces (first varies slowest), with the
        column of `var` held at `value`, inserted at its position in domain.vars
        (boundaries.py:28-38,54-59)"""
        mesh = flatten_and_stack(multimesh(self.get_not_dims(var)))
        col = np.full(len(mesh), float(value))
        return np.insert(mesh, self.domain.vars.index(var), col, axis=1)

    def _function_values(self, funs, groups):
        vals = []
        for f, names in zip(funs, groups):
            args = [get_linspace(self.get_dict(v)) for v in names]
            inp = flatten_and_stack(multimesh(args))
            vals.append(np.asarray(f(*inp.T), dtype=np.float64))
        return np.reshape(vals, (-1, 1))


class dirichletBC(BC):
    """boundaries.py:41-59"""

    def __init__(self, domain, val, var, target):
        super().__init__()
        self.domain = domain
        self.val = val
        self.var = var
        self.dict_ = self.get_dict(var)
        self.dicts_ = self._others(var)
        self.target = self.dict_[var + target]
        self.compile()
        self.isDirichlect = True

    def create_input(self):
        return self._face_mesh(self.var, self.target)


class FunctionDirichletBC(BC):
    """boundaries.py:62-101; n_values draws that many boundary points WITH replacement (:87-89)"""

    def __init__(self, domain, fun, var, target, func_inputs, n_values=None):
        super().__init__()
        self.domain = domain
        self.fun = fun
        self.var = var
        self.func_inputs = func_inputs
        self.n_values = n_values
        self.dict_ = self.get_dict(var)
        self.dicts_ = self._others(var)
        self.target = target
        self.targets = self.dict_[var + target]
        self.nums = None
        self.compile()
        self.create_target()
        self.isDirichlect = True

    def create_input(self):
        mesh = self._face_mesh(self.var, self.dict_[self.var + self.target])
        if self.n_values is not None:
            self.nums = np.random.randint(0, high=len(mesh), size=self.n_values)
            mesh = mesh[self.nums]
        return mesh

    def create_target(self):
        vals = self._function_values(self.fun, self.func_inputs)
        if self.nums is not None:
            vals = vals[self.nums]
        self.val = convertTensor(vals)


class IC(BC):
    """boundaries.py:163-200; t = 0 is appended as the LAST column"""

    def __init__(self, domain, fun, var, n_values=None):
        super().__init__()
        self.isInit = True
        self.n_values = n_values
        self.domain = domain
        self.fun = fun
        self.vars = var
        self.nums = None
        self.dicts_ = self._others(domain.time_var)
        self.dict_ = self.get_dict(domain.time_var)
        self.compile()
        self.create_target()

    def create_input(self):
        mesh = flatten_and_stack(multimesh(self.get_not_dims(self.domain.time_var)))
        mesh = np.concatenate((mesh, np.zeros((len(mesh), 1))), axis=1)
        if self.n_values is not None:
            self.nums = np.random.randint(0, high=len(mesh), size=self.n_values)
            mesh = mesh[self.nums]
        return mesh

    def create_target(self):
        vals = self._function_values(self.fun, self.vars)
        if self.nums is not None:
            vals = vals[self.nums]
        self.val = convertTensor(vals)

    def loss(self):
        return MSE(self.preds, self.val)


class _PairedBC(BC):
    def _bootstrap(self, n_points):
        # boundaries.py:129-132,230-233: an unseeded draw WITH replacement
        size = self.n_values if self.n_values is not None else n_points
        self.nums = np.random.randint(0, high=n_points, size=size)

    def unroll(self, inp):
        """[n_var] arrays [d, n, 1] of float64 columns (boundaries.py:241-246)"""
        return [np.asarray([np.reshape(col, (-1, 1))[self.nums] for col in mesh.T]) for mesh in inp]

    def u_x_model(self, u_model, inputs):
        return [model(u_model, *inputs) for model in self.deriv_model]


class periodicBC(_PairedBC):
    """boundaries.py:205-246"""

    def __init__(self, domain, var, deriv_model, n_values=None):
        super().__init__()
        self.n_values = n_values
        self.domain = domain
        self.var = var
        self.deriv_model = [get_tf_model(m) for m in deriv_model]
        self.isPeriodic = True
        self.compile()

    def compile(self):
        self.upper, self.lower = [], []
        for var in self.var:
            self.dict_ = self.get_dict(var)
            self.upper.append(self._face_mesh(var, self.dict_["range"][1]))
            self.lower.append(self._face_mesh(var, self.dict_["range"][0]))
        self._bootstrap(len(self.upper[0]))
        self.upper = self.unroll(self.upper)
        self.lower = self.unroll(self.lower)


class FunctionNeumannBC(_PairedBC):
    """boundaries.py:103-156"""

    def __init__(self, domain, fun, var, target, deriv_model, func_inputs, n_values=None):
        super().__init__()
        self.n_values = n_values
        self.domain = domain
        self.fun = fun
        self.var = var
        self.target = target
        self.deriv_model = [get_tf_model(m) for m in deriv_model]
        self.isNeumann = True
        self.func_inputs = func_inputs
        self.compile()
        self.create_target()

    def compile(self):
        self.input = []
        for var in self.var:
            self.dict_ = self.get_dict(var)
            value = self.dict_[var + self.target] if isinstance(self.target, str) else self.target
            self.input.append(self._face_mesh(var, value))
        self._bootstrap(len(self.input[0]))
        self.input = self.unroll(self.input)

    def create_target(self):
        self.val = convertTensor(self._function_values(self.fun, self.func_inputs)[self.nums])
