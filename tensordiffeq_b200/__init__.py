"""tensordiffeq_b200 -- a B200-native engine behind the TensorDiffEq API.

Same module names as the reference package (tensordiffeq/__init__.py:3); in addition the
top-level re-exports the reference left commented out (__init__.py:5-10) are provided, since
its own examples use them (examples/AC-dist-new.py:78,81)."""
from . import models, optimizers, networks, plotting, utils, domains, boundaries, fit, helpers, sampling, tf  # noqa: F401
from .models import CollocationSolverND, DiscoveryModel  # noqa: F401
from .boundaries import dirichletBC, periodicBC, IC, FunctionDirichletBC, FunctionNeumannBC  # noqa: F401
from .domains import DomainND  # noqa: F401
from .utils import constant, LatinHypercubeSample, tensor  # noqa: F401
from .plotting import get_griddata  # noqa: F401
from .helpers import find_L2_error  # noqa: F401

__all__ = ["models", "networks", "plotting", "utils", "helpers", "optimizers", "boundaries", "domains", "fit",
           "sampling", "tf"]
__version__ = "0.1.0"
