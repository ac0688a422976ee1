"""Post-processing helpers of tensordiffeq/plotting.py that do not need matplotlib."""
import numpy as np


def get_griddata(grid, data, dims):
    """plotting.py:152-153"""
    from scipy.interpolate import griddata
    return griddata(grid, data, dims, method='cubic')


def _no_matplotlib(name):
    def f(*a, **k):
        raise NotImplementedError(
            "tensordiffeq_b200.plotting.%s: figure drawing (tensordiffeq/plotting.py) is outside this engine's scope "
            "and matplotlib is not a dependency; use model.predict() and plot with the tool of your choice" % name)
    f.__name__ = name
    return f


newfig = _no_matplotlib("newfig")
plot_solution_domain1D = _no_matplotlib("plot_solution_domain1D")
plot_weights = _no_matplotlib("plot_weights")
plot_residuals = _no_matplotlib("plot_residuals")
plot_glam_values = _no_matplotlib("plot_glam_values")
