"""Post-processing helpers of tensordiffeq/plotting.py that do not need matplotlib."""
import numpy as np


def get_griddata(grid, data, dims):
    """plotting.py:152-153"""
    from scipy.interpolate import griddata
    return griddata(grid, data, dims, method='cubic')
