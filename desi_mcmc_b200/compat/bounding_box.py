"""CelestePy/util/bound/bounding_box.py:9-31 through the library's host code
(the device prologue evaluates the same formula per item)."""
from __future__ import annotations

import numpy as np

from .. import _ffi


def calc_bounding_radius(weights, means, covars, error, center=np.array([0, 0])):
    means = np.ascontiguousarray(means, dtype=np.float64)
    covars = np.ascontiguousarray(covars, dtype=np.float64)
    center = np.ascontiguousarray(center, dtype=np.float64)
    if means.shape[0] != covars.shape[0]:
        raise ValueError("means and covariances must have same first dimension")
    return float(_ffi.load().cg_calc_bounding_radius(means.ctypes.data, covars.ctypes.data,
                                                     means.shape[0], float(error), center.ctypes.data))
