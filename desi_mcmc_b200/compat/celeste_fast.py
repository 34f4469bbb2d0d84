"""CelestePy/celeste_fast.pyx call surface.  These tables are a few hundred
flops on tiny host arrays; they run in the library's host code (C ABI), the
device kernels build the same tables in their prologue."""
from __future__ import annotations

import numpy as np

from .. import _ffi


def _c(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def gen_galaxy_prof_psf_mixture_params(W, v_s, image_ws, image_means, image_covars,
                                       gal_prof_amp, gal_prof_sigs):
    """celeste_fast.pyx:100-140 -> (weights, means, covars)."""
    L = _ffi.load()
    W, v_s, iw, im, ic = _c(W), _c(v_s), _c(image_ws), _c(image_means), _c(image_covars)
    pa, ps = _c(gal_prof_amp), _c(gal_prof_sigs)
    K_psf, K_prof = iw.shape[0], pa.shape[0]
    K = K_psf * K_prof
    weights, means, covars = np.zeros(K), np.zeros((K, 2)), np.zeros((K, 2, 2))
    _ffi.check(L.cg_galaxy_prof_psf_mixture_params(
        W.ctypes.data, v_s.ctypes.data, iw.ctypes.data, im.ctypes.data, ic.ctypes.data, K_psf,
        pa.ctypes.data, ps.ctypes.data, K_prof, weights.ctypes.data, means.ctypes.data,
        covars.ctypes.data))
    return weights, means, covars


def gen_galaxy_psf_mixture_params(thetas, W, v_s, image_ws, image_means, image_covars,
                                  gal_exp_amp, gal_exp_sigs, gal_dev_amp, gal_dev_sigs):
    """celeste_fast.pyx:29-94 -> (weights, means, covars)."""
    L = _ffi.load()
    th, W, v_s = _c(thetas), _c(W), _c(v_s)
    iw, im, ic = _c(image_ws), _c(image_means), _c(image_covars)
    ea, es, da, ds = _c(gal_exp_amp), _c(gal_exp_sigs), _c(gal_dev_amp), _c(gal_dev_sigs)
    K_psf = iw.shape[0]
    K = K_psf * (ea.shape[0] + da.shape[0])
    weights, means, covars = np.zeros(K), np.zeros((K, 2)), np.zeros((K, 2, 2))
    _ffi.check(L.cg_galaxy_psf_mixture_params(
        th.ctypes.data, W.ctypes.data, v_s.ctypes.data, iw.ctypes.data, im.ctypes.data,
        ic.ctypes.data, K_psf, ea.ctypes.data, es.ctypes.data, ea.shape[0], da.ctypes.data,
        ds.ctypes.data, da.shape[0], weights.ctypes.data, means.ctypes.data, covars.ctypes.data))
    return weights, means, covars
