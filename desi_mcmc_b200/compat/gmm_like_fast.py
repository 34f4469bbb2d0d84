"""CelestePy/util/like/gmm_like_fast.pyx call surface on the CUDA path."""
from __future__ import annotations

import numpy as np
import torch

from .. import engine
from .._ffi import check


def _f64(a, name):
    a = np.asarray(a)
    if a.dtype != np.float64:
        raise ValueError("Buffer dtype mismatch, expected 'FLOAT_t' but got '%s' (%s)" % (a.dtype, name))
    return np.ascontiguousarray(a)


def gmm_like_2d(probs, x, ws, mus, sigs):
    """gmm_like_fast.pyx:130-176: probs[n] = sum_k ws[k] N(x[n]; mus[k], sigs[k]),
    written in place (probs is zeroed first, :158-159)."""
    probs_in = probs
    x, ws, mus, sigs = _f64(x, "x"), _f64(ws, "ws"), _f64(mus, "mus"), _f64(sigs, "sigs")
    if mus.shape[0] != sigs.shape[0] or mus.shape[0] != ws.shape[0]:
        raise ValueError("Means, covariances and weights must have same first dimension!")   # :146-147
    if mus.shape[1] != sigs.shape[1] or mus.shape[1] != sigs.shape[2]:
        raise ValueError("Means and inverse covariance shapes don't jive!")                  # :148-149
    if x.ndim != 2 or x.shape[1] != 2 or mus.shape[1] != 2:
        raise ValueError("must be two-dimensional!")
    if probs_in.shape[0] != x.shape[0]:
        raise ValueError("probs and x must have the same first dimension")
    dev = engine.device()
    d = lambda a: torch.from_numpy(a).to(dev)
    out = engine.gmm_like_2d_device(d(x), d(ws), d(mus), d(sigs))
    probs_in[:] = out.cpu().numpy()
    return None


def multivariate_normal_2d_covinv_logdet_logpdf(lls, x, mean, covinv, logdet):
    """gmm_like_fast.pyx:34-77 (logdet is a C float in the reference, :39)."""
    x, mean, covinv = _f64(x, "x"), _f64(mean, "mean"), _f64(covinv, "covinv")
    if mean.shape[0] != x.shape[1]:
        raise ValueError("data dimension 1 doesn't equal mean dimension 0")
    if mean.shape[0] != covinv.shape[0] or mean.shape[0] != covinv.shape[1]:
        raise ValueError("inverse covariance and mean shape aren't equal!")
    if mean.shape[0] != 2:
        raise ValueError("must be two-dimensional!")
    dev = engine.device()
    d_x = torch.from_numpy(x).to(dev)
    d_l = torch.empty(x.shape[0], dtype=torch.float64, device=dev)
    check(engine.lib().cg_mvn2d_logpdf(d_l.data_ptr(), d_x.data_ptr(), mean.ctypes.data,
                                       covinv.ctypes.data, float(logdet), x.shape[0],
                                       engine.stream_ptr()))
    lls[:] = d_l.cpu().numpy()
    return lls


def gmm_like_2d_wrapper(x, ws, mus, sigs, probs=None):
    """CelestePy/util/like/__init__.py:7-11."""
    if probs is None:
        probs = np.zeros(x.shape[0], dtype=np.float64)
    gmm_like_2d(probs, x, ws, mus, sigs)
    return probs
