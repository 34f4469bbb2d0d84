"""Shared helpers of the compat layer: one (source, image) render through the
engine, returned the way the reference returns patches (float64 host array,
(y0, y1), (x0, x1))."""
from __future__ import annotations

import numpy as np

from .. import _ffi, engine

BANDS = ['u', 'g', 'r', 'i', 'z']


def star_record(u, fluxes=None):
    arr = np.zeros(1, dtype=_ffi.SRC_DTYPE)
    arr["type"] = 0
    arr["u"] = np.asarray(u, dtype=np.float64)
    if fluxes is not None:
        arr["flux"] = np.asarray(fluxes, dtype=np.float64)
    return arr


def galaxy_record(th, u, fluxes=None):
    arr = np.zeros(1, dtype=_ffi.SRC_DTYPE)
    arr["type"] = 1
    arr["u"] = np.asarray(u, dtype=np.float64)
    arr["theta"], arr["sigma"], arr["phi"], arr["rho"] = [float(v) for v in th[0:4]]
    if fluxes is not None:
        arr["flux"] = np.asarray(fluxes, dtype=np.float64)
    return arr


def fluxes_of(src):
    f = src.fluxes
    if isinstance(f, dict):
        return np.array([f[b] for b in BANDS], dtype=np.float64)
    return np.asarray(f, dtype=np.float64)


def record_of(src):
    """SrcParams-like -> one CgSource record (stars with a temperature take the
    black-body route, celeste.py:38-40)."""
    return engine.pack_sources([engine.SrcParams(u=src.u, a=src.a, b=src.b, t=src.t, theta=src.theta,
                                                 phi=src.phi, sigma=src.sigma, rho=src.rho,
                                                 fluxes=None if src.fluxes is None else fluxes_of(src))])


def limits_box(xlim, ylim):
    return (int(xlim[0]), int(xlim[1]), int(ylim[0]), int(ylim[1]))


def render_one(image, record, flags, xlim=None, ylim=None):
    """-> (patch float64 (h, w) or None when nothing was rendered, ylim, xlim)."""
    box = None
    if xlim is not None and ylim is not None:
        box = [limits_box(xlim, ylim)]
    items = engine.make_items([0], [0], flags=flags, box=box)
    patches, boxes = engine.render_patches([image], record, items=items)
    b = boxes[0]
    h, w = int(b["y1"] - b["y0"]), int(b["x1"] - b["x0"])
    patch = patches[0, :h * w].cpu().numpy().astype(np.float64).reshape(h, w)
    return patch, (int(b["y0"]), int(b["y1"])), (int(b["x0"]), int(b["x1"]))


def full_frame(image, patch, ylim, xlim, grid=None):
    """The return_patch=False variants: the patch pasted into a frame-sized grid."""
    if grid is None:
        grid = np.zeros(image.shape, dtype=np.float64)
    grid[ylim[0]:ylim[1], xlim[0]:xlim[1]] = patch
    return grid, (0, grid.shape[0]), (0, grid.shape[1])
