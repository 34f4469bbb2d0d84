"""CelestePy/celeste_sample_sources.pyx call surface: NativePatch and the
photon allocation, on the device (cg_sample_source_counts).

Random numbers: the reference reaches into a numpy RandomState's MT19937 and
consumes it serially over pixels.  The device uses a counter-based stream per
pixel seeded from `random_state` (one 64-bit draw), so a seeded call is
reproducible but not stream-identical to the reference -- the allocation has
the same distribution (tests/test_gpu_parity.py)."""
from __future__ import annotations

import numpy as np
import torch

from .. import _ffi, engine


class NativePatch(object):
    """:31-41: a patch of a field, lower left (x0, y0), upper right (x1, y1)."""

    def __init__(self, data, ylim, xlim):
        self.data = data
        self.x0, self.x1 = int(xlim[0]), int(xlim[1])
        self.y0, self.y1 = int(ylim[0]), int(ylim[1])


def _seed_from(random_state):
    random_state = random_state or np.random.RandomState()
    return int(random_state.randint(0, 2 ** 31 - 1)) * (2 ** 31) + int(random_state.randint(0, 2 ** 31 - 1))


def sample_source_counts(src_imgs, samp_imgs, nelec, src_boxes, epsilon, random_state=None):
    """:61-122 for one image given as arrays: src_imgs / samp_imgs are lists of
    NativePatch, nelec the (H, W) photon counts, src_boxes [S, 4] = (x0, x1, y0,
    y1), epsilon the sky rate.  Returns (samp_imgs, noise_sum); samp_imgs'
    data arrays are filled in place."""
    nelec = np.ascontiguousarray(nelec, dtype=np.float64)
    H, W = nelec.shape
    S = len(src_imgs)
    img = engine.Image(band='r', nelec=nelec, calib=1., gain=1., epsilon=float(epsilon), rho_n=[0., 0.],
                       phi_n=[0., 0.], Ups_n=np.eye(2), psf_weights=[1.], psf_means=np.zeros((1, 2)),
                       psf_covars=np.eye(2)[None])
    boxes = np.zeros(max(S, 1), dtype=_ffi.BOX_DTYPE)
    max_px = 1
    for s in range(S):
        boxes[s] = tuple(int(v) for v in np.asarray(src_boxes)[s])
        max_px = max(max_px, int((boxes[s]["x1"] - boxes[s]["x0"]) * (boxes[s]["y1"] - boxes[s]["y0"])))
    rates = np.zeros((max(S, 1), max_px), dtype=np.float32)
    for s, p in enumerate(src_imgs):
        d = np.asarray(p.data, dtype=np.float32)
        rates[s, :d.size] = d.ravel()
    dev = engine.device()
    d_rates = torch.from_numpy(rates).to(dev)
    d_samp = torch.empty_like(d_rates)
    d_boxes = engine.to_device(boxes)
    L = engine.lib()
    tiles = int(L.cg_field_tiles(H, W))
    ws = torch.empty(int(L.cg_alloc_workspace_bytes(1, tiles)), dtype=torch.uint8, device=dev)
    noise = torch.zeros(1, dtype=torch.float64, device=dev)
    harr, _ = engine._handles([img])
    _ffi.check(L.cg_sample_source_counts(harr, 1, S, d_boxes.data_ptr(), d_rates.data_ptr(), max_px,
                                         _seed_from(random_state), d_samp.data_ptr(), noise.data_ptr(),
                                         ws.data_ptr(), ws.numel(), engine.stream_ptr()))
    out = d_samp.cpu().numpy()
    for s, p in enumerate(samp_imgs):
        h, w = p.y1 - p.y0, p.x1 - p.x0
        np.asarray(p.data)[...] = out[s, :h * w].reshape(h, w)
    return samp_imgs, float(noise.item())


def sample_binomial(n, p, random_state):
    """:159-166."""
    return int(engine.binomial_draws(int(n), float(p), _seed_from(random_state), 1)[0])
