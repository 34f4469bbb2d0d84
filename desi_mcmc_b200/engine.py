"""Host side of the B200 render + likelihood engine.

torch owns every device buffer (images, source arrays, patches, workspaces)
and supplies the stream; all arithmetic happens in libceleste_b200.so through
the C ABI (``_ffi``).  Nothing here computes a likelihood on the host.

Data contract mirrors
  * ``Image``     <- CelestePy/fits_image.py:16-155 (FitsImage fields the hot
                     path reads; no FITS I/O)
  * ``SrcParams`` <- CelestePy/celeste_src.py:12-157
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterable, Sequence

import numpy as np
import torch

from . import _ffi
from ._ffi import (BOX_DTYPE, CG_ITEM_FIXED_BOX, CG_ITEM_FLUX_BAND, CG_ITEM_FLUX_UNIT,
                   CG_LL_CONDITIONAL, CG_LL_POISSON_BG, CG_MAX_IMAGES, CG_NUM_BANDS,
                   CG_SRC_EIGEN, CG_SRC_PLANCK, ITEM_DTYPE, SRC_DTYPE, check)

BANDS = ["u", "g", "r", "i", "z"]

_state = {"device": None, "lib": None, "ws": {}, "filters": False}


# ---------------------------------------------------------------------------
# life cycle
# ---------------------------------------------------------------------------
def init(device: int | None = None):
    """Load the library and bind it to one GPU (one process per GPU)."""
    lib = _ffi.load()
    if not torch.cuda.is_available():
        raise RuntimeError("desi_mcmc_b200 needs a CUDA device: the render/likelihood path "
                           "has no CPU fallback")
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if _state["device"] == device:
        return lib
    torch.cuda.set_device(device)
    torch.zeros(1, device="cuda:%d" % device)      # create the primary context torch shares
    check(lib.cg_init(device))
    _state.update(device=device, lib=lib, ws={}, filters=False)
    return lib


def lib():
    if _state["lib"] is None:
        return init()
    return _state["lib"]


def device() -> torch.device:
    lib()
    return torch.device("cuda", _state["device"])


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def sm_count() -> int:
    return lib().cg_sm_count()


def launch_count(reset: bool = False) -> int:
    return int(lib().cg_launch_count(1 if reset else 0))


def last_eval_count() -> int:
    """pixel x component evaluations of the last engine call (synchronises)."""
    n = int(lib().cg_last_eval_count(stream_ptr()))
    if n < 0:
        raise MemoryError(_ffi.last_error())
    return n


def last_issued_count() -> int:
    """lane-evaluations the render kernel issued for the last call (after
    component culling, idle lanes of a warp included)."""
    return int(lib().cg_last_issued_count(stream_ptr()))


def set_option(option: int, value: int) -> None:
    check(lib().cg_set_option(int(option), int(value)))


def profile_render(enable: bool) -> None:
    """Start (and reset) / stop per-launch CUDA-event timing of render_kernel."""
    check(lib().cg_profile_render(1 if enable else 0))


def profile_render_read():
    """-> (total milliseconds, launches) of the render_kernel launches recorded."""
    ms, n = C.c_double(0), C.c_int64(0)
    check(lib().cg_profile_render_read(C.byref(ms), C.byref(n)))
    return ms.value, n.value


def pipe_probe(which: int = 0) -> float:
    """Measured MUFU.EX2 (0) or FP32-FMA (1) instructions*lanes per second."""
    v = C.c_double(0)
    check(lib().cg_pipe_probe(which, C.byref(v), stream_ptr()))
    return v.value


def _workspace(kind: int, n_items: int, n_img: int, max_patch_px: int, n_tiles: int = 0):
    need = int(lib().cg_workspace_bytes(kind, int(n_items), int(n_img), int(max_patch_px), int(n_tiles)))
    key = kind
    ws = _state["ws"].get(key)
    if ws is None or ws.numel() < need:
        ws = torch.empty(max(need, 1 << 20), dtype=torch.uint8, device=device())
        _state["ws"][key] = ws
    return ws


def to_device(a: np.ndarray) -> torch.Tensor:
    """Host POD array -> device bytes (pageable host memory; see PinnedStage
    for the steady-state path)."""
    a = np.ascontiguousarray(a)
    t = torch.from_numpy(a.view(np.uint8).reshape(-1))
    return t.to(device(), non_blocking=False)


class PinnedStage(object):
    """A pinned host buffer + device twin for per-step H2D of POD arrays."""

    def __init__(self, nbytes: int):
        self.host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        self.dev = torch.empty(nbytes, dtype=torch.uint8, device=device())
        self.np = self.host.numpy()

    def upload(self, a: np.ndarray) -> torch.Tensor:
        raw = a.view(np.uint8).reshape(-1)
        n = raw.size
        self.np[:n] = raw
        self.dev[:n].copy_(self.host[:n], non_blocking=True)
        return self.dev


# ---------------------------------------------------------------------------
# data contract
# ---------------------------------------------------------------------------
class SrcParams(object):
    """celeste_src.py:12-157.  ``fluxes`` may be a length-5 array (ugriz) or a
    band->value dict, as the reference accepts both."""

    def __init__(self, u, a=None, b=None, t=None, v=None, theta=None, phi=None,
                 sigma=None, rho=None, fluxes=None, ell=None, d=None, header=None, flux_row=None):
        self.a = a
        self.flux_row = flux_row          # row of the bound flux table (eigen-spectrum route), or None
        self.u = np.asarray(u, dtype=np.float64)
        self.b, self.t = b, t
        self.v = v
        self.theta, self.phi, self.sigma, self.rho = theta, phi, sigma, rho
        if isinstance(fluxes, dict):
            fluxes = [fluxes[c] for c in BANDS]
        self.fluxes = None if fluxes is None else np.asarray(fluxes, dtype=np.float64)
        self.ell, self.d, self.header = ell, d, header

    @property
    def flux_dict(self):
        return {c: v for c, v in zip(BANDS, self.fluxes)}

    @property
    def shape(self):
        return np.array([self.theta, self.sigma, self.phi, self.rho])

    def is_star(self):
        return self.a == 0

    def is_galaxy(self):
        return self.a == 1


def pack_sources(srcs: Sequence, out: np.ndarray | None = None) -> np.ndarray:
    """SrcParams-like objects -> CgSource array (include/celeste_b200.h).
    A star with a truthy temperature takes the black-body route, exactly the
    reference's test ``if src.t`` (celeste.py:38)."""
    n = len(srcs)
    arr = np.zeros(n, dtype=SRC_DTYPE) if out is None else out[:n]
    for i, s in enumerate(srcs):
        r = arr[i]
        if s.a not in (0, 1):
            raise ValueError("source type unknown")
        r["type"] = int(s.a)
        r["u"] = s.u
        flags = 0
        if s.fluxes is not None:
            r["flux"] = [s.fluxes[c] for c in BANDS] if isinstance(s.fluxes, dict) else s.fluxes
        if getattr(s, "flux_row", None) is not None:
            flags |= CG_SRC_EIGEN
            r["flux_row"] = int(s.flux_row)
        elif s.a == 0 and s.t:
            flags |= CG_SRC_PLANCK
            r["t"] = s.t
            r["b"] = s.b
        if s.a == 1:
            r["theta"], r["sigma"], r["phi"], r["rho"] = s.theta, s.sigma, s.phi, s.rho
        r["flags"] = flags
    return arr


def pack_source_fields(type_, u, flux, theta=None, sigma=None, phi=None, rho=None,
                       t=None, b=None) -> np.ndarray:
    """Vectorised packer for arrays of parameters (proposal batches)."""
    u = np.asarray(u, dtype=np.float64)
    n = u.shape[0]
    arr = np.zeros(n, dtype=SRC_DTYPE)
    arr["type"] = type_
    arr["u"] = u
    arr["flux"] = flux
    for name, val in (("theta", theta), ("sigma", sigma), ("phi", phi), ("rho", rho)):
        if val is not None:
            arr[name] = val
    if t is not None:
        arr["t"] = t
        arr["b"] = b
        arr["flags"] = np.where((np.asarray(type_) == 0) & (np.asarray(t) != 0), CG_SRC_PLANCK, 0)
    return arr


class Image(object):
    """Device-resident mirror of FitsImage (fits_image.py:16-155): the caller
    passes what the FITS header/frames would hold.  ``nelec`` is uploaded once
    (float32: photo-electron counts are integers below 2**24) and stays in HBM."""

    def __init__(self, band, nelec, calib, gain, rho_n, phi_n, Ups_n, psf_weights, psf_means,
                 psf_covars, sky=None, epsilon=None, invvar=None):
        L = lib()
        self.band = band if isinstance(band, str) else BANDS[band]
        self.band_index = BANDS.index(self.band)
        if isinstance(nelec, torch.Tensor):
            self.d_nelec = nelec.to(device(), dtype=torch.float32).contiguous()
        else:
            self.d_nelec = torch.from_numpy(np.ascontiguousarray(nelec, dtype=np.float32)).to(device())
        self.shape = tuple(self.d_nelec.shape)
        if len(self.shape) != 2:
            raise ValueError("nelec must be 2-d")
        self.d_mask = None
        if invvar is not None:
            if isinstance(invvar, torch.Tensor):
                self.d_mask = invvar.to(device(), dtype=torch.float32).contiguous()
            else:
                self.d_mask = torch.from_numpy(np.ascontiguousarray(invvar, dtype=np.float32)).to(device())
            if tuple(self.d_mask.shape) != self.shape:
                raise ValueError("invvar shape must equal nelec shape")
        self.rho_n = np.asarray(rho_n, dtype=np.float64)
        self.phi_n = np.asarray(phi_n, dtype=np.float64)
        self.Ups_n = np.asarray(Ups_n, dtype=np.float64).reshape(2, 2)
        self.Ups_n_inv = np.linalg.inv(self.Ups_n)
        self.kappa = float(gain)
        if epsilon is None:
            if sky is None:
                raise ValueError("pass sky or epsilon")
            epsilon = float(sky) * self.kappa            # fits_image.py:113,119
        self.epsilon = float(epsilon)
        self.calib = float(calib)
        self.weights = np.asarray(psf_weights, dtype=np.float64)
        self.means = np.asarray(psf_means, dtype=np.float64).reshape(-1, 2)
        self.covars = np.asarray(psf_covars, dtype=np.float64).reshape(-1, 2, 2)
        K = self.weights.shape[0]
        if not (self.means.shape[0] == K == self.covars.shape[0]):
            raise ValueError("PSF weights, means and covariances must have same first dimension")
        if K > _ffi.CG_MAX_PSF:
            raise ValueError("at most %d PSF components" % _ffi.CG_MAX_PSF)
        d = _ffi.CgImageDesc()
        d.H, d.W = self.shape
        d.band, d.n_psf = self.band_index, K
        d.epsilon, d.kappa, d.calib = self.epsilon, self.kappa, self.calib
        for k in range(K):
            d.psf_w[k] = self.weights[k]
            d.psf_mu[k][0], d.psf_mu[k][1] = self.means[k]
            d.psf_cov[k][0] = self.covars[k, 0, 0]
            d.psf_cov[k][1] = self.covars[k, 1, 1]
            d.psf_cov[k][2] = self.covars[k, 0, 1]
        d.rho_n[0], d.rho_n[1] = self.rho_n
        d.phi_n[0], d.phi_n[1] = self.phi_n
        for j, v in enumerate(self.Ups_n.ravel()):
            d.ups[j] = v
        d.d_nelec = self.d_nelec.data_ptr()
        d.d_mask = self.d_mask.data_ptr() if self.d_mask is not None else None
        h = C.c_void_p()
        check(L.cg_image_create(C.byref(d), C.byref(h)))
        self._h = h
        self.R = float(L.cg_image_radius(h))             # fits_image.py:151-155

    # host conveniences (float64, a handful of flops; fits_image.py:166-216)
    def equa2pixel(self, s_equa):
        phi1rad = self.phi_n[1] / 180. * np.pi
        s_iwc = np.array([(s_equa[0] - self.phi_n[0]) * np.cos(phi1rad), (s_equa[1] - self.phi_n[1])])
        return np.dot(self.Ups_n_inv, s_iwc) + self.rho_n

    def pixel2equa(self, s_pixel):
        phi1rad = self.phi_n[1] / 180. * np.pi
        s_iwc = np.dot(self.Ups_n, np.asarray(s_pixel, dtype=np.float64) - self.rho_n)
        return np.array([s_iwc[0] / np.cos(phi1rad) + self.phi_n[0], s_iwc[1] + self.phi_n[1]])

    def nmgy2counts(self, flux):
        return (flux / self.calib) * self.kappa

    def cd_at_pixel(self, x, y):
        """fits_image.py:196-216: finite-difference CD matrix (10 px step)."""
        ra0, dec0 = self.pixel2equa(np.array([x, y]))
        step = 10.
        rax, decx = self.pixel2equa(np.array([x + step, y]))
        ray, decy = self.pixel2equa(np.array([x, y + step]))
        cosd = np.cos(dec0 * (np.pi / 180.))
        return np.array([[(rax - ra0) / step * cosd, (ray - ra0) / step * cosd],
                         [(decx - dec0) / step, (decy - dec0) / step]])

    def contains(self, s_equa, pad=50):
        """fits_image.py:157-164."""
        px, py = self.equa2pixel(s_equa)
        return (px > -pad) and (px < self.shape[1] + pad) and (py > -pad) and (py < self.shape[0] + pad)

    @property
    def psf(self):
        """fits_image.py:219-221: the PSF as a mixture-of-Gaussians object."""
        if getattr(self, "_psf", None) is None:
            from .compat.mog import MixtureOfGaussians
            self._psf = MixtureOfGaussians(self.means, self.covars, self.weights)
        return self._psf

    @property
    def invcovars(self):
        return self.psf.icovs

    @property
    def logdets(self):
        return np.log(self.psf.dets)

    @property
    def invvar(self):
        """host copy of the mask (all ones when the image has none)."""
        if self.d_mask is None:
            return np.ones(self.shape)
        return self.d_mask.cpu().numpy().astype(np.float64)

    def set_epsilon(self, epsilon: float):
        """Field.resample_photons re-draws epsilon every sweep (models.py:156-160)."""
        check(lib().cg_image_set_epsilon(self._h, float(epsilon)))
        self.epsilon = float(epsilon)

    @property
    def nelec(self) -> torch.Tensor:
        return self.d_nelec

    def set_nelec(self, nelec):
        """Overwrite the resident pixel counts in place (same device buffer)."""
        if not isinstance(nelec, torch.Tensor):
            nelec = torch.from_numpy(np.ascontiguousarray(nelec, dtype=np.float32))
        self.d_nelec.copy_(nelec.to(dtype=torch.float32).reshape(self.shape))

    def close(self):
        if getattr(self, "_h", None) is not None and _state["lib"] is not None:
            _state["lib"].cg_image_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _handles(images: Sequence[Image]):
    n = len(images)
    if n < 1 or n > CG_MAX_IMAGES:
        raise ValueError("between 1 and %d images per call" % CG_MAX_IMAGES)
    arr = (C.c_void_p * n)(*[im._h for im in images])
    return arr, n


def _srcs_dev(srcs):
    """Accepts a packed CgSource array (numpy), a device byte tensor, or a
    sequence of SrcParams; returns (device tensor, count)."""
    if isinstance(srcs, torch.Tensor):
        return srcs, srcs.numel() // SRC_DTYPE.itemsize
    if isinstance(srcs, np.ndarray) and srcs.dtype == SRC_DTYPE:
        _need_filters(srcs)
        return to_device(srcs), srcs.shape[0]
    arr = pack_sources(list(srcs))
    _need_filters(arr)                       # black-body stars need the filter curves on the device
    return to_device(arr), arr.shape[0]


def read_filter_table(path: str | None = None):
    """planck.py:31-55: per band, wavelengths (m, ascending), sensitivities and
    h_vec = S/(hc/lambda)*dLambda.  Default table: the SDSS ugriz response
    curves packed in desi_mcmc_b200/data/filter_curves.npz (Angstrom); a path
    to a text file in the reference's "band<TAB>angstrom<TAB>sens" layout is
    also accepted."""
    h, c = 6.6260693e-34, 299792458
    wl, sens = {}, {}
    if path is None:
        z = np.load(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
        for b in BANDS:
            wl[b] = z["wl_" + b] * 1e-4 * 1e-6
            sens[b] = z["sens_" + b]
    else:
        acc_w = {b: [] for b in BANDS}
        acc_s = {b: [] for b in BANDS}
        with open(path) as f:
            for line in f:
                band, wavelength, sensitivity = line.strip().split("\t")
                acc_w[band].append(float(wavelength) * 1e-4 * 1e-6)
                acc_s[band].append(float(sensitivity))
        for b in BANDS:
            wl[b], sens[b] = np.array(acc_w[b]), np.array(acc_s[b])
    hv = {}
    for b in BANDS:
        order = wl[b].argsort()
        wl[b], sens[b] = wl[b][order], sens[b][order]
        dLam = np.abs(wl[b][2] - wl[b][1])
        hv[b] = sens[b] / (h * c / wl[b]) * dLam
    return wl, sens, hv


def load_filter_curves(path: str | None = None):
    """Hand the filter tables to the device (cg_set_filter_curves)."""
    wl, sens, hv = read_filter_table(path)
    wl_all = np.ascontiguousarray(np.concatenate([wl[b] for b in BANDS]))
    hv_all = np.ascontiguousarray(np.concatenate([hv[b] for b in BANDS]))
    s_all = np.ascontiguousarray(np.concatenate([sens[b] for b in BANDS]))
    cnt = np.asarray([wl[b].size for b in BANDS], dtype=np.int32)
    check(lib().cg_set_filter_curves(wl_all.ctypes.data, hv_all.ctypes.data, s_all.ctypes.data,
                                     cnt.ctypes.data))
    _state["filters"] = True


def _need_filters(arr_or_srcs):
    if _state["filters"]:
        return
    need = False
    if isinstance(arr_or_srcs, np.ndarray):
        need = bool(np.any(arr_or_srcs["flags"] & CG_SRC_PLANCK))
    if need:
        load_filter_curves()


# ---------------------------------------------------------------------------
# engine calls
# ---------------------------------------------------------------------------
def make_items(src, img, flags=0, box=None, aux=-1) -> np.ndarray:
    """Build a CgItem array from broadcastable columns; box = (x0, x1, y0, y1)."""
    src = np.atleast_1d(np.asarray(src, dtype=np.int32))
    n = src.shape[0]
    it = np.zeros(n, dtype=ITEM_DTYPE)
    it["src"] = src
    it["img"] = img
    it["flags"] = flags
    it["aux"] = aux
    if box is not None:
        it["box"] = np.asarray(box, dtype=np.int32).reshape(-1, 4)
        it["flags"] |= CG_ITEM_FIXED_BOX
    return it


def render_patches(images: Sequence[Image], srcs, items: np.ndarray | None = None,
                   default_flags: int = CG_ITEM_FLUX_BAND, bound_cap: float = 0.0,
                   max_patch_px: int | None = None):
    """A5/A6/A10: rate (or unit-flux) patches for (source, image) items.
    Returns (patches [n_items, max_patch_px] float32 device tensor,
    boxes [n_items] BOX_DTYPE host array).  items=None renders the cross
    product, item = src*n_img + img."""
    L = lib()
    harr, n_img = _handles(images)
    if isinstance(srcs, np.ndarray):
        _need_filters(srcs)
    d_srcs, n_src = _srcs_dev(srcs)
    if items is None:
        n_items = n_src * n_img
        d_items = None
    else:
        n_items = items.shape[0]
        d_items = to_device(items)
    if max_patch_px is None:
        max_patch_px = max(im.shape[0] * im.shape[1] for im in images)
        if bound_cap > 0:
            side = int(2 * bound_cap + 3)
            max_patch_px = min(max_patch_px, side * side)
    if n_items == 0:
        return (torch.zeros((0, max_patch_px), dtype=torch.float32, device=device()),
                np.zeros(0, dtype=BOX_DTYPE))
    patches = torch.zeros((n_items, max_patch_px), dtype=torch.float32, device=device())
    boxes = torch.zeros(n_items * BOX_DTYPE.itemsize, dtype=torch.uint8, device=device())
    ws = _workspace(_ffi.CG_WS_RENDER, n_items, n_img, max_patch_px)
    check(L.cg_render_patches(harr, n_img, d_srcs.data_ptr(), n_src,
                              d_items.data_ptr() if d_items is not None else None, n_items,
                              default_flags, float(bound_cap), max_patch_px,
                              patches.data_ptr(), boxes.data_ptr(), ws.data_ptr(), ws.numel(),
                              stream_ptr()))
    last_eval_count()          # surfaces footprint/work-list overflow as an exception
    return patches, boxes.cpu().numpy().view(BOX_DTYPE)


def binomial_draws(n: int, p: float, seed: int, count: int) -> np.ndarray:
    """sample_binomial (celeste_sample_sources.pyx:161-166): count device draws."""
    out = torch.zeros(count, dtype=torch.float64, device=device())
    check(lib().cg_binomial_draws(int(n), float(p), int(seed), int(count), out.data_ptr(), stream_ptr()))
    return out.cpu().numpy()


class PhotonAllocation(object):
    """Result of resample_photons: device patches in the item = src*n_img + img
    layout.  rates / samp: [n_items, max_patch_px] float32; boxes: host BOX_DTYPE
    array; noise_sums: host float64 [n_img]."""

    def __init__(self, rates, samp, boxes, d_boxes, noise_sums, n_src, n_img):
        self.rates, self.samp, self.boxes, self.d_boxes = rates, samp, boxes, d_boxes
        self.noise_sums, self.n_src, self.n_img = noise_sums, n_src, n_img

    def patch(self, src: int, img: int, which: str = "samp") -> np.ndarray:
        it = src * self.n_img + img
        b = self.boxes[it]
        h, w = int(b["y1"] - b["y0"]), int(b["x1"] - b["x0"])
        buf = self.samp if which == "samp" else self.rates
        return buf[it, :h * w].cpu().numpy().astype(np.float64).reshape(h, w)


def resample_photons(images: Sequence[Image], srcs, seed: int, bound_cap: float = 0.0,
                     max_patch_px: int | None = None, flags: int = CG_ITEM_FLUX_BAND) -> PhotonAllocation:
    """N1 (celeste_mcmc.py:98-150 + celeste_sample_sources.pyx:61-122) for all
    images of a field: render every source's rate patch, then split every
    pixel's photons over the overlapping sources and the sky."""
    L = lib()
    harr, n_img = _handles(images)
    if isinstance(srcs, np.ndarray):
        _need_filters(srcs)
    d_srcs, n_src = _srcs_dev(srcs)
    if max_patch_px is None:
        side = int(2 * bound_cap + 3) if bound_cap > 0 else 203
        max_patch_px = side * side
    n_items = n_src * n_img
    rates = torch.zeros((max(n_items, 1), max_patch_px), dtype=torch.float32, device=device())
    samp = torch.empty_like(rates)
    d_boxes = torch.zeros(max(n_items, 1) * BOX_DTYPE.itemsize, dtype=torch.uint8, device=device())
    if n_items:
        ws = _workspace(_ffi.CG_WS_RENDER, n_items, n_img, max_patch_px)
        check(L.cg_render_patches(harr, n_img, d_srcs.data_ptr(), n_src, None, n_items, flags,
                                  float(bound_cap), max_patch_px, rates.data_ptr(), d_boxes.data_ptr(),
                                  ws.data_ptr(), ws.numel(), stream_ptr()))
    tiles = n_img * max(int(L.cg_field_tiles(im.shape[0], im.shape[1])) for im in images)
    need = int(L.cg_alloc_workspace_bytes(n_img, tiles))
    aws = _state["ws"].get("alloc")
    if aws is None or aws.numel() < need:
        aws = torch.empty(need, dtype=torch.uint8, device=device())
        _state["ws"]["alloc"] = aws
    noise = torch.zeros(n_img, dtype=torch.float64, device=device())
    check(L.cg_sample_source_counts(harr, n_img, n_src, d_boxes.data_ptr(), rates.data_ptr(), max_patch_px,
                                    int(seed) & 0xFFFFFFFFFFFFFFFF, samp.data_ptr(), noise.data_ptr(),
                                    aws.data_ptr(), aws.numel(), stream_ptr()))
    if n_items:
        last_eval_count()
    boxes = d_boxes.cpu().numpy().view(BOX_DTYPE)[:n_items]
    return PhotonAllocation(rates, samp, boxes, d_boxes, noise.cpu().numpy(), n_src, n_img)


def field_workspace(images: Sequence[Image], n_src: int, max_patch_px: int):
    n_img = len(images)
    tiles = n_img * max(int(lib().cg_field_tiles(im.shape[0], im.shape[1])) for im in images)
    return _workspace(_ffi.CG_WS_FIELD, max(n_src, 1) * n_img, n_img, max_patch_px, tiles)


def field_loglik(images: Sequence[Image], srcs, bound_cap: float = 0.0,
                 max_patch_px: int | None = None, return_model: bool = False,
                 out: torch.Tensor | None = None, sync: bool = True):
    """A7/A8 (celeste.py:203-252): per-image Poisson log-likelihood of the full
    frames under lambda = epsilon + sum_s patch_s.  Returns a float64 device
    tensor [n_img] (and the model images when return_model)."""
    L = lib()
    harr, n_img = _handles(images)
    if isinstance(srcs, np.ndarray):
        _need_filters(srcs)
    d_srcs, n_src = _srcs_dev(srcs)
    if max_patch_px is None:
        side = int(2 * bound_cap + 3) if bound_cap > 0 else 203
        max_patch_px = side * side
    ws = field_workspace(images, n_src, max_patch_px)
    ll = out if out is not None else torch.zeros(n_img, dtype=torch.float64, device=device())
    models, mptr = None, None
    if return_model:
        models = [torch.zeros(im.shape, dtype=torch.float32, device=device()) for im in images]
        mptr = (C.c_void_p * n_img)(*[m.data_ptr() for m in models])
    check(L.cg_field_loglik(harr, n_img, d_srcs.data_ptr(), n_src, float(bound_cap), max_patch_px,
                            ll.data_ptr(), mptr, ws.data_ptr(), ws.numel(), stream_ptr()))
    if sync:
        last_eval_count()
    return (ll, models) if return_model else ll


class FieldSet(object):
    """A batch of independent fields (models.py:75-79 loops the fields of a
    CelesteBase) scored by ONE cg_fields_loglik call: one prepare, one render
    and one gather launch for all of them.  Buffers are allocated once.

    fields: list of (images, n_src_max).  loglik() takes, per field, either a
    device byte tensor of packed CgSource or a host CgSource array (copied
    through a pinned stage -- the per-step host->device traffic)."""

    def __init__(self, image_lists: Sequence[Sequence[Image]], n_src_max: Sequence[int],
                 bound_cap: float = 0.0, max_patch_px: int | None = None, d_srcs: torch.Tensor | None = None,
                 h_srcs: torch.Tensor | None = None, ll: torch.Tensor | None = None):
        """d_srcs / h_srcs / ll: optional views into rank-level buffers (device source bytes, pinned
        source bytes, device log-likelihoods) so that several sets share one contiguous source array
        and one log-likelihood vector (engine.SweepPayload reads them as a whole)."""
        self.L = lib()
        self.fields = [list(imgs) for imgs in image_lists]
        F = len(self.fields)
        if F < 1 or F > _ffi.CG_MAX_FIELDS:
            raise ValueError("between 1 and %d fields per set" % _ffi.CG_MAX_FIELDS)
        flat = [im for imgs in self.fields for im in imgs]
        self.harr, self.n_img = _handles(flat)
        self.img_off = np.zeros(F + 1, dtype=np.int32)
        self.img_off[1:] = np.cumsum([len(f) for f in self.fields])
        self.n_src_max = np.asarray(n_src_max, dtype=np.int64)
        self.cap_off = np.zeros(F + 1, dtype=np.int64)
        self.cap_off[1:] = np.cumsum(self.n_src_max)
        self.bound_cap = float(bound_cap)
        if max_patch_px is None:
            side = int(2 * bound_cap + 3) if bound_cap > 0 else 203
            max_patch_px = side * side
        self.max_patch_px = int(max_patch_px)
        n_items = int(sum(int(n) * len(f) for n, f in zip(self.n_src_max, self.fields)))
        tiles = self.n_img * max(int(self.L.cg_field_tiles(im.shape[0], im.shape[1])) for im in flat)
        need = int(self.L.cg_workspace_bytes(_ffi.CG_WS_FIELD, max(n_items, 1), self.n_img,
                                             self.max_patch_px, tiles))
        self.ws = torch.empty(need, dtype=torch.uint8, device=device())
        nb = int(self.cap_off[-1]) * SRC_DTYPE.itemsize
        self.d_srcs = d_srcs if d_srcs is not None else torch.zeros(max(nb, 1), dtype=torch.uint8, device=device())
        self.h_srcs = h_srcs if h_srcs is not None else torch.zeros(max(nb, 1), dtype=torch.uint8).pin_memory()
        if self.d_srcs.numel() < nb or self.h_srcs.numel() < nb:
            raise ValueError("source buffers too small for n_src_max")
        self.h_np = self.h_srcs.numpy()
        self.ll = ll if ll is not None else torch.zeros(self.n_img, dtype=torch.float64, device=device())
        if self.ll.numel() != self.n_img:
            raise ValueError("ll must have one entry per image")
        self.h_ll = torch.zeros(self.n_img, dtype=torch.float64).pin_memory()
        self.src_off = np.zeros(F + 1, dtype=np.int32)

    def set_sources(self, per_field: Sequence[np.ndarray]) -> int:
        """Host CgSource arrays -> pinned stage -> device (async); returns bytes copied."""
        o = 0
        for f, arr in enumerate(per_field):
            n = arr.shape[0]
            if n > self.n_src_max[f]:
                raise ValueError("field %d: more sources than n_src_max" % f)
            raw = arr.view(np.uint8).reshape(-1)
            self.h_np[o:o + raw.size] = raw
            self.src_off[f] = o // SRC_DTYPE.itemsize
            o += raw.size
        self.src_off[len(per_field)] = o // SRC_DTYPE.itemsize
        self.d_srcs[:o].copy_(self.h_srcs[:o], non_blocking=True)
        return o

    def pinned_sources(self, f: int, n: int | None = None) -> np.ndarray:
        """A CgSource view INTO the pinned staging buffer for field f (n sources,
        default n_src_max[f]; fields are packed back to back, so set the counts
        of fields 0..f-1 first).  A sampler that keeps its state here pays no
        host copy per step: upload() is the only per-step host work."""
        n = int(self.n_src_max[f]) if n is None else int(n)
        if n > self.n_src_max[f]:
            raise ValueError("field %d: more sources than n_src_max" % f)
        start = int(self.src_off[f])
        self.src_off[f + 1:] = np.maximum(self.src_off[f + 1:], start + n)
        self.src_off[f + 1] = start + n
        b0 = start * SRC_DTYPE.itemsize
        return self.h_np[b0:b0 + n * SRC_DTYPE.itemsize].view(SRC_DTYPE)

    def upload(self) -> int:
        """Pinned staging buffer -> device (async on the current stream); returns bytes."""
        o = int(self.src_off[len(self.fields)]) * SRC_DTYPE.itemsize
        self.d_srcs[:o].copy_(self.h_srcs[:o], non_blocking=True)
        return o

    def loglik_device(self) -> torch.Tensor:
        """Sources already resident (set_sources called before); nothing synchronises."""
        check(self.L.cg_fields_loglik(self.harr, self.img_off.ctypes.data, self.d_srcs.data_ptr(),
                                      self.src_off.ctypes.data, len(self.fields), self.bound_cap,
                                      self.max_patch_px, self.ll.data_ptr(), None,
                                      self.ws.data_ptr(), self.ws.numel(), stream_ptr()))
        return self.ll

    def loglik(self, per_field: Sequence[np.ndarray]) -> np.ndarray:
        """End to end: host source arrays in, per-image log-likelihoods (host) out."""
        self.set_sources(per_field)
        self.loglik_device()
        self.h_ll.copy_(self.ll, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        if np.isnan(self.h_ll.numpy()).any():      # a device-side error poisons the results (field_tile_final_kernel)
            last_eval_count()
            raise RuntimeError("the field log-likelihood is NaN")
        return self.h_ll.numpy()


class FieldSweep(object):
    """Every field set of a rank scored as ONE step (models.py:75-79 loops the
    fields; they are independent, so their launches may overlap).

    The sets are enqueued round-robin on ``n_streams`` side streams forked from
    the current stream, so the HBM-bound gather of one set runs under the
    XU-bound render of the next; ``capture()`` records the whole step once as a
    CUDA graph and ``run()`` replays it (one driver call per step instead of
    4 launches + 2 memsets per set).  The number of sources per field is baked
    into the captured launches: call ``capture()`` again after ``set_sources``
    changed a count (parameter VALUES may change freely)."""

    def __init__(self, sets: Sequence["FieldSet"], n_streams: int = 2, after=None, io: bool = False,
                 parallel=None):
        """after: optional callable enqueued on the current stream after the sets
        of a step (e.g. the all-reduce of the sweep statistics); it is captured
        into the graph with them.  io=True makes the step self-contained end to
        end: it starts with every set's upload() (pinned source stage -> device)
        and ends with the copy of the per-image log-likelihoods to the pinned
        h_ll, so that a replay is ONE driver call including the transfers."""
        self.sets = list(sets)
        self.after = after
        self.parallel = parallel       # callable enqueued on a forked stream beside the field kernels
        self._par_stream = torch.cuda.Stream(device=device()) if parallel is not None else None
        self.io = bool(io)
        self.n_streams = max(1, min(int(n_streams), len(self.sets)))
        self._side = [torch.cuda.Stream(device=device()) for _ in range(self.n_streams)] \
            if self.n_streams > 1 else []
        self.graph = None
        self._counts = None

    def _enqueue(self):
        if self.io:
            for fs in self.sets:
                fs.upload()
        self._enqueue_compute()
        if self.io:
            for fs in self.sets:
                fs.h_ll.copy_(fs.ll, non_blocking=True)

    def _enqueue_compute(self):
        cur = torch.cuda.current_stream()
        if self.parallel is not None:          # small independent work (prior statistics) in the shadow
            self._par_stream.wait_stream(cur)  # of the latency-bound set-up kernels of the field path
            with torch.cuda.stream(self._par_stream):
                self.parallel()
        if not self._side:
            for fs in self.sets:
                fs.loglik_device()
        else:
            for s in self._side:
                s.wait_stream(cur)
            for i, fs in enumerate(self.sets):
                with torch.cuda.stream(self._side[i % len(self._side)]):
                    fs.loglik_device()
            for s in self._side:
                cur.wait_stream(s)
        if self.parallel is not None:
            cur.wait_stream(self._par_stream)
        if self.after is not None:
            self.after()

    def capture(self):
        self._enqueue()                      # warm-up outside the capture (lazy occupancy queries)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._enqueue()
        self.graph = g
        self._counts = [fs.src_off.copy() for fs in self.sets]
        return self

    def run(self):
        """Enqueue one step on the current stream; nothing synchronises."""
        if self.graph is not None:
            for fs, c in zip(self.sets, self._counts):
                if not np.array_equal(fs.src_off, c):
                    raise RuntimeError("source counts changed since capture(): capture again")
            self.graph.replay()
        else:
            self._enqueue()

    def loglik_all(self) -> torch.Tensor:
        return torch.cat([fs.ll for fs in self.sets])


def items_loglik(images: Sequence[Image], srcs, items: np.ndarray, mode: int = CG_LL_POISSON_BG,
                 aux: torch.Tensor | None = None, aux_offsets: np.ndarray | None = None,
                 bound_cap: float = 0.0, max_patch_px: int = 0, use_fp64: bool = False):
    """A13/A14/A15 building block: per-item log-likelihood over its box.
    Returns (item_ll float64 device tensor, boxes host array)."""
    L = lib()
    harr, n_img = _handles(images)
    d_srcs, n_src = _srcs_dev(srcs)
    n_items = items.shape[0]
    d_items = to_device(items)
    if max_patch_px <= 0:
        max_patch_px = max(im.shape[0] * im.shape[1] for im in images)
    ws = _workspace(_ffi.CG_WS_ITEMS, n_items, n_img, max_patch_px)
    ll = torch.zeros(n_items, dtype=torch.float64, device=device())
    boxes = torch.zeros(n_items * BOX_DTYPE.itemsize, dtype=torch.uint8, device=device())
    d_off = None
    if aux is not None:
        if aux_offsets is None:
            raise ValueError("aux_offsets required with aux")
        d_off = to_device(np.ascontiguousarray(aux_offsets, dtype=np.int64))
    check(L.cg_items_loglik(harr, n_img, d_srcs.data_ptr(), n_src, d_items.data_ptr(), n_items,
                            mode, float(bound_cap), max_patch_px,
                            aux.data_ptr() if aux is not None else None,
                            d_off.data_ptr() if d_off is not None else None,
                            ll.data_ptr(), boxes.data_ptr(), 1 if use_fp64 else 0,
                            ws.data_ptr(), ws.numel(), stream_ptr()))
    last_eval_count()
    return ll, boxes.cpu().numpy().view(BOX_DTYPE)


class ProposalScorer(object):
    """A15: batched Source.calculate_acceptance_logprob + accept test
    (sources.py:258,275-306) for P proposals of one source over its cached
    bounding boxes / background patches.  Buffers are allocated once; score()
    takes HOST arrays (packed proposals, log-terms, log-uniforms) and returns
    host arrays -- the per-step H2D/D2H the end-to-end numbers include."""

    #: fp32 error of a score per unit of sqrt(sum of squared counts over the boxes): measured 4e-8
    #: (|delta_fp32 - delta_fp64| <= 2.6e-3 on config 3, 64x64x5 pixels at ~450 e-), carried with a 4x margin
    GUARD_PER_RMS_COUNT = 1.6e-7

    def __init__(self, images: Sequence[Image], boxes, backgrounds=None, max_proposals: int = 4096,
                 guard="auto"):
        """guard: half-width of the band around a tie in which a decision is re-made in float64.
        "auto" (default): max(1e-2, GUARD_PER_RMS_COUNT * sqrt(sum nelec^2 over the boxes)) -- the
        fp32 error of a score grows with the photon counts it multiplies, so a bright source gets a
        wider band; a number fixes it; 0 disables the re-score."""
        self.L = lib()
        self.images = list(images)
        self.harr, self.n_img = _handles(self.images)
        self.P_max = int(max_proposals)
        if guard == "auto":
            ss = 0.0
            for im, bx in zip(self.images, boxes):
                x0, x1, y0, y1 = [int(v) for v in bx]
                ss += float(im.d_nelec[max(y0, 0):y1, max(x0, 0):x1].double().square().sum().item())
            guard = max(1e-2, self.GUARD_PER_RMS_COUNT * np.sqrt(ss))
        self.guard = float(guard)
        b = np.zeros(self.n_img, dtype=BOX_DTYPE)
        for i, bx in enumerate(boxes):          # (x0, x1, y0, y1)
            b[i] = tuple(int(v) for v in bx)
        self.boxes = b
        max_px = int(max((bb["x1"] - bb["x0"]) * (bb["y1"] - bb["y0"]) for bb in b))
        self.max_px = max_px
        self.d_bg, self.bg_off = None, None
        if backgrounds is not None:
            offs, flat, o = [], [], 0
            for i, bg in enumerate(backgrounds):
                bg = np.ascontiguousarray(bg, dtype=np.float32)
                bb = b[i]
                if bg.shape != (bb["y1"] - bb["y0"], bb["x1"] - bb["x0"]):
                    raise ValueError("background %d shape %s does not match its box" % (i, bg.shape))
                offs.append(o)
                flat.append(bg.ravel())
                o += bg.size
            self.bg_off = np.asarray(offs, dtype=np.int64)
            self.d_bg = torch.from_numpy(np.concatenate(flat)).to(device())
        n_items = (self.P_max + 1) * self.n_img
        need = int(self.L.cg_workspace_bytes(_ffi.CG_WS_SCORE, n_items, self.n_img, max_px, 0))
        self.ws = torch.empty(need, dtype=torch.uint8, device=device())
        self.stage_src = PinnedStage((self.P_max + 1) * SRC_DTYPE.itemsize)
        self.stage_terms = PinnedStage(self.P_max * 8)
        self.stage_u = PinnedStage(self.P_max * 8)
        self.d_delta = torch.zeros(self.P_max, dtype=torch.float64, device=device())
        self.d_accept = torch.zeros(self.P_max, dtype=torch.uint8, device=device())
        self.d_nguard = torch.zeros(1, dtype=torch.int32, device=device())
        self.h_delta = torch.zeros(self.P_max, dtype=torch.float64).pin_memory()
        self.h_accept = torch.zeros(self.P_max, dtype=torch.uint8).pin_memory()
        self.h_nguard = torch.zeros(1, dtype=torch.int32).pin_memory()

    def score_device(self, d_srcs: torch.Tensor, P: int, d_terms, d_log_u):
        """All inputs already in HBM; nothing synchronises."""
        check(self.L.cg_score_proposals(
            self.harr, self.n_img, d_srcs.data_ptr(), int(P), self.boxes.ctypes.data,
            self.d_bg.data_ptr() if self.d_bg is not None else None,
            self.bg_off.ctypes.data if self.bg_off is not None else None,
            d_terms.data_ptr() if d_terms is not None else None, d_log_u.data_ptr(),
            self.guard, self.d_delta.data_ptr(), self.d_accept.data_ptr(),
            self.d_nguard.data_ptr(), self.ws.data_ptr(), self.ws.numel(), stream_ptr()))

    def score(self, cur_and_props: np.ndarray, log_u: np.ndarray, log_terms: np.ndarray | None = None):
        """cur_and_props: CgSource array, [0] = current state, [1:] = proposals.
        Returns (delta [P] float64, accept [P] bool, n_guard)."""
        P = cur_and_props.shape[0] - 1
        if P > self.P_max:
            raise ValueError("more proposals than max_proposals")
        if log_u.shape[0] != P:
            raise ValueError("log_u must have one entry per proposal")
        d_srcs = self.stage_src.upload(cur_and_props)
        d_u = self.stage_u.upload(np.ascontiguousarray(log_u, dtype=np.float64))
        d_t = None
        if log_terms is not None:
            d_t = self.stage_terms.upload(np.ascontiguousarray(log_terms, dtype=np.float64))
        self.score_device(d_srcs, P, d_t, d_u)
        self.h_delta[:P].copy_(self.d_delta[:P], non_blocking=True)
        self.h_accept[:P].copy_(self.d_accept[:P], non_blocking=True)
        self.h_nguard.copy_(self.d_nguard, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return (self.h_delta[:P].numpy().copy(), self.h_accept[:P].numpy().astype(bool),
                int(self.h_nguard[0]))


def gmm_like_2d_device(d_x, d_ws, d_mus, d_sigs, d_probs=None):
    """A2 on device tensors (float64)."""
    N, K = d_x.shape[0], d_ws.shape[0]
    if d_probs is None:
        d_probs = torch.empty(N, dtype=torch.float64, device=device())
    check(lib().cg_gmm_like_2d(d_probs.data_ptr(), d_x.data_ptr(), d_ws.data_ptr(), d_mus.data_ptr(),
                               d_sigs.data_ptr(), N, K, stream_ptr()))
    return d_probs


def planck_fluxes(T, b) -> np.ndarray:
    """A16: photons_expected_brightness(T[i], b[i], band) for all 5 bands
    (planck.py:155-158) -> [n, 5] float64."""
    if not _state["filters"]:
        load_filter_curves()
    T = np.atleast_1d(np.asarray(T, dtype=np.float64))
    b = np.broadcast_to(np.asarray(b, dtype=np.float64), T.shape).copy()
    d_T = torch.from_numpy(T).to(device())
    d_b = torch.from_numpy(b).to(device())
    out = torch.zeros((T.shape[0], CG_NUM_BANDS), dtype=torch.float64, device=device())
    check(lib().cg_planck_fluxes(d_T.data_ptr(), d_b.data_ptr(), T.shape[0], out.data_ptr(), stream_ptr()))
    return out.cpu().numpy()


def eigen_fluxes(z, w, m, B, lam0, fluxes=None, fluxes_ivar=None):
    """A17: batched project_to_bands / pixel_likelihood
    (experiments/redshift/redshift_utils.py:304-323,
    quasar_infer_photometry.py:31-46).  Returns (mu [C,5], ll [C] or None)."""
    if not _state["filters"]:
        load_filter_curves()
    z = np.atleast_1d(np.asarray(z, dtype=np.float64))
    Cn = z.shape[0]
    w = np.ascontiguousarray(np.asarray(w, dtype=np.float64).reshape(Cn, -1))
    m = np.broadcast_to(np.asarray(m, dtype=np.float64), z.shape).copy()
    B = np.ascontiguousarray(B, dtype=np.float64)
    K, V = B.shape
    if w.shape[1] != K:
        raise ValueError("weights and basis disagree on K")
    dev = device()
    d = lambda a: torch.from_numpy(np.array(a, dtype=np.float64, order="C")).to(dev)   # a writable copy
    d_z, d_w, d_m, d_B, d_l = d(z), d(w), d(m), d(B), d(lam0)
    mu = torch.zeros((Cn, CG_NUM_BANDS), dtype=torch.float64, device=dev)
    ll, d_f, d_iv = None, None, None
    if fluxes is not None:
        d_f = d(np.broadcast_to(np.asarray(fluxes, dtype=np.float64), (Cn, CG_NUM_BANDS)))
        d_iv = d(np.broadcast_to(np.asarray(fluxes_ivar, dtype=np.float64), (Cn, CG_NUM_BANDS)))
        ll = torch.zeros(Cn, dtype=torch.float64, device=dev)
    check(lib().cg_eigen_fluxes(d_z.data_ptr(), d_w.data_ptr(), d_m.data_ptr(), d_B.data_ptr(),
                                d_l.data_ptr(), Cn, K, V,
                                d_f.data_ptr() if d_f is not None else None,
                                d_iv.data_ptr() if d_iv is not None else None,
                                mu.data_ptr(), ll.data_ptr() if ll is not None else None, stream_ptr()))
    return mu.cpu().numpy(), (ll.cpu().numpy() if ll is not None else None)


# ---------------------------------------------------------------------------
# N3: empirical priors (util/dists/mog.py, flux_prior.py) on the device
# ---------------------------------------------------------------------------
def mog_logpdf(x, means, icovs, dets, pis, return_resp: bool = False):
    """mog_loglike (util/dists/mog.py:5-21) for D-dimensional points, float64.
    x [N, D] (or [D]); returns log-densities [N] (and responsibilities [N, K])."""
    x = np.asarray(x, dtype=np.float64)
    one = x.ndim == 1
    xx = np.atleast_2d(x)
    means = np.ascontiguousarray(means, dtype=np.float64)
    K, D = means.shape
    if xx.shape[1] != D:
        raise ValueError("x and means disagree on the dimension")
    dev = device()
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    d_x, d_m, d_i, d_d, d_p = d(xx), d(means), d(icovs), d(dets), d(pis)
    N = xx.shape[0]
    logp = torch.zeros(max(N, 1), dtype=torch.float64, device=dev)
    resp = torch.zeros((max(N, 1), K), dtype=torch.float64, device=dev) if return_resp else None
    check(lib().cg_mog_logpdf(d_x.data_ptr(), N, D, d_m.data_ptr(), d_i.data_ptr(), d_d.data_ptr(),
                              d_p.data_ptr(), K, logp.data_ptr(),
                              resp.data_ptr() if resp is not None else None, stream_ptr()))
    out = logp[:N].cpu().numpy()
    if one:
        out = out[0]
    return (out, resp[:N]) if return_resp else out


def prior_sufficient_stats(resp: torch.Tensor, colors) -> torch.Tensor:
    """[sum r_k | sum r_k c | sum r_k c c^T (upper triangle)] over the points:
    the block of dist.SweepStats a rank contributes for one prior (device tensor)."""
    dev = device()
    c = torch.from_numpy(np.ascontiguousarray(np.atleast_2d(colors), dtype=np.float64)).to(dev)
    N, K = resp.shape
    D = c.shape[1]
    out = torch.zeros(K * (1 + D + D * (D + 1) // 2), dtype=torch.float64, device=dev)
    check(lib().cg_prior_stats(resp.data_ptr(), c.data_ptr(), N, K, D, out.data_ptr(), stream_ptr()))
    return out


# ---------------------------------------------------------------------------
# device-resident Gibbs sweep over a crowded field (BASELINE config 4)
# ---------------------------------------------------------------------------
def bind_flux_table(table: torch.Tensor | None) -> None:
    """cg_set_flux_table: band fluxes [rows, 5] (float64, nanomaggies, device) for sources
    flagged CG_SRC_EIGEN (flux_row = row).  Pass None to unbind."""
    if table is None:
        check(lib().cg_set_flux_table(None, 0))
        _state["flux_table"] = None
        return
    if table.dtype != torch.float64 or table.dim() != 2 or table.shape[1] != CG_NUM_BANDS or not table.is_cuda:
        raise ValueError("flux table must be a float64 CUDA tensor of shape [rows, 5]")
    table = table.contiguous()
    check(lib().cg_set_flux_table(table.data_ptr(), table.shape[0]))
    _state["flux_table"] = table            # keep it alive while it is bound


class EigenSpectra(object):
    """Device state of the eigen-spectrum flux model (experiments/redshift/redshift_utils.py:304-323,
    quasar_infer_photometry.py:31-46): per source z, simplex weights w over the K basis spectra and
    the magnitude scale m.  project() evaluates mu = project_to_bands(w.B, lam0 (1+z)) m / (1+z) for
    every row into the bound flux table -- the render reads it from HBM (no host hop per band)."""

    def __init__(self, z, w, m, B, lam0):
        if not _state["filters"]:
            load_filter_curves()
        dev = device()
        d = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
        self.z, self.w, self.m, self.B, self.lam0 = d(np.atleast_1d(z)), d(np.atleast_2d(w)), d(np.atleast_1d(m)), d(B), d(lam0)
        self.n = self.z.shape[0]
        self.K, self.V = self.B.shape
        if self.w.shape != (self.n, self.K) or self.m.shape[0] != self.n or self.lam0.shape[0] != self.V:
            raise ValueError("z, w, m, B, lam0 disagree on their sizes")
        self.mu = torch.zeros((self.n, CG_NUM_BANDS), dtype=torch.float64, device=dev)
        bind_flux_table(self.mu)

    def project(self) -> torch.Tensor:
        check(lib().cg_eigen_fluxes(self.z.data_ptr(), self.w.data_ptr(), self.m.data_ptr(), self.B.data_ptr(),
                                    self.lam0.data_ptr(), self.n, self.K, self.V, None, None,
                                    self.mu.data_ptr(), None, stream_ptr()))
        return self.mu


def static_boxes(images: Sequence[Image], u: np.ndarray, pixel_radius: float) -> np.ndarray:
    """make_bbox_dict (sources.py:434-455) for every (source, image): boxes [n_src, n_img] BOX_DTYPE."""
    u = np.asarray(u, dtype=np.float64).reshape(-1, 2)
    out = np.zeros((u.shape[0], len(images)), dtype=BOX_DTYPE)
    for b, im in enumerate(images):
        phi1rad = im.phi_n[1] / 180. * np.pi
        iwc = np.stack([(u[:, 0] - im.phi_n[0]) * np.cos(phi1rad), u[:, 1] - im.phi_n[1]])
        px, py = np.dot(im.Ups_n_inv, iwc) + im.rho_n[:, None]
        H, W = im.shape
        out["x0"][:, b] = np.maximum(0, np.floor(px - pixel_radius).astype(np.int64))
        out["x1"][:, b] = np.minimum(W, np.ceil(px + pixel_radius).astype(np.int64))
        out["y0"][:, b] = np.maximum(0, np.floor(py - pixel_radius).astype(np.int64))
        out["y1"][:, b] = np.minimum(H, np.ceil(py + pixel_radius).astype(np.int64))
    return out


def active_neighbours(boxes: np.ndarray):
    """get_active_sources (sources.py:457-472) for every (source, image): CSR (offsets [n_src*n_img+1],
    indices) of the OTHER sources whose box intersects -- the reference's test, as written:
    |xa0 - xb0| * 2 < wa + wb and |ya0 - yb0| * 2 < ha + hb.  Sorted by source index."""
    S, n_img = boxes.shape
    offs = np.zeros(S * n_img + 1, dtype=np.int32)
    idx = []
    for b in range(n_img):
        x0, y0 = boxes["x0"][:, b].astype(np.int64), boxes["y0"][:, b].astype(np.int64)
        w = boxes["x1"][:, b].astype(np.int64) - x0
        h = boxes["y1"][:, b].astype(np.int64) - y0
        order = np.argsort(x0, kind="stable")
        xs = x0[order]
        wmax = int(w.max()) if S else 0
        for s in range(S):
            lo = np.searchsorted(xs, x0[s] - wmax, side="left")
            hi = np.searchsorted(xs, x0[s] + wmax, side="right")
            c = order[lo:hi]
            ok = (np.abs(x0[c] - x0[s]) * 2 < w[c] + w[s]) & (np.abs(y0[c] - y0[s]) * 2 < h[c] + h[s]) & (c != s)
            nb = np.sort(c[ok])
            idx.append((s * n_img + b, nb))
    idx.sort(key=lambda t: t[0])
    flat = []
    for k, (t, nb) in enumerate(idx):
        offs[t + 1] = nb.size
        flat.append(nb)
    offs = np.cumsum(offs, dtype=np.int64).astype(np.int32)
    return offs, (np.concatenate(flat).astype(np.int32) if flat else np.zeros(0, dtype=np.int32))


class GibbsSweep(object):
    """One field's sources advanced by full Gibbs sweeps ENTIRELY on the device
    (CelesteBase.resample_model, models.py:75-83, for all sources at once):

      1. eigen-spectrum fluxes projected into the flux table (A17; Planck stars inside prepare, A16)
      2. photon allocation for the five bands (models.py:123-160)            cg_field_sample_photons
      3. epsilon ~ Gamma per image (models.py:156-160)                       cg_resample_epsilon
      4. fluxes ~ Gamma per source and band (sources.py:327-350)             cg_resample_fluxes
      5. P proposals per source scored by the conditional likelihood given its photons
         (Source.log_likelihood, sources.py:134-183), accept/reject on the device from host-supplied
         log-uniforms, first accepted proposal taken (exact: the sources are independent given z)
      6. optionally P more proposals scored by the full MH ratio over the cached boxes and
         backgrounds (calculate_acceptance_logprob, sources.py:275-306; backgrounds =
         generate_background_patch, :476-483, re-rendered every sweep)
      7. the field's per-image log-likelihood (celeste.py:237-252)

    Nothing synchronises inside sweep(); stats() reads the results back."""

    DEFAULT_SCALES = (0.05 * 1.1e-4, 0.02, 0.02, 0.02, 3.0, 0.03)    # synth.make_proposals

    def __init__(self, images: Sequence[Image], srcs: np.ndarray, bound_cap: float = 25.0, proposals: int = 16,
                 pixel_radius: float = 25.0, scales=None, guard: float = 1e-2, eigen: EigenSpectra | None = None,
                 background_pass: bool = True, a0: float = 5., b0: float = .005, seed: int = 0):
        self.L = lib()
        self.images = list(images)
        self.harr, self.n_img = _handles(self.images)
        _need_filters(srcs)
        self.n_src = int(srcs.shape[0])
        self.P = int(proposals)
        self.bound_cap, self.pixel_radius = float(bound_cap), float(pixel_radius)
        self.guard, self.a0, self.b0 = float(guard), float(a0), float(b0)
        self.eigen = eigen
        self.background_pass = bool(background_pass)
        self.scales = np.asarray(self.DEFAULT_SCALES if scales is None else scales, dtype=np.float64)
        self.seed, self.n_sweeps = int(seed), 0
        dev = device()
        side = int(max(2 * bound_cap + 3, 2 * pixel_radius + 2))
        self.side = side
        self.max_patch_px = side * side
        n_items = self.n_src * self.n_img
        self.n_items = n_items
        self.d_cur = to_device(srcs)
        self.d_states = torch.zeros(self.n_src * (self.P + 1) * SRC_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        self.samp = torch.zeros((max(n_items, 1), self.max_patch_px), dtype=torch.float32, device=dev)
        self.d_boxes = torch.zeros(max(n_items, 1) * BOX_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        self.noise = torch.zeros(self.n_img, dtype=torch.float64, device=dev)
        self.counts = torch.zeros(max(n_items, 1), dtype=torch.float64, device=dev)
        self.unit = torch.zeros(max(n_items, 1), dtype=torch.float64, device=dev)
        self.eps = torch.zeros(self.n_img, dtype=torch.float64, device=dev)
        self.ll = torch.zeros(self.n_img, dtype=torch.float64, device=dev)
        n_sp = max(self.n_src * self.P, 1)
        self.delta = [torch.zeros(n_sp, dtype=torch.float64, device=dev) for _ in range(2)]
        self.accept = [torch.zeros(n_sp, dtype=torch.uint8, device=dev) for _ in range(2)]
        self.n_guard = torch.zeros(2, dtype=torch.int32, device=dev)
        self.n_acc = torch.zeros(2, dtype=torch.int32, device=dev)
        self.stage_u = [PinnedStage(n_sp * 8) for _ in range(2)]
        tiles = self.n_img * max(int(self.L.cg_field_tiles(im.shape[0], im.shape[1])) for im in self.images)
        need = lambda kind: int(self.L.cg_sweep_workspace_bytes(kind, self.n_src, self.n_img, self.P,
                                                                self.max_patch_px, tiles))
        self.ws_alloc = torch.empty(need(_ffi.CG_WS_SWEEP_ALLOC), dtype=torch.uint8, device=dev)
        self.ws_score = torch.empty(need(_ffi.CG_WS_SWEEP_SCORE), dtype=torch.uint8, device=dev)
        self.ws_field = field_workspace(self.images, self.n_src, self.max_patch_px)
        self.workspace_bytes = self.ws_alloc.numel() + self.ws_score.numel() + self.ws_field.numel() + \
            self.samp.numel() * 4
        if self.background_pass:
            boxes = static_boxes(self.images, srcs["u"], self.pixel_radius)
            self.h_static = boxes
            w = (boxes["x1"] - boxes["x0"]).max() if n_items else 0
            h = (boxes["y1"] - boxes["y0"]).max() if n_items else 0
            if max(w, h) > side:
                raise ValueError("static boxes exceed the patch side")
            offs, idx = active_neighbours(boxes)
            self.h_nbr = (offs, idx)
            self.d_static = to_device(boxes.reshape(-1))
            self.d_nbr_off = torch.from_numpy(offs).to(dev)
            self.d_nbr_idx = torch.from_numpy(idx if idx.size else np.zeros(1, dtype=np.int32)).to(dev)
            self.bg = torch.zeros((max(n_items, 1), self.max_patch_px), dtype=torch.float32, device=dev)
            self.ws_bg = torch.empty(need(_ffi.CG_WS_SWEEP_BACKGROUND), dtype=torch.uint8, device=dev)
            self.workspace_bytes += self.ws_bg.numel() + self.bg.numel() * 4

    # ---- the individual steps (each one C-ABI call; usable on their own) ---------------
    def allocate(self, seed: int):
        check(self.L.cg_field_sample_photons(
            self.harr, self.n_img, self.d_cur.data_ptr(), self.n_src, self.bound_cap, self.max_patch_px,
            int(seed) & 0xFFFFFFFFFFFFFFFF, self.samp.data_ptr(), self.d_boxes.data_ptr(), self.noise.data_ptr(),
            self.counts.data_ptr(), self.unit.data_ptr(), self.ws_alloc.data_ptr(), self.ws_alloc.numel(),
            stream_ptr()))

    def resample_epsilon(self, seed: int):
        check(self.L.cg_resample_epsilon(self.harr, self.n_img, self.noise.data_ptr(), self.a0, self.b0,
                                         int(seed) & 0xFFFFFFFFFFFFFFFF, self.eps.data_ptr(), stream_ptr()))

    def resample_fluxes(self, seed: int):
        check(self.L.cg_resample_fluxes(self.harr, self.n_img, self.d_cur.data_ptr(), self.n_src,
                                        self.counts.data_ptr(), self.unit.data_ptr(), self.a0, self.b0,
                                        int(seed) & 0xFFFFFFFFFFFFFFFF, stream_ptr()))

    def propose(self, seed: int):
        check(self.L.cg_make_proposals(self.d_cur.data_ptr(), self.n_src, self.P, self.scales.ctypes.data,
                                       int(seed) & 0xFFFFFFFFFFFFFFFF, self.d_states.data_ptr(), stream_ptr()))

    def score(self, which: int, log_u: np.ndarray):
        """which = 0: conditional on the allocated photons; 1: full MH ratio with backgrounds."""
        d_u = self.stage_u[which].upload(np.ascontiguousarray(log_u, dtype=np.float64))
        if which == 0:
            mode, boxes, aux = CG_LL_CONDITIONAL, self.d_boxes, self.samp
        else:
            mode, boxes, aux = CG_LL_POISSON_BG, self.d_static, self.bg
        check(self.L.cg_score_proposals_multi(
            self.harr, self.n_img, self.d_states.data_ptr(), self.n_src, self.P, boxes.data_ptr(), mode,
            aux.data_ptr(), self.max_patch_px, None, d_u.data_ptr(), self.guard, self.delta[which].data_ptr(),
            self.accept[which].data_ptr(), self.n_guard[which:].data_ptr(), self.ws_score.data_ptr(),
            self.ws_score.numel(), stream_ptr()))

    def select(self, which: int):
        check(self.L.cg_select_proposals(self.d_states.data_ptr(), self.n_src, self.P,
                                         self.accept[which].data_ptr(), self.d_cur.data_ptr(),
                                         self.n_acc[which:].data_ptr(), stream_ptr()))

    def backgrounds(self):
        check(self.L.cg_background_patches(
            self.harr, self.n_img, self.d_cur.data_ptr(), self.n_src, self.d_static.data_ptr(),
            self.d_nbr_off.data_ptr(), self.d_nbr_idx.data_ptr(), self.max_patch_px, self.side,
            self.bg.data_ptr(), self.ws_bg.data_ptr(), self.ws_bg.numel(), stream_ptr()))

    def slice_locations(self, seed: int, n_uniforms: int = 256, sigma: float = 1.0):
        """Source.resample_location (sources.py:308-319) for every source given the allocated photons, the
        slice samplers run on the device in ONE launch (cg_slice_sample_locations).  The uniforms come from
        one host RandomState(seed): row s of a [n_src, n_uniforms] block is source s's stream.  Near-ties
        are decided in float64 inside the kernel; a source that exhausts its block or meets a NaN (status
        != 0, not expected with the default block) raises.
        Returns (status, n_evals) host arrays."""
        rng = np.random.RandomState(int(seed) & 0x7fffffff)
        order = np.where(rng.uniform(size=(self.n_src, 1)) < .5, [[0, 1]], [[1, 0]]).astype(np.int32)
        uni = rng.random_sample((self.n_src, n_uniforms))
        ll, used, status, evals = slice_sample_locations(self.images, self.d_cur, self.n_src, self.d_boxes, self.samp,
                                                         self.max_patch_px, uni, order, sigma=sigma, guard=self.guard)
        self.slice_stats = dict(ll=ll, used=used, status=status, evals=evals,
                                refined=slice_sample_locations.last_refined,
                                kernel_ms=slice_sample_locations.last_kernel_ms)
        if np.any(status != 0):
            raise RuntimeError("on-device slice sampler: %d sources stopped with status %s (raise n_uniforms)" %
                               (int((status != 0).sum()), sorted(set(status[status != 0].tolist()))))
        return status, evals

    def hmc_galaxies(self, seed: int, step: float = 1e-5, n_steps: int = 50, momentum: np.ndarray = None):
        """hmc_sample_galaxy_params (celeste_mcmc.py:281-323) for every galaxy given the allocated photons, the
        whole trajectory on the device in ONE launch (cg_hmc_galaxy_params) on the sweep's boxes.  Normals
        and the acceptance uniform come from one host RandomState(seed).  momentum: [n_gal, 11] carried
        from the previous call (None: drawn).  Returns (momentum, accepted)."""
        rng = np.random.RandomState(int(seed) & 0x7fffffff)
        types = self.sources()["type"]
        gal = np.nonzero(types == 1)[0].astype(np.int32)
        G = gal.size
        sm = np.sqrt(HMC_GALAXY_MASS)
        p_old = sm * rng.randn(G, 11) if momentum is None else np.asarray(momentum, dtype=np.float64)
        p_new = .9 * p_old + np.sqrt(1. - .9 ** 2) * sm * rng.randn(G, 11)
        log_u = np.log(rng.rand(G))
        ke = np.sum(p_old * p_old / HMC_GALAXY_MASS, axis=1)
        p_out, h, acc, status = hmc_galaxy_params(self.images, self.d_cur, G, self.d_boxes, self.samp, self.max_patch_px,
                                                  p_new, ke, log_u, step=step, n_steps=n_steps, sel=gal, header_cd=False)
        self.hmc_stats = dict(h=h, accepted=acc, status=status, kernel_ms=hmc_galaxy_params.last_kernel_ms,
                              likelihoods=G * (11 * (n_steps + 1) + 2))
        return p_out, acc

    def loglik(self):
        check(self.L.cg_field_loglik(self.harr, self.n_img, self.d_cur.data_ptr(), self.n_src, self.bound_cap,
                                     self.max_patch_px, self.ll.data_ptr(), None, self.ws_field.data_ptr(),
                                     self.ws_field.numel(), stream_ptr()))

    # ---- one sweep -------------------------------------------------------------------------
    def sweep(self, log_u=None):
        """log_u: optional pair of host arrays [n_src * P] (conditional pass, background pass);
        default: drawn from numpy RandomState(seed + sweep index)."""
        k = self.n_sweeps
        base = (self.seed + 1) * 1000003 + k * 16
        n_sp = self.n_src * self.P
        if log_u is None:
            rng = np.random.RandomState((self.seed + k) & 0x7fffffff)
            log_u = (np.log(rng.uniform(size=n_sp)), np.log(rng.uniform(size=n_sp)))
        if self.eigen is not None:
            self.eigen.project()
        self.allocate(base + 1)
        self.resample_epsilon(base + 2)
        self.resample_fluxes(base + 3)
        if self.P > 0 and self.n_src > 0:
            self.propose(base + 4)
            self.score(0, log_u[0])
            self.select(0)
            if self.background_pass:
                self.backgrounds()
                self.propose(base + 5)
                self.score(1, log_u[1])
                self.select(1)
        self.loglik()
        self.n_sweeps += 1

    def stats(self) -> dict:
        """Synchronises; refreshes the host copies of epsilon."""
        torch.cuda.current_stream().synchronize()
        eps = self.eps.cpu().numpy()
        for im, e in zip(self.images, eps):
            check(self.L.cg_image_sync_epsilon(im._h, stream_ptr(), None))
            im.epsilon = float(e)
        delta = [d.cpu().numpy() for d in self.delta]
        if any(np.isnan(d).any() for d in delta) or np.isnan(self.ll.cpu().numpy()).any():
            last_eval_count()                 # raises with the device's error message when it set one
            raise RuntimeError("a device error poisoned the sweep results")
        return dict(loglik=self.ll.cpu().numpy(), epsilon=eps, noise=self.noise.cpu().numpy(),
                    accepted=self.n_acc.cpu().numpy(), guard=self.n_guard.cpu().numpy())

    def sources(self) -> np.ndarray:
        return self.d_cur.cpu().numpy().view(SRC_DTYPE).copy()

    def states(self) -> np.ndarray:
        return self.d_states.cpu().numpy().view(SRC_DTYPE).reshape(self.n_src, self.P + 1).copy()

    def boxes(self) -> np.ndarray:
        return self.d_boxes.cpu().numpy().view(BOX_DTYPE)[:self.n_items].reshape(self.n_src, self.n_img).copy()


class SweepPayload(object):
    """What a rank contributes to the per-sweep all-reduce (dist.SweepStats layout), computed on the
    device from its resident sources and written straight into the collective's send buffer:
    sum of the per-image log-likelihoods, sky photons per band, star / galaxy counts and the
    sufficient statistics of the star and galaxy colour priors (cg_flux_prior_stats), SURVEY.md 8e."""

    def __init__(self, d_srcs: torch.Tensor, n_src: int, ll: torch.Tensor, star_prior, gal_prior,
                 noise: torch.Tensor | None = None, img_bands=None, n_buffers: int = 2):
        """star_prior / gal_prior: (means [K,5], covs [K,5,5], pis [K]) host arrays."""
        self.L = lib()
        dev = device()
        self.d_srcs, self.n_src, self.ll = d_srcs, int(n_src), ll
        self.priors = []
        for means, covs, pis in (star_prior, gal_prior):
            means = np.ascontiguousarray(means, dtype=np.float64)
            covs = np.ascontiguousarray(covs, dtype=np.float64)
            K = means.shape[0]
            d = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
            self.priors.append(dict(K=K, means=d(means), icovs=d(np.linalg.inv(covs)), dets=d(np.linalg.det(covs)),
                                    pis=d(pis)))
        self.K = self.priors[0]["K"]
        if self.priors[1]["K"] != self.K:
            raise ValueError("both priors must have the same number of components")
        self.block = self.K * (1 + 5 + 15)
        self.size = 8 + 2 * self.block
        self.blocks = [torch.zeros(self.block, dtype=torch.float64, device=dev) for _ in range(2)]
        nb = int(self.L.cg_flux_prior_scratch_bytes(self.n_src, self.K))
        self.scratch = [torch.empty(nb, dtype=torch.uint8, device=dev) for _ in range(2)]
        self.vec = [torch.zeros(self.size, dtype=torch.float64, device=dev) for _ in range(n_buffers)]
        self.noise = noise
        self.d_bands = None
        if noise is not None:
            self.d_bands = torch.from_numpy(np.ascontiguousarray(img_bands, dtype=np.int32)).to(dev)

    def prior_stats(self):
        """two launches of three small kernels each; independent of the field kernels"""
        for t, (pr, blk, scr) in enumerate(zip(self.priors, self.blocks, self.scratch)):
            check(self.L.cg_flux_prior_stats(self.d_srcs.data_ptr(), self.n_src, t, pr["means"].data_ptr(),
                                             pr["icovs"].data_ptr(), pr["dets"].data_ptr(), pr["pis"].data_ptr(),
                                             pr["K"], scr.data_ptr(), scr.numel(), None, blk.data_ptr(), stream_ptr()))

    def pack(self, k: int = 0) -> torch.Tensor:
        check(self.L.cg_sweep_stats(self.ll.data_ptr(), self.ll.numel(),
                                    self.noise.data_ptr() if self.noise is not None else None,
                                    self.d_bands.data_ptr() if self.d_bands is not None else None,
                                    self.noise.numel() if self.noise is not None else 0,
                                    self.d_srcs.data_ptr(), self.n_src, self.blocks[0].data_ptr(),
                                    self.blocks[1].data_ptr(), self.block, self.vec[k].data_ptr(), stream_ptr()))
        return self.vec[k]


HMC_GALAXY_MASS = np.array([.1, .1, 1., .1, .001, .001, 1., 1., 1., 1., 1.])    # celeste_mcmc.py:300-303


def hmc_galaxy_params(images: Sequence[Image], d_srcs: torch.Tensor, n_src: int, d_boxes: torch.Tensor,
                      photons: torch.Tensor, max_patch_px: int, p_in: np.ndarray, ke_curr: np.ndarray,
                      log_u: np.ndarray, step: float = 1e-5, n_steps: int = 50, mass: np.ndarray = None,
                      sel: np.ndarray = None, header_cd: bool = True):
    """N2: cg_hmc_galaxy_params -- one HMC update (celeste_mcmc.py:281-323, util/infer/hmc.py with
    num_iter = 1) of every galaxy in ONE launch.  d_srcs: device CgSource bytes (updated in place where the
    proposal is accepted); d_boxes: device CgBox bytes [.. * n_img]; photons [.. * n_img, max_patch_px]
    float32 device patches shaped like the boxes; p_in [n_src, 11] the momentum after the host's refresh,
    ke_curr [n_src] sum(P P / mass) of the momentum before it, log_u [n_src].
    sel: optional indices of the n_src galaxies inside d_srcs / d_boxes / photons (a mixed source list);
    header_cd: the Cython route's header CD matrix (galaxy_source_like) or, False, cd_at_pixel (live route).
    Returns host arrays (p_out [n_src, 11], h [n_src, 2] = -H before / proposed, accepted, status)."""
    harr, n_img = _handles(images)
    dev = device()
    d_sel = None if sel is None else torch.from_numpy(np.ascontiguousarray(sel, dtype=np.int32)).to(dev)
    mass = np.ascontiguousarray(HMC_GALAXY_MASS if mass is None else mass, dtype=np.float64)
    p_in = np.ascontiguousarray(p_in, dtype=np.float64)
    if p_in.shape != (n_src, 11) or mass.shape != (11,) or np.shape(ke_curr) != (n_src,) or np.shape(log_u) != (n_src,):
        raise ValueError("p_in [n_src, 11], ke_curr [n_src], log_u [n_src], mass [11]")
    d_p = torch.from_numpy(p_in).to(dev)
    d_ke = torch.from_numpy(np.ascontiguousarray(ke_curr, dtype=np.float64)).to(dev)
    d_lu = torch.from_numpy(np.ascontiguousarray(log_u, dtype=np.float64)).to(dev)
    p_out = torch.zeros((n_src, 11), dtype=torch.float64, device=dev)
    h = torch.zeros((n_src, 2), dtype=torch.float64, device=dev)
    acc = torch.zeros(n_src, dtype=torch.int32, device=dev)
    status = torch.zeros(n_src, dtype=torch.int32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(lib().cg_hmc_galaxy_params(harr, n_img, d_srcs.data_ptr(), int(n_src),
                                     d_sel.data_ptr() if d_sel is not None else None,
                                     _ffi.CG_ITEM_HEADER_CD if header_cd else 0, d_boxes.data_ptr(),
                                     photons.data_ptr(), int(max_patch_px), d_p.data_ptr(), d_ke.data_ptr(),
                                     d_lu.data_ptr(), float(step), mass.ctypes.data, int(n_steps),
                                     p_out.data_ptr(), h.data_ptr(), acc.data_ptr(), status.data_ptr(), stream_ptr()))
    e1.record()
    out = p_out.cpu().numpy(), h.cpu().numpy(), acc.cpu().numpy(), status.cpu().numpy()
    hmc_galaxy_params.last_kernel_ms = e0.elapsed_time(e1)
    return out


def slice_sample_locations(images: Sequence[Image], d_srcs: torch.Tensor, n_src: int, d_boxes: torch.Tensor,
                           photons: torch.Tensor, max_patch_px: int, uniforms: np.ndarray, order: np.ndarray,
                           sigma: float = 1.0, guard: float = 1e-2):
    """N2: cg_slice_sample_locations -- Source.resample_location for every source in ONE launch.
    d_srcs: device CgSource bytes (updated in place for the sources that finish); d_boxes: device CgBox
    bytes [n_src * n_img]; photons: [n_src * n_img, max_patch_px] float32 device patches; uniforms
    [n_src, B] float64 host array in the sampler's draw order; order [n_src, 2] int32 coordinate order.
    Returns host arrays (ll, n_used, status, n_evals)."""
    harr, n_img = _handles(images)
    dev = device()
    uniforms = np.ascontiguousarray(uniforms, dtype=np.float64)
    if uniforms.shape[0] != n_src or order.shape != (n_src, 2):
        raise ValueError("uniforms / order must have one row per source")
    with np.errstate(divide="ignore"):
        logs = np.log(uniforms)
    d_u = torch.from_numpy(uniforms).to(dev)
    d_lu = torch.from_numpy(logs).to(dev)
    d_ord = torch.from_numpy(np.ascontiguousarray(order, dtype=np.int32)).to(dev)
    ll = torch.zeros(n_src, dtype=torch.float64, device=dev)
    used = torch.zeros(n_src, dtype=torch.int32, device=dev)
    status = torch.zeros(n_src, dtype=torch.int32, device=dev)
    evals = torch.zeros(n_src, dtype=torch.int32, device=dev)
    refined = torch.zeros(n_src, dtype=torch.int32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(lib().cg_slice_sample_locations(harr, n_img, d_srcs.data_ptr(), int(n_src), d_boxes.data_ptr(),
                                          photons.data_ptr(), int(max_patch_px), d_u.data_ptr(), d_lu.data_ptr(),
                                          d_ord.data_ptr(), uniforms.shape[1], float(sigma), float(guard),
                                          ll.data_ptr(), used.data_ptr(), status.data_ptr(), evals.data_ptr(),
                                          refined.data_ptr(), stream_ptr()))
    e1.record()
    slice_sample_locations.last_refined = refined.cpu().numpy()
    slice_sample_locations.last_kernel_ms = e0.elapsed_time(e1)
    return ll.cpu().numpy(), used.cpu().numpy(), status.cpu().numpy(), evals.cpu().numpy()
