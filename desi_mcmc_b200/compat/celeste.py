"""CelestePy/celeste.py call surface on the CUDA path.  `image` arguments are
desi_mcmc_b200.engine.Image objects (the FitsImage mirror); `src` arguments are
anything with SrcParams's fields."""
from __future__ import annotations

import numpy as np

from .. import _ffi, engine
from . import celeste_galaxy_conditionals as gal_funs
from ._util import BANDS as _B, fluxes_of, full_frame, record_of, render_one, star_record

BANDS = np.array(['u', 'g', 'r', 'i', 'z'], dtype=object)


def _off_image(u, image):
    """celeste.py:130-132 (v_s[0] is compared against both dimensions, as written)."""
    v_s = image.equa2pixel(u)
    return (v_s[0] < -50 or v_s[0] > 2 * image.shape[0] or
            v_s[1] < -50 or v_s[0] > 2 * image.shape[1])


def gen_point_source_psf_image(u, image, xlim=None, ylim=None, check_overlap=True,
                               return_patch=True, psf_grid=None, pixel_grid=None):
    """:114-176 -> (patch, ylim, xlim); (None, None, None) off the image.
    pixel_grid is accepted for signature parity; the kernel generates the
    coordinates of (xlim, ylim) itself."""
    if check_overlap and _off_image(u, image):
        return None, None, None
    patch, yl, xl = render_one(image, star_record(u), _ffi.CG_ITEM_FLUX_UNIT, xlim, ylim)
    if xlim is not None and ylim is not None:
        yl, xl = ylim, xlim
    if return_patch:
        return patch, yl, xl
    return full_frame(image, patch, (int(yl[0]), int(yl[1])), (int(xl[0]), int(xl[1])), psf_grid)


def gen_point_source_psf_image_with_fluxes(src_params, fits_image, return_patch=True, psf_grid=None):
    """:72-82."""
    if _off_image(src_params.u, fits_image):
        raise TypeError("source does not overlap the image")     # the reference fails on None *= flux
    rec = star_record(src_params.u, fluxes_of(src_params))
    return render_one(fits_image, rec, _ffi.CG_ITEM_FLUX_BAND)


def gen_galaxy_psf_image(src, image, return_patch=True, check_overlap=True):
    """:99-111."""
    assert src.a == 1, "generating glaxay psf image for non galaxy."
    th = np.array([src.theta, src.sigma, src.phi, src.rho])
    return gal_funs.gen_galaxy_psf_image(th, src.u, image, check_overlap=check_overlap,
                                         unconstrained=False, return_patch=return_patch)


def gen_src_image_with_fluxes(src, img):
    """:84-96 -> (rate patch, ylim, xlim)."""
    if src.a == 0:
        return gen_point_source_psf_image_with_fluxes(src, img)
    patch, yl, xl = render_one(img, record_of(_flux_only(src)), _ffi.CG_ITEM_FLUX_BAND)
    return patch, (np.float64(yl[0]), np.float64(yl[1])), (np.float64(xl[0]), np.float64(xl[1]))


class _flux_only(object):
    """view of a source that hides its temperature (flux route only)."""

    def __init__(self, src):
        self.__dict__.update(u=src.u, a=src.a, b=None, t=None, theta=src.theta, phi=src.phi,
                             sigma=src.sigma, rho=src.rho, fluxes=src.fluxes)


def gen_src_image(src, image, return_patch=True):
    """:27-62: expected-photon image of one source (black-body stars included)."""
    if src.a == 0:
        if _off_image(src.u, image):
            raise TypeError("source does not overlap the image")
        patch, yl, xl = render_one(image, record_of(src), _ffi.CG_ITEM_FLUX_BAND)
    elif src.a == 1:
        patch, yl, xl = render_one(image, record_of(src), _ffi.CG_ITEM_FLUX_BAND)
    elif src.a is None and src.fluxes is not None:
        # :50-53: expected_photons = kappa * flux (no calibration)
        patch, yl, xl = render_one(image, star_record(src.u), _ffi.CG_ITEM_FLUX_UNIT)
        patch = patch * (image.kappa * fluxes_of(src)[image.band_index])
    else:
        raise Exception("No way to compute expected photons without at least fluxes or brightness")
    if return_patch:
        return patch
    return full_frame(image, patch, yl, xl)[0]


def gen_src_psf_image(src, image):
    if src.a == 0:
        return gen_point_source_psf_image(src.u, image)
    elif src.a == 1:
        return gen_galaxy_psf_image(src, image)
    raise NotImplementedError("not implemented!")


def gen_psf_src_image_bound(src, img):
    """:192-199."""
    if src.a == 0:
        return img.R
    return gal_funs.gen_galaxy_psf_image_bound(src, img)


def gen_model_image(srcs, image):
    """:203-219: lambda = epsilon + sum_s patch_s on the full frame (one
    prepare / render / gather pass on the device)."""
    _, models = engine.field_loglik([image], engine.pack_sources(list(srcs)) if len(srcs) else
                                    np.zeros(0, dtype=_ffi.SRC_DTYPE), return_model=True)
    return models[0].cpu().numpy().astype(np.float64)


def gen_src_prob_layers(srcs, img):
    """:222-234: responsibility of the sky (layer 0) and of each source."""
    H, W = img.shape
    layers = np.zeros((len(srcs) + 1, H, W))
    for s, src in enumerate(srcs):
        layers[s + 1] = gen_src_image(src, img, False)
    total = layers[1:].sum(axis=0) + img.epsilon
    layers[0] = img.epsilon
    return layers / total


def celeste_likelihood(srcs, image):
    """:237-240."""
    return float(celeste_likelihood_per_image(srcs, [image])[0])


def celeste_likelihood_per_image(srcs, images):
    arr = engine.pack_sources(list(srcs)) if len(srcs) else np.zeros(0, dtype=_ffi.SRC_DTYPE)
    return engine.field_loglik(list(images), arr).cpu().numpy()


def celeste_likelihood_multi_image(srcs, images):
    """:243-252: all images in one device pass."""
    return float(np.sum(celeste_likelihood_per_image(srcs, images)))
