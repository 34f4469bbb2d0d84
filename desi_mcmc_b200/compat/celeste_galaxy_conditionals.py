"""CelestePy/celeste_galaxy_conditionals.py call surface on the CUDA path."""
from __future__ import annotations

import numpy as np
import torch

from .. import _ffi, engine
from . import mixture_profiles as mp
from ._util import BANDS, full_frame, galaxy_record, render_one
from .bounding_box import calc_bounding_radius
from .mog import MixtureOfGaussians

galaxy_profs_dstn = [mp.get_exp_mixture(), mp.get_dev_mixture()]
galaxy_profs = [MixtureOfGaussians(means=g.mean, covs=g.var, pis=g.amp) for g in galaxy_profs_dstn]
galaxy_prof_dict = dict(zip(['exp', 'dev'], galaxy_profs_dstn))


def gen_galaxy_ra_dec_basis(sig_s, rho_s, phi_s):
    """:90-107 (host; four flops -- the device prologue recomputes it per item)."""
    phi = (90. - phi_s) * np.pi / 180.
    re_deg = max(1. / 30, sig_s) / 3600.
    cp, sp = np.cos(phi), np.sin(phi)
    return re_deg * np.array([[cp, sp * rho_s], [-sp, cp * rho_s]])


def gen_galaxy_transformation(sig_s, rho_s, phi_s, Ups_n):
    """:109-125."""
    G = gen_galaxy_ra_dec_basis(sig_s, rho_s, phi_s)
    T = np.dot(np.linalg.inv(G), Ups_n)
    return np.linalg.inv(T)


def gen_galaxy_psf_image(th, u_s, img, xlim=None, ylim=None, check_overlap=True,
                         unconstrained=True, return_patch=True):
    """:185-214, the live route: exp/dev mixture, affine map from
    cd_at_pixel, PSF convolution, footprint at error 1e-5 (floor/ceil), all in
    the device prologue; 42-component render."""
    patch, yl, xl = render_one(img, galaxy_record(th, u_s), _ffi.CG_ITEM_FLUX_UNIT, xlim, ylim)
    if xlim is not None and ylim is not None:
        yl, xl = ylim, xlim
    else:   # the reference returns numpy floats from np.max / np.min (:208-211)
        yl, xl = (np.float64(yl[0]), np.float64(yl[1])), (np.float64(xl[0]), np.float64(xl[1]))
    return patch, yl, xl


def gen_galaxy_psf_image_bound(src, img):
    """:217-232 (host: 42 closed-form radii; the device computes the same
    bound for its own footprints)."""
    theta_s, sig_s, phi_s, rho_s = src.shape
    px, py = img.equa2pixel(src.u)
    galmix = MixtureOfGaussians.convex_combine(galaxy_profs, [theta_s, 1. - theta_s])
    Tinv = gen_galaxy_transformation(sig_s, rho_s, phi_s, img.cd_at_pixel(px, py))
    cmix = galmix.apply_affine(Tinv, np.array([px, py])).convolve(img.psf)
    return calc_bounding_radius(cmix.pis, cmix.means, cmix.covs, error=1e-5, center=np.array([px, py]))


def gen_galaxy_prof_psf_image(prof_type, R, u, img, return_patch=True, xlim=None, ylim=None):
    """:134-182, the Cython route for one profile.  R is the matrix
    gen_galaxy_transformation returns; the device needs the shape parameters
    it came from, which are recovered from R and the image's CD matrix
    (R = Ups^-1 G, G = re [[c, s rho], [-s, c rho]])."""
    assert prof_type in galaxy_prof_dict, "unknown galaxy profile type"
    G = np.dot(img.Ups_n, np.asarray(R, dtype=np.float64))
    re_deg = np.hypot(G[0, 0], G[1, 0])
    rho = np.hypot(G[0, 1], G[1, 1]) / re_deg
    phi = 90. - np.degrees(np.arctan2(-G[1, 0], G[0, 0]))
    th = [1., re_deg * 3600., phi, rho]
    flags = _ffi.CG_ITEM_FLUX_UNIT | _ffi.CG_ITEM_TRUNC_BOX | _ffi.CG_ITEM_HEADER_CD | \
        (_ffi.CG_ITEM_PROF_EXP if prof_type == 'exp' else _ffi.CG_ITEM_PROF_DEV)
    patch, yl, xl = render_one(img, galaxy_record(th, u), flags, xlim, ylim)
    if return_patch:
        return patch, yl, xl
    return full_frame(img, patch, yl, xl)


def _cond_items(th, images, n_variants):
    """CgSource records + items for galaxy_source_like of n_variants parameter
    vectors (rows of th) over all images: the joint exp+dev mixture with the
    truncated Cython-route footprint and the header CD matrix."""
    th = np.atleast_2d(np.asarray(th, dtype=np.float64))
    rec = np.zeros(th.shape[0], dtype=_ffi.SRC_DTYPE)
    rec["type"] = 1
    rec["theta"], rec["sigma"], rec["phi"], rec["rho"] = th[:, 0], th[:, 1], th[:, 2], th[:, 3]
    rec["u"] = th[:, 4:6]
    for i, img in enumerate(images):
        rec["flux"][:, img.band_index] = th[:, 6 + BANDS.index(img.band)]
    return rec


def galaxy_source_like_batch(ths, Z_s, images, use_fp64=False):
    """galaxy_source_like (:15-42) for every row of ths in ONE launch: the
    conditional Poisson log-likelihood sum_n sum_m z log(lam) - lam with
    lam = flux_n (theta f_exp + (1-theta) f_dev) on the truncated footprint
    of the Cython route.  Z_s[n] must be shaped like that footprint.
    use_fp64: float64 tables / exp / log on the device -- needed when the values
    are differenced at delta 1e-5 (galaxy_source_like_grad)."""
    ths = np.atleast_2d(np.asarray(ths, dtype=np.float64))
    P, n_img = ths.shape[0], len(images)
    rec = _cond_items(ths, images, P)
    flags = _ffi.CG_ITEM_TRUNC_BOX | _ffi.CG_ITEM_HEADER_CD
    src_idx = np.repeat(np.arange(P), n_img)
    img_idx = np.tile(np.arange(n_img), P)
    items = engine.make_items(src_idx, img_idx, flags=flags, aux=img_idx)
    zs = [np.ascontiguousarray(z, dtype=np.float32).ravel() for z in Z_s]
    offs = np.zeros(n_img, dtype=np.int64)
    offs[1:] = np.cumsum([z.size for z in zs])[:-1]
    aux = torch.from_numpy(np.concatenate(zs)).to(engine.device())
    ll, boxes = engine.items_loglik(images, rec, items, mode=_ffi.CG_LL_CONDITIONAL_SUM, aux=aux,
                                    aux_offsets=offs, use_fp64=use_fp64)
    # EVERY variant's footprint must have the shape of its photon patch: a finite-difference variant
    # whose int() truncation lands one pixel off would index the patch with the wrong row stride
    # (the reference raises a shape mismatch in `Z_s[n] * np.log(lam)`, :41)
    for p in range(P):
        for n in range(n_img):
            b = boxes[p * n_img + n]
            shp = (int(b["y1"] - b["y0"]), int(b["x1"] - b["x0"]))
            ok = (shp == tuple(np.shape(Z_s[n]))) if np.ndim(Z_s[n]) == 2 else (shp[0] * shp[1] == np.size(Z_s[n]))
            if not ok:
                raise ValueError("operands could not be broadcast together: Z_s[%d] has shape %s, the footprint of "
                                 "parameter vector %d on image %d is %s" %
                                 (n, tuple(np.shape(Z_s[n])), p, n, (int(b["y1"] - b["y0"]), int(b["x1"] - b["x0"]))))
    return ll.cpu().numpy().reshape(P, n_img).sum(axis=1)


def galaxy_source_like(th, Z_s, images, check_overlap=True, unconstrained=True):
    """:15-42.  th = [theta, sigma, phi, rho, u0, u1, b_u, b_g, b_r, b_i, b_z]."""
    return float(galaxy_source_like_batch(np.asarray(th)[None, :], Z_s, images)[0])


def galaxy_source_like_grad(th, Z_s, images, check_overlap=True, unconstrained=False):
    """:44-88.  The reference differentiates theta and the fluxes analytically
    and sigma, phi, rho, u by central differences (delta 1e-5): ten extra
    likelihood evaluations.  Here all eleven parameter vectors (and the two
    single-profile renders the analytic part needs) go to the device in one
    batch."""
    th = np.asarray(th, dtype=np.float64)
    theta_s, sig_s, phi_s, rho_s = th[0:4]
    u_s = th[4:6]
    bs = dict(zip(BANDS, th[-5:]))
    grad_theta_s = 0.
    grad_bs = dict(zip(BANDS, np.zeros(len(BANDS))))
    for n, img in enumerate(images):
        R_s = gen_galaxy_transformation(sig_s, rho_s, phi_s, img.Ups_n)
        f_exp, _, _ = gen_galaxy_prof_psf_image('exp', R_s, u_s, img)
        f_dev, _, _ = gen_galaxy_prof_psf_image('dev', R_s, u_s, img)
        f_nms = theta_s * f_exp + (1. - theta_s) * f_dev
        grad_theta_s += np.sum((np.asarray(Z_s[n]) / f_nms - bs[img.band]) * (f_exp - f_dev))
        grad_bs[img.band] += 1. / bs[img.band] * np.sum(Z_s[n]) - np.sum(f_nms)
    variants = []
    for th_i in [1, 2, 3, 4, 5]:
        de = np.zeros(th.shape)
        de[th_i] = 1e-5
        variants += [th + de, th - de]
    # float64 on the device: an fp32 likelihood of ~1e5 carries ~1e-2 of rounding noise,
    # i.e. +-500 after division by 2e-5
    lls = galaxy_source_like_batch(np.array(variants), Z_s, images, use_fp64=True)
    grad_RU = (lls[0::2] - lls[1::2]) / (2. * 1e-5)
    return np.concatenate([[grad_theta_s], grad_RU, np.array([grad_bs[b] for b in BANDS])])


def gen_galaxy_prof_psf_image_bound(prof_type, R, u, img, ERROR=.01):
    """:236-256: bounding radius of one profile convolved with the image PSF."""
    from . import celeste_fast, mixture_profiles
    from .bounding_box import calc_bounding_radius
    prof = {'exp': mixture_profiles.get_exp_mixture(), 'dev': mixture_profiles.get_dev_mixture()}
    assert prof_type in prof, "unknown galaxy profile type"
    v_s = np.asarray(img.equa2pixel(u), dtype=np.float64)
    R = np.asarray(R, dtype=np.float64)
    weights, means, covars = celeste_fast.gen_galaxy_prof_psf_mixture_params(
        W=np.dot(R, R.T), v_s=v_s, image_ws=img.weights, image_means=img.means,
        image_covars=img.covars, gal_prof_amp=prof[prof_type].amp,
        gal_prof_sigs=prof[prof_type].var[:, 0, 0])
    return calc_bounding_radius(weights, means, covars, ERROR, center=v_s)


# ---- drop-in prior probability functions (:262-275, util/like/like_list.py:18-27) ----
def fast_inv_gamma_lnpdf(x, a0, b0):
    if x <= 0.:
        return -np.inf
    return (-a0 - 1.) * np.log(x) - b0 / x


def galaxy_shape_prior_unconstrained(logit_theta, log_sig, logit_phi, logit_rho):
    return -(1. / 50.) * (logit_theta * logit_theta + log_sig * log_sig +
                          logit_phi * logit_phi + logit_rho * logit_rho)


def galaxy_shape_prior_constrained(theta, sig, phi, rho):
    if theta <= 0. or theta >= 1. or sig <= 0. or phi <= 0. or phi >= np.pi or rho <= 0. or rho >= 1.:
        return -np.inf
    return fast_inv_gamma_lnpdf(sig * sig, a0=1., b0=1.)


def constrain_params(th):
    """:288-299."""
    return 1. / (1. + np.exp(-th[0])), np.exp(th[1]), np.pi / (1. + np.exp(-th[2])), 1. / (1. + np.exp(-th[3]))


def unconstrain_params(th_constrained):
    """:301-313."""
    return np.log(th_constrained[0] / (1. - th_constrained[0])), np.log(th_constrained[1]), \
        np.log(th_constrained[2] / (np.pi - th_constrained[2])), \
        np.log(th_constrained[3] / (1. - th_constrained[3]))
