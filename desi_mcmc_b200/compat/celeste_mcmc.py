"""CelestePy/celeste_mcmc.py call surface on the CUDA path.

Kept: the functions the live Gibbs sweep reaches --
``sample_source_photons_single_image_cython`` (:98-150), ``temp_bright_like``
(:374-390), ``loc_like`` (:393-404), ``unif_lnpdf`` (:406-409),
``sample_star_params`` (:334-371) -- plus a batched form of the photon
allocation for all images of a field (one render + one allocation launch for
the five bands instead of five Python passes).

``celeste_gibbs_sample`` (:13-96) references undefined names in the reference
(``pixel_loc``, ``get_bounding_boxes_idx``) and cannot run there either; the
live sweep is ``models.CelesteBase.resample_model``.  Of ``sample_galaxy_params``
(:167-331) the working parts are provided as functions of their own:
``sample_galaxy_fluxes`` (gamma conditional), ``slice_sample_skew`` and
``hmc_sample_galaxy_params`` (util/infer/hmc.py leapfrog over the device
likelihood and its batched finite-difference gradient).
"""
from __future__ import annotations

import numpy as np

from .. import _ffi, engine
from . import celeste, planck
from . import celeste_galaxy_conditionals as gal
from . import celeste_sample_sources as css
from ._util import BANDS, fluxes_of, record_of
from .hmc import hmc
from .slicesample import slicesample, slicesample_step_out


def _seed(random_state=None):
    return css._seed_from(random_state)


def sample_source_photons_field(imgs, srcs, random_state=None, bound_cap=0.0):
    """All images of a field at once.  Returns (samp_imgs[img][src], noise_sums[img])
    with the same per-image content as sample_source_photons_single_image_cython."""
    packed = np.concatenate([record_of(s) for s in srcs]) if len(srcs) else \
        np.zeros(0, dtype=_ffi.SRC_DTYPE)
    alloc = engine.resample_photons(imgs, packed, seed=_seed(random_state), bound_cap=bound_cap)
    n_img = len(imgs)
    out = []
    for i in range(n_img):
        per_src = []
        for s in range(len(srcs)):
            b = alloc.boxes[s * n_img + i]
            if b["x1"] <= b["x0"] or b["y1"] <= b["y0"]:
                per_src.append(None)            # gen_src_image_with_fluxes returned (None, None, None)
                continue
            per_src.append(css.NativePatch(alloc.patch(s, i, "samp"),
                                           (int(b["y0"]), int(b["y1"])), (int(b["x0"]), int(b["x1"]))))
        out.append(per_src)
    return out, [float(v) for v in alloc.noise_sums]


def sample_source_photons_single_image_cython(img, srcs, random_state=None):
    """:98-150.  Render every source's rate patch on `img`, then split each
    pixel's photons over the sources whose patch box contains it and the sky
    (celeste_sample_sources.pyx:61-122).  Returns (samp_imgs, noise_sum)."""
    samp, noise = sample_source_photons_field([img], srcs, random_state)
    return samp[0], noise[0]


def unif_lnpdf(x, a0, b0):
    """:406-409."""
    if x <= a0 or x >= b0:
        return -np.inf
    return 0.


def temp_bright_like(th, fs_sum, src_image_sums, imgs):
    """:374-390: black-body temperature / brightness conditional of a star.
    The five band integrals are one device call (planck.py:155 per band in the
    reference)."""
    ll = 0
    exp_b = planck.photons_expected_brightness_batch(np.array([th[0]]), np.array([th[1]]))[0]
    for i, img in enumerate(imgs):
        expected_num_photons = fs_sum[i] * exp_b[BANDS.index(img.band)]
        if expected_num_photons > 0:
            ll += np.sum(src_image_sums[i]) * np.log(expected_num_photons) - expected_num_photons
    ll += unif_lnpdf(th[0], 20., 20000.)
    ll += unif_lnpdf(th[1], 0., 1.)
    return ll


def loc_like(u, src, imgs, src_images):
    """:393-404: location conditional given the source's photon images.  All
    images the source reaches are rendered and reduced in one launch."""
    src.u = u
    use = [n for n, img in enumerate(imgs) if img.contains(src.u)]
    if not use:
        return 0
    rec = record_of(src)
    zs, boxes, ims = [], [], []
    for n in use:
        z = np.asarray(src_images[n].data if hasattr(src_images[n], "data") else src_images[n])
        p = src_images[n]
        if hasattr(p, "x0"):
            boxes.append((p.x0, p.x1, p.y0, p.y1))
        else:
            boxes.append((0, z.shape[1], 0, z.shape[0]))
        zs.append(np.ascontiguousarray(z, dtype=np.float32).ravel())
        ims.append(imgs[n])
    offs = np.zeros(len(zs), dtype=np.int64)
    offs[1:] = np.cumsum([z.size for z in zs])[:-1]
    import torch
    aux = torch.from_numpy(np.concatenate(zs)).to(engine.device())
    k = len(ims)
    items = engine.make_items([0] * k, np.arange(k), box=boxes, aux=np.arange(k))
    # sum log(f)*z over {z > 0, f > 0} - sum f   (:401-403) = CG_LL_CONDITIONAL_SUM
    ll, _ = engine.items_loglik(ims, rec, items, mode=_ffi.CG_LL_CONDITIONAL_SUM, aux=aux, aux_offsets=offs)
    return float(ll.sum().item())


def sample_star_params(src, src_images, imgs, subiter=2, verbose=False):
    """:334-371: slice-sample (T, b) then the location of a star given its photon images."""
    for gibbs_iter in range(subiter):
        sum_fs = np.zeros(len(imgs))
        for n, img in enumerate(imgs):
            patch, _, _ = celeste.gen_point_source_psf_image(src.u, img)
            sum_fs[n] = min(1., np.sum(patch)) if patch is not None else 0.
        src_image_sums = np.array([np.sum(np.asarray(getattr(im, "data", im))) for im in src_images])
        th = np.array([src.t, src.b], dtype=np.float64)
        # the reference's keyword set (xx=, llh_func=, step=, x_l=) is that of
        # slicesample_step_out (util/infer/slicesample.py:4)
        for it in range(2):
            th, ll = slicesample_step_out(
                xx=th, llh_func=lambda t: temp_bright_like(t, sum_fs, src_image_sums, imgs),
                step=[1000, .1], step_out=True, x_l=[0., 0.])
        src.t, src.b = th[0], th[1]
        u, ll = slicesample_step_out(xx=np.asarray(src.u, dtype=np.float64),
                                     llh_func=lambda u: loc_like(u, src, imgs, src_images),
                                     step=[.1, .1], step_out=False)
        src.u = u


def sample_galaxy_fluxes(src, src_imgs, imgs, random_state=None):
    """:172-209: conjugate gamma update of the five fluxes of a galaxy."""
    rng = random_state or np.random
    a_0, b_0 = 5., .005
    band_counts, psf_sums = np.zeros(5), np.zeros(5)
    th = (src.theta, src.sigma, src.phi, src.rho)
    for n, img in enumerate(imgs):
        bi = BANDS.index(img.band)
        band_counts[bi] += np.sum(np.asarray(getattr(src_imgs[n], "data", src_imgs[n])))
        psf_ns, _, _ = gal.gen_galaxy_psf_image(th, src.u, img)
        if psf_ns is not None:
            psf_sums[bi] += psf_ns.sum() * img.kappa / img.calib
    fluxes = rng.gamma(a_0 + band_counts, 1. / (b_0 + psf_sums))
    src.fluxes = fluxes
    return fluxes


def slice_sample_skew(src, src_imgs, imgs, logprior=None):
    """:211-246: slice sampler over (theta, sigma, phi, rho) given the photon
    images, location and fluxes; every evaluation is one batched device call
    (galaxy_source_like over the five bands)."""
    f = fluxes_of(src)

    def skew_likelihood(th):
        if not (0. < th[0] < 1. and th[1] > 0. and 0. < th[3] <= 1.):
            return -np.inf
        lp = 0. if logprior is None else logprior(th)
        if not np.isfinite(lp):
            return -np.inf
        full = np.concatenate([th, np.asarray(src.u, dtype=np.float64), f])
        return gal.galaxy_source_like(full, src_imgs, imgs, unconstrained=False) + lp

    th_curr = np.array([src.theta, src.sigma, src.phi, src.rho], dtype=np.float64)
    th_curr, llh = slicesample(th_curr, skew_likelihood, step_out=True, doubling_step=True,
                               numdir=len(th_curr), compwise=False)
    th_curr[2] = (th_curr[2] + np.pi) % np.pi
    src.theta, src.sigma, src.phi, src.rho = th_curr
    return th_curr


def hmc_sample_galaxy_params(src, src_imgs, imgs, p_curr=None, step_sz=1e-5, steps_per_sample=50):
    """:273-319: HMC over (theta, sigma, phi, rho, u, fluxes) given the photon
    images.  Each leapfrog gradient is galaxy_source_like_grad: its ten
    finite-difference likelihoods x five bands are ONE device batch.  Z_s must
    be shaped like the Cython-route footprints (galaxy_source_like's contract).
    Returns the new momentum (the reference keeps it in the global p_cached)."""
    ll_fun = lambda th: gal.galaxy_source_like(th, src_imgs, imgs, check_overlap=True, unconstrained=True)
    ll_jac = lambda th: gal.galaxy_source_like_grad(th, src_imgs, imgs)
    th = np.concatenate(([src.theta, src.sigma, src.phi, src.rho], np.asarray(src.u, dtype=np.float64),
                         fluxes_of(src)))
    mass = np.array([.1, .1, 1., .1, .001, .001, 1., 1., 1., 1., 1.])
    if p_curr is None:
        p_curr = np.sqrt(mass) * np.random.randn(th.shape[0])
    th_samp, p_samp, step_sz, _ = hmc(x_curr=th, llhfunc=ll_fun, grad_llhfunc=ll_jac, eps=step_sz,
                                      num_steps=steps_per_sample, mass=mass, p_curr=p_curr,
                                      refresh_alpha=.9, adaptive_step_sz=False)
    src.theta, src.sigma, src.phi, src.rho = th_samp[0:4]
    src.u = th_samp[4:6]
    src.fluxes = dict(zip(BANDS, th_samp[6:]))
    return p_samp


def hmc_sample_galaxy_params_on_device(srcs, src_imgs_list, imgs, p_curr=None, step_sz=1e-5,
                                       steps_per_sample=50, random_state=None):
    """hmc_sample_galaxy_params (:281-323) for EVERY galaxy of `srcs` in ONE launch
    (cg_hmc_galaxy_params): the 50 leapfrog steps and their 51 x 11 + 2 float64
    likelihoods run on the device, one thread block per galaxy.  src_imgs_list[s][n] is galaxy s's photon
    image on imgs[n], shaped like the Cython-route footprint at the current parameters
    (galaxy_source_like's contract; checked).  Randomness comes from `random_state` (default numpy.random)
    in the order of the per-galaxy host loop: [the initial momentum if p_curr is None,] the refresh
    normals, the acceptance uniform, galaxy after galaxy.  Where the reference would re-derive the
    footprint at every evaluation (and raise once its shape differs from Z_s) the device keeps the
    footprint of the start point.  Updates the sources in place; returns (p_samp [n, 11], accepted [n])."""
    import torch
    rs = np.random if random_state is None else random_state
    srcs = list(srcs)
    S, n_img = len(srcs), len(imgs)
    if S == 0:
        return np.zeros((0, 11)), np.zeros(0, dtype=bool)
    mass = engine.HMC_GALAXY_MASS
    th = np.array([np.concatenate(([s.theta, s.sigma, s.phi, s.rho], np.asarray(s.u, dtype=np.float64), fluxes_of(s)))
                   for s in srcs])
    rec = gal._cond_items(th, imgs, S)
    flags = _ffi.CG_ITEM_TRUNC_BOX | _ffi.CG_ITEM_HEADER_CD
    src_idx, img_idx = np.repeat(np.arange(S), n_img), np.tile(np.arange(n_img), S)
    zmax = max(int(np.size(z)) for zs in src_imgs_list for z in zs)
    _, boxes = engine.render_patches(imgs, rec, engine.make_items(src_idx, img_idx, flags=flags | _ffi.CG_ITEM_FLUX_UNIT),
                                     max_patch_px=max(zmax, 1))     # the footprints at the start points
    stride = 1
    for k in range(S * n_img):
        b = boxes[k]
        shp = (int(b["y1"] - b["y0"]), int(b["x1"] - b["x0"]))
        z = np.asarray(src_imgs_list[k // n_img][k % n_img])
        if z.shape != shp:
            raise ValueError("operands could not be broadcast together: Z_s[%d] of galaxy %d has shape %s, its "
                             "footprint is %s" % (k % n_img, k // n_img, z.shape, shp))
        stride = max(stride, shp[0] * shp[1])
    host = np.zeros((S * n_img, stride), dtype=np.float32)
    for k in range(S * n_img):
        z = np.asarray(src_imgs_list[k // n_img][k % n_img], dtype=np.float32).ravel()
        host[k, :z.size] = z
    photons = torch.from_numpy(host).to(engine.device())
    dbox = np.zeros(S * n_img, dtype=_ffi.BOX_DTYPE)
    for k in range(S * n_img):
        dbox[k] = (int(boxes[k]["x0"]), int(boxes[k]["x1"]), int(boxes[k]["y0"]), int(boxes[k]["y1"]))
    # the host generator, in the order of hmc.py:62,67,73 for one galaxy after the other
    P_old, P_new, log_u = np.zeros((S, 11)), np.zeros((S, 11)), np.zeros(S)
    refresh_alpha = .9
    for s in range(S):
        P_old[s] = np.sqrt(mass) * rs.randn(11) if p_curr is None else np.asarray(p_curr[s], dtype=np.float64)
        P_new[s] = refresh_alpha * P_old[s] + np.sqrt(1.0 - refresh_alpha ** 2) * np.sqrt(mass) * rs.randn(11)
        log_u[s] = np.log(rs.rand())
    ke = np.array([np.sum((1. / mass) * P_old[s] * P_old[s]) for s in range(S)])
    d_srcs = engine.to_device(rec)
    p_out, h, acc, status = engine.hmc_galaxy_params(imgs, d_srcs, S, engine.to_device(dbox), photons, stride,
                                                     P_new, ke, log_u, step=step_sz, n_steps=steps_per_sample)
    if np.any(status != 0):
        raise FloatingPointError("non-finite energy on the HMC trajectory of galaxies %s" % np.nonzero(status)[0].tolist())
    new = d_srcs.cpu().numpy().view(_ffi.SRC_DTYPE)
    for s, src in enumerate(srcs):
        src.theta, src.sigma, src.phi, src.rho = (float(new["theta"][s]), float(new["sigma"][s]),
                                                  float(new["phi"][s]), float(new["rho"][s]))
        src.u = np.array(new["u"][s], dtype=np.float64)
        src.fluxes = dict(zip(BANDS, [float(v) for v in new["flux"][s]]))
    hmc_sample_galaxy_params_on_device.last = dict(h=h, kernel_ms=engine.hmc_galaxy_params.last_kernel_ms)
    return p_out, acc.astype(bool)


def sample_source_params(src, src_imgs, imgs, verbose=False):
    """:153-164."""
    if src.a == 0:
        sample_star_params(src, src_imgs, imgs, subiter=2, verbose=verbose)
    elif src.a == 1:
        for _ in range(2):
            sample_galaxy_fluxes(src, src_imgs, imgs)
            slice_sample_skew(src, src_imgs, imgs)
    else:
        raise Exception("Must be star or galaxy to sample")


def printif(statement, condition):
    if condition:
        print(statement)
