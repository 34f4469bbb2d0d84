"""CPU ORACLE -- TEST INFRASTRUCTURE ONLY.

Python-3 / NumPy float64 restatement of the reference's (HIPS/DESI-MCMC
"CelestePy") render + Poisson log-likelihood path.  The reference's own Python
cannot be imported here (Python-2 syntax, autograd/fitsio/tractor imports), so
each function below restates one reference function and cites the file:line it
follows (paths relative to /root/reference).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline
legs may import this module.  The product package (``desi_mcmc_b200``) never
does; it fails loudly when its CUDA library is missing.

Pinning status
  * ``gmm_like_2d`` / ``mog_loglike``: pinned by the reference's own KAT
    (CelestePy/test/test_gmm.py:63-105) and by the compiled reference
    (``oracle/_ref``, built by ``oracle/build_ref.py``), see
    tests/test_oracle.py.
  * ``gen_galaxy_prof_psf_mixture_params``: pinned against the compiled
    reference ``celeste_fast`` module.
  * every other function: the reference ships no test or golden vector for it
    ("parity unpinned" by the reference's own tests); the
    restatement is cross-checked where two reference routes compute the same
    quantity (Cython route vs live MoG route).
"""
from __future__ import annotations

import math
import os

import numpy as np
from scipy.special import logsumexp

BANDS = ["u", "g", "r", "i", "z"]

# --------------------------------------------------------------------------
# mixture_profiles.py:9-19  (Hogg & Lang exp / deV mixture-of-Gaussian fits)
# --------------------------------------------------------------------------
exp_amp = np.array([2.34853813e-03, 3.07995260e-02, 2.23364214e-01,
                    1.17949102e+00, 4.33873750e+00, 5.99820770e+00])
exp_var = np.array([1.20078965e-03, 8.84526493e-03, 3.91463084e-02,
                    1.39976817e-01, 4.60962500e-01, 1.50159566e+00])
exp_amp = exp_amp / np.sum(exp_amp)

dev_amp = np.array([4.26347652e-02, 2.40127183e-01, 6.85907632e-01, 1.51937350e+00,
                    2.83627243e+00, 4.46467501e+00, 5.72440830e+00, 5.60989349e+00])
dev_var = np.array([2.23759216e-04, 1.00220099e-03, 4.18731126e-03, 1.69432589e-02,
                    6.84850479e-02, 2.87207080e-01, 1.33320254e+00, 8.40215071e+00])
dev_amp = dev_amp / np.sum(dev_amp)


# --------------------------------------------------------------------------
# util/dists/mog.py
# --------------------------------------------------------------------------
def mog_loglike(x, means, icovs, dets, pis):
    """util/dists/mog.py:5-21."""
    xx = np.atleast_2d(x)
    centered = xx[:, :, np.newaxis] - means.T[np.newaxis, :, :]
    solved = np.einsum('ijk,lji->lki', icovs, centered)
    logprobs = (-0.5 * np.sum(solved * centered, axis=1) - np.log(2 * np.pi)
                - 0.5 * np.log(dets) + np.log(pis))
    logprob = logsumexp(logprobs, axis=1)
    if len(np.shape(x)) == 1:
        return logprob[0]
    return logprob


class MixtureOfGaussians(object):
    """util/dists/mog.py:38-112."""

    def __init__(self, means, covs, pis):
        self.K, self.D = means.shape
        self.update_params(means, covs, pis)

    def update_params(self, means, covs, pis):          # mog.py:48-57
        assert covs.shape[1] == covs.shape[2] == self.D
        assert self.K == covs.shape[0] == len(pis)
        self.means = means
        self.covs = covs
        self.pis = pis
        self.dets = np.array([np.linalg.det(c) for c in self.covs])
        self.icovs = np.array([np.linalg.inv(c) for c in self.covs])

    def logpdf(self, x):                                # mog.py:59-61
        return mog_loglike(x, means=self.means, icovs=self.icovs,
                           dets=self.dets, pis=self.pis)

    def convolve(self, mog):                            # mog.py:75-81
        means = np.reshape(self.means[:, None] + mog.means[None, :], (-1, 2))
        weights = np.reshape(self.pis[:, None] * mog.pis[None, :], (-1,))
        covs = np.reshape(self.covs[:, None] + mog.covs[None, :], (-1, 2, 2))
        return MixtureOfGaussians(means, covs, weights)

    def apply_affine(self, A, b):                       # mog.py:84-92
        return MixtureOfGaussians(
            means=np.dot(self.means, A.T) + b,
            covs=np.array([np.dot(np.dot(A, c), A.T) for c in self.covs]),
            pis=self.pis)

    @staticmethod
    def convex_combine(mogs, mixing_weights):           # mog.py:94-100
        return MixtureOfGaussians(
            means=np.vstack([m.means for m in mogs]),
            covs=np.vstack([m.covs for m in mogs]),
            pis=np.concatenate([w * m.pis for w, m in zip(mixing_weights, mogs)]))

    def evaluate_grid(self, xlim, ylim):                # mog.py:102-112
        assert (ylim[1] > ylim[0]) and (xlim[1] > xlim[0]), "bad limits."
        y_grid = np.arange(ylim[0], ylim[1], dtype=np.float64)
        x_grid = np.arange(xlim[0], xlim[1], dtype=np.float64)
        xx, yy = np.meshgrid(x_grid, y_grid, indexing='xy')
        pts = np.column_stack((xx.ravel(order='C'), yy.ravel(order='C')))
        lls = mog_loglike(pts, self.means, self.icovs, self.dets, self.pis)
        return np.reshape(np.exp(lls), xx.shape)


def _profile_mog(amp, var):
    """celeste_galaxy_conditionals.py:128-131 (profile objects: zero means,
    isotropic var*I covariances, mixture_profiles.py:21-25,44-53)."""
    K = amp.size
    covs = np.zeros((K, 2, 2))
    covs[:, 0, 0] = var
    covs[:, 1, 1] = var
    return MixtureOfGaussians(means=np.zeros((K, 2)), covs=covs, pis=amp.copy())


galaxy_profs = [_profile_mog(exp_amp, exp_var), _profile_mog(dev_amp, dev_var)]
galaxy_prof_dict = {"exp": (exp_amp, exp_var), "dev": (dev_amp, dev_var)}


# --------------------------------------------------------------------------
# util/bound/bounding_box.py
# --------------------------------------------------------------------------
CHI2_PPF = {0.001: 13.815510557964274, 1e-5: 23.025850929940457}


def chi2_ppf_2dof(p):
    """scipy.stats.chi2.ppf(p, 2) == -2 log(1-p) (closed form for 2 dof)."""
    return -2.0 * math.log1p(-p)


def calc_bounding_radius(weights, means, covars, error, center=np.array([0, 0])):
    """util/bound/bounding_box.py:9-31."""
    import scipy.stats
    minbound = -np.inf
    Rsq = scipy.stats.chi2.ppf(1 - error, 2)
    for i in range(len(weights)):
        sigma1 = np.sqrt(covars[i, 0, 0])
        sigma2 = np.sqrt(covars[i, 1, 1])
        rho = covars[i, 0, 1] / (sigma1 * sigma2)
        A11 = sigma1
        A21 = rho * sigma2
        A22 = sigma2 * np.sqrt(1 - rho ** 2.)
        An = 1. / Rsq * (1 / A11 ** 2. + A21 ** 2. / A22 ** 2.)
        Bn = 1. / Rsq * (-2 * A21 / (A11 * A22 ** 2.))
        Cn = 1. / Rsq * 1. / A22 ** 2.
        majaxis = (0.5 * (An + Cn - np.sqrt(Bn ** 2. + (An - Cn) ** 2.))) ** (-0.5)
        dist = np.sqrt((means[i, 0] - center[0]) ** 2. + (means[i, 1] - center[1]) ** 2.)
        minbound = max(minbound, majaxis + dist)
    return minbound


# --------------------------------------------------------------------------
# celeste_fast.pyx / gmm_like_fast.pyx restated (A1 / A2)
# --------------------------------------------------------------------------
def gen_galaxy_prof_psf_mixture_params(W, v_s, image_ws, image_means, image_covars,
                                       gal_prof_amp, gal_prof_sigs):
    """celeste_fast.pyx:100-140 (PSF-major ordering c = k*K_prof + j)."""
    K_psf, K_prof = image_ws.shape[0], gal_prof_amp.shape[0]
    weights = np.zeros(K_psf * K_prof)
    means = np.zeros((K_psf * K_prof, 2))
    covars = np.zeros((K_psf * K_prof, 2, 2))
    cnt = 0
    for k in range(K_psf):
        for j in range(K_prof):
            weights[cnt] = image_ws[k] * gal_prof_amp[j]
            means[cnt, 0] = v_s[0] + image_means[k, 0]
            means[cnt, 1] = v_s[1] + image_means[k, 1]
            covars[cnt] = image_covars[k] + gal_prof_sigs[j] * W
            cnt += 1
    return weights, means, covars


def gen_galaxy_psf_mixture_params(thetas, W, v_s, image_ws, image_means, image_covars,
                                  gal_exp_amp, gal_exp_sigs, gal_dev_amp, gal_dev_sigs):
    """celeste_fast.pyx:29-94 (joint exp+dev table, K = K_psf*(K_exp+K_dev))."""
    K_psf = image_ws.shape[0]
    num = K_psf * (gal_exp_amp.shape[0] + gal_dev_amp.shape[0])
    weights = np.zeros(num)
    means = np.zeros((num, 2))
    covars = np.zeros((num, 2, 2))
    cnt = 0
    for k in range(K_psf):
        for i in range(2):
            amp, var = (gal_exp_amp, gal_exp_sigs) if i == 0 else (gal_dev_amp, gal_dev_sigs)
            for j in range(amp.shape[0]):
                weights[cnt] = image_ws[k] * thetas[i] * amp[j]
                means[cnt, 0] = v_s[0] + image_means[k, 0]
                means[cnt, 1] = v_s[1] + image_means[k, 1]
                covars[cnt] = image_covars[k] + var[j] * W
                cnt += 1
    return weights, means, covars


LOG2PI = 1.8378770664093453     # gmm_like_fast.pyx:29


def gmm_like_2d(x, ws, mus, sigs, probs=None):
    """gmm_like_fast.pyx:130-176 through the wrapper util/like/__init__.py:7-11.
    Same operation order per component as the reference's inner loop."""
    if mus.shape[0] != sigs.shape[0] or mus.shape[0] != ws.shape[0]:
        raise ValueError("Means, covariances and weights must have same first dimension!")
    if mus.shape[1] != sigs.shape[1] or mus.shape[1] != sigs.shape[2]:
        raise ValueError("Means and inverse covariance shapes don't jive!")
    if probs is None:
        probs = np.zeros(x.shape[0], dtype=np.float64)
    probs[:] = 0.0
    for k in range(mus.shape[0]):
        detk = sigs[k, 0, 0] * sigs[k, 1, 1] - sigs[k, 0, 1] * sigs[k, 1, 0]
        i00 = sigs[k, 1, 1] / detk
        i11 = sigs[k, 0, 0] / detk
        i01 = -1 * sigs[k, 0, 1] / detk
        c0 = x[:, 0] - mus[k, 0]
        c1 = x[:, 1] - mus[k, 1]
        quad = c0 * c0 * i00 + c1 * c1 * i11 + 2. * c0 * c1 * i01
        probs += np.exp(-LOG2PI - .5 * np.log(detk) - .5 * quad) * ws[k]
    return probs


def multivariate_normal_2d_covinv_logdet_logpdf(lls, x, mean, covinv, logdet):
    """gmm_like_fast.pyx:34-77.  NB the reference declares ``float logdet``
    (C single precision, :39), so logdet is rounded to float32 on entry."""
    if mean.shape[0] != x.shape[1]:
        raise ValueError("data dimension 1 doesn't equal mean dimension 0")
    if mean.shape[0] != covinv.shape[0] or mean.shape[0] != covinv.shape[1]:
        raise ValueError("inverse covariance and mean shape aren't equal!")
    if mean.shape[0] != 2:
        raise ValueError("must be two-dimensional!")
    logdet = float(np.float32(logdet))
    mean_sig = (mean[0] * mean[0] * covinv[0, 0] + mean[1] * mean[1] * covinv[1, 1]
                + 2. * mean[0] * mean[1] * covinv[0, 1])
    xs0 = x[:, 0] * covinv[0, 0] + x[:, 1] * covinv[1, 0]
    xs1 = x[:, 0] * covinv[0, 1] + x[:, 1] * covinv[1, 1]
    quad = (x[:, 0] * xs0 + x[:, 1] * xs1) - 2. * (mean[0] * xs0 + mean[1] * xs1) + mean_sig
    lls[:] = -.5 * 2 * LOG2PI - .5 * logdet - .5 * quad
    return lls


# --------------------------------------------------------------------------
# fits_image.py  (header-free constructor; no FITS I/O)
# --------------------------------------------------------------------------
class SyntheticImage(object):
    """The data contract of FitsImage (fits_image.py:16-155) without the FITS /
    tractor / astropy plumbing: the caller passes what the header would hold."""

    def __init__(self, band, nelec, calib, gain, sky, rho_n, phi_n, Ups_n,
                 psf_weights, psf_means, psf_covars, invvar=None, epsilon=None):
        self.band = band
        self.nelec = np.asarray(nelec, dtype=np.float64)
        self.shape = self.nelec.shape
        self.rho_n = np.asarray(rho_n, dtype=np.float64)      # fits_image.py:99
        self.phi_n = np.asarray(phi_n, dtype=np.float64)      # :100
        self.Ups_n = np.asarray(Ups_n, dtype=np.float64)      # :101-102
        self.Ups_n_inv = np.linalg.inv(self.Ups_n)            # :103
        self.kappa = float(gain)                              # :112,118
        self.epsilon = float(sky) * self.kappa if epsilon is None else float(epsilon)  # :113,119
        self.calib = float(calib)                             # :116,122
        self.weights = np.asarray(psf_weights, dtype=np.float64)   # :130
        self.means = np.asarray(psf_means, dtype=np.float64)       # :131
        self.covars = np.asarray(psf_covars, dtype=np.float64)     # :133-138
        K = self.weights.shape[0]
        self.invcovars = np.zeros((K, 2, 2))
        self.logdets = np.zeros(K)
        for i in range(K):
            self.invcovars[i] = np.linalg.inv(self.covars[i])      # :141
            _, self.logdets[i] = np.linalg.slogdet(self.covars[i])  # :144-145
        self.psf = MixtureOfGaussians(means=self.means, covs=self.covars, pis=self.weights)  # :147
        self.R = calc_bounding_radius(self.weights, self.means, self.covars, 0.001)  # :151-155
        self.invvar = np.ones_like(self.nelec) if invvar is None else np.asarray(invvar, dtype=np.float64)

    def equa2pixel(self, s_equa):                              # fits_image.py:166-174
        phi1rad = self.phi_n[1] / 180. * np.pi
        s_iwc = np.array([(s_equa[0] - self.phi_n[0]) * np.cos(phi1rad),
                          (s_equa[1] - self.phi_n[1])])
        return np.dot(self.Ups_n_inv, s_iwc) + self.rho_n

    def pixel2equa(self, s_pixel):                             # fits_image.py:176-181
        phi1rad = self.phi_n[1] / 180. * np.pi
        s_iwc = np.dot(self.Ups_n, s_pixel - self.rho_n)
        return np.array([s_iwc[0] / np.cos(phi1rad) + self.phi_n[0],
                         s_iwc[1] + self.phi_n[1]])

    def nmgy2counts(self, flux):                               # fits_image.py:183-184
        return (flux / self.calib) * self.kappa

    def cd_at_pixel(self, x, y):                               # fits_image.py:196-216
        ra0, dec0 = self.pixel2equa(np.array([x, y]))
        step = 10.
        rax, decx = self.pixel2equa(np.array([x + step, y]))
        ray, decy = self.pixel2equa(np.array([x, y + step]))
        cosd = np.cos(dec0 * (np.pi / 180.))
        return np.array([[(rax - ra0) / step * cosd, (ray - ra0) / step * cosd],
                         [(decx - dec0) / step, (decy - dec0) / step]])


class SrcParams(object):
    """celeste_src.py:12-157 (the fields the hot path reads)."""

    def __init__(self, u, a=None, b=None, t=None, theta=None, phi=None,
                 sigma=None, rho=None, fluxes=None):
        self.u = np.asarray(u, dtype=np.float64)
        self.a, self.b, self.t = a, b, t
        self.theta, self.phi, self.sigma, self.rho = theta, phi, sigma, rho
        self.fluxes = None if fluxes is None else np.asarray(fluxes, dtype=np.float64)

    @property
    def flux_dict(self):                                       # celeste_src.py:141-143
        return {c: v for c, v in zip(BANDS, self.fluxes)}

    @property
    def shape(self):                                           # celeste_src.py:154-156
        return np.array([self.theta, self.sigma, self.phi, self.rho])

    def is_star(self):
        return self.a == 0

    def is_galaxy(self):
        return self.a == 1


# --------------------------------------------------------------------------
# celeste_galaxy_conditionals.py
# --------------------------------------------------------------------------
def gen_galaxy_ra_dec_basis(sig_s, rho_s, phi_s):
    """celeste_galaxy_conditionals.py:90-107."""
    phi = (90. - phi_s) * np.pi / 180.
    re_deg = max(1. / 30, sig_s) / 3600.
    cp, sp = np.cos(phi), np.sin(phi)
    return re_deg * np.array([[cp, sp * rho_s], [-sp, cp * rho_s]])


def gen_galaxy_transformation(sig_s, rho_s, phi_s, Ups_n):
    """celeste_galaxy_conditionals.py:109-125."""
    G = gen_galaxy_ra_dec_basis(sig_s, rho_s, phi_s)
    T = np.dot(np.linalg.inv(G), Ups_n)
    return np.linalg.inv(T)


def _grid(xlim, ylim):
    y_grid = np.arange(ylim[0], ylim[1], dtype=np.float64)
    x_grid = np.arange(xlim[0], xlim[1], dtype=np.float64)
    xx, yy = np.meshgrid(x_grid, y_grid, indexing='xy')
    return xx, np.column_stack((xx.ravel(order='C'), yy.ravel(order='C')))


def gen_galaxy_prof_psf_image(prof_type, R, u, img, xlim=None, ylim=None):
    """celeste_galaxy_conditionals.py:134-182 (Cython route: A1 + A12 + A2)."""
    amp, var = galaxy_prof_dict[prof_type]
    v_s = img.equa2pixel(u)
    weights, means, covars = gen_galaxy_prof_psf_mixture_params(
        np.dot(R, R.T), v_s, img.weights, img.means, img.covars, amp, var)
    bound = calc_bounding_radius(weights, means, covars, 0.00001, center=v_s)
    if xlim is not None and ylim is not None:
        minx_b, maxx_b = xlim
        miny_b, maxy_b = ylim
    else:
        minx_b, maxx_b = max(0, int(v_s[0] - bound)), min(int(v_s[0] + bound + 1), img.nelec.shape[1])
        miny_b, maxy_b = max(0, int(v_s[1] - bound)), min(int(v_s[1] + bound + 1), img.nelec.shape[0])
    xx, pts = _grid((minx_b, maxx_b), (miny_b, maxy_b))
    patch = gmm_like_2d(pts, weights, means, covars)
    return patch.reshape(xx.shape, order='C'), (miny_b, maxy_b), (minx_b, maxx_b)


def galaxy_mixture(th, u_s, img):
    """The 42-component convolved mixture of the live route,
    celeste_galaxy_conditionals.py:193-201."""
    theta_s, sig_s, phi_s, rho_s = th[0:4]
    px, py = img.equa2pixel(u_s)
    galmix = MixtureOfGaussians.convex_combine(galaxy_profs, [theta_s, 1. - theta_s])
    Tinv = gen_galaxy_transformation(sig_s, rho_s, phi_s, img.cd_at_pixel(px, py))
    amix = galmix.apply_affine(Tinv, np.array([px, py]))
    return amix.convolve(img.psf), px, py


def gen_galaxy_psf_image_bound(th, u_s, img):
    """celeste_galaxy_conditionals.py:217-232."""
    cmix, px, py = galaxy_mixture(th, u_s, img)
    return calc_bounding_radius(cmix.pis, cmix.means, cmix.covs, error=1e-5,
                                center=np.array([px, py]))


def gen_galaxy_psf_image(th, u_s, img, xlim=None, ylim=None):
    """celeste_galaxy_conditionals.py:185-214 (live route)."""
    cmix, px, py = galaxy_mixture(th, u_s, img)
    if xlim is None and ylim is None:
        bound = calc_bounding_radius(cmix.pis, cmix.means, cmix.covs, error=1e-5,
                                     center=np.array([px, py]))
        xlim = (np.max([0, np.floor(px - bound)]),
                np.min([img.nelec.shape[1], np.ceil(px + bound)]))
        ylim = (np.max([0, np.floor(py - bound)]),
                np.min([img.nelec.shape[0], np.ceil(py + bound)]))
    return cmix.evaluate_grid(xlim, ylim), ylim, xlim


def galaxy_source_like(th, Z_s, images):
    """celeste_galaxy_conditionals.py:15-42.  th = [theta, sigma, phi, rho,
    u0, u1, b_u..b_z]; Z_s = per-image photon patches covering the footprint the
    Cython route picks."""
    ll = 0.
    theta_s, sig_s, phi_s, rho_s = th[0:4]
    u_s = th[4:6]
    bs = dict(zip(BANDS, th[-5:]))
    for n, img in enumerate(images):
        R_s = gen_galaxy_transformation(sig_s, rho_s, phi_s, img.Ups_n)
        f_exp, _, _ = gen_galaxy_prof_psf_image('exp', R_s, u_s, img)
        f_dev, _, _ = gen_galaxy_prof_psf_image('dev', R_s, u_s, img)
        f = theta_s * f_exp + (1. - theta_s) * f_dev
        lam = (bs[img.band] / img.calib) * img.kappa * f
        ll += np.sum(Z_s[n] * np.log(lam) - lam)
    return ll


# --------------------------------------------------------------------------
# celeste.py
# --------------------------------------------------------------------------
def gen_point_source_psf_image(u, image, xlim=None, ylim=None, check_overlap=True,
                               return_patch=True, psf_grid=None, pixel_grid=None):
    """celeste.py:114-176 (including the reference's overlap test, which
    compares v_s[0] against both image dimensions, :130-132)."""
    v_s = image.equa2pixel(u)
    does_not_overlap = check_overlap and \
        (v_s[0] < -50 or v_s[0] > 2 * image.nelec.shape[0] or
         v_s[1] < -50 or v_s[0] > 2 * image.nelec.shape[1])
    if does_not_overlap:
        return None, None, None
    if xlim is None and ylim is None:
        bound = image.R
        minx_b, maxx_b = max(0, int(v_s[0] - bound)), min(int(v_s[0] + bound + 1), image.nelec.shape[1])
        miny_b, maxy_b = max(0, int(v_s[1] - bound)), min(int(v_s[1] + bound + 1), image.nelec.shape[0])
    else:
        miny_b, maxy_b = ylim
        minx_b, maxx_b = xlim
    if pixel_grid is None:
        _, pixel_grid = _grid((minx_b, maxx_b), (miny_b, maxy_b))
    grid_shape = (int(maxy_b - miny_b), int(maxx_b - minx_b))
    patch = np.exp(mog_loglike(pixel_grid, means=image.means + v_s, icovs=image.invcovars,
                               dets=np.exp(image.logdets), pis=image.weights))
    if return_patch:
        return patch.reshape(grid_shape, order='C'), (miny_b, maxy_b), (minx_b, maxx_b)
    if psf_grid is None:
        psf_grid = np.zeros(image.nelec.shape, dtype=np.float64)
    psf_grid[int(miny_b):int(maxy_b), int(minx_b):int(maxx_b)] = patch.reshape(grid_shape, order='C')
    return psf_grid, (0, psf_grid.shape[0]), (0, psf_grid.shape[1])


def gen_src_image_with_fluxes(src, img, xlim=None, ylim=None):
    """celeste.py:72-96 (rate patch = unit patch * flux/calib*kappa)."""
    if src.a == 0:
        f_s, ylim, xlim = gen_point_source_psf_image(src.u, img, xlim=xlim, ylim=ylim)
        if f_s is None:
            return None, None, None
        f_s = f_s * ((src.flux_dict[img.band] / img.calib) * img.kappa)
    elif src.a == 1:
        psf_img, ylim, xlim = gen_galaxy_psf_image(
            [src.theta, src.sigma, src.phi, src.rho], src.u, img, xlim=xlim, ylim=ylim)
        f_s = ((src.flux_dict[img.band] / img.calib) * img.kappa) * psf_img
    else:
        raise ValueError("source type unknown")
    return f_s, ylim, xlim


def gen_src_image(src, image):
    """celeste.py:26-62 with return_patch=False: full-frame expected-photon
    image of one source (star: temperature route if src.t else flux route)."""
    full = np.zeros(image.nelec.shape)
    if src.a == 0:
        if src.t:
            expected = photons_expected_brightness(src.t, src.b, image.band)
        else:
            expected = image.nmgy2counts(src.flux_dict[image.band])
        f_s, ylim, xlim = gen_point_source_psf_image(src.u, image)
    elif src.a == 1:
        expected = image.nmgy2counts(src.flux_dict[image.band])
        f_s, ylim, xlim = gen_galaxy_psf_image(
            np.array([src.theta, src.sigma, src.phi, src.rho]), src.u, image)
    else:
        raise ValueError("source type unknown")
    if f_s is None:
        return full
    full[int(ylim[0]):int(ylim[1]), int(xlim[0]):int(xlim[1])] = f_s * expected
    return full


def gen_model_image(srcs, image):
    """celeste.py:203-219."""
    f_s = np.zeros(image.nelec.shape)
    for src in srcs:
        f_s += gen_src_image(src, image)
    return image.epsilon + f_s


def celeste_likelihood(srcs, image):
    """celeste.py:237-240."""
    lambdas = gen_model_image(srcs, image)
    return np.sum(image.nelec * np.log(lambdas) - lambdas)


def celeste_likelihood_multi_image(srcs, images):
    """celeste.py:243-252."""
    ll = 0
    for img in images:
        ll += celeste_likelihood(srcs, img)
    return ll


# --------------------------------------------------------------------------
# sources.py
# --------------------------------------------------------------------------
def poisson_loglike(data, model_img, mask):
    """sources.py:6-12."""
    assert model_img.shape == mask.shape
    assert data.shape == model_img.shape
    good_pix = (model_img > 0.) & (mask != 0)
    return np.sum(np.log(model_img[good_pix]) * data[good_pix]) - np.sum(model_img[good_pix])


def compute_scatter_on_pixels(src, fits_image, u=None, shape=None, xlim=None, ylim=None):
    """sources.py:351-388."""
    u = src.u if u is None else u
    if src.is_star():
        return gen_point_source_psf_image(u, fits_image, xlim=xlim, ylim=ylim)
    if src.is_galaxy():
        shape = src.shape if shape is None else shape
        return gen_galaxy_psf_image(shape, u, fits_image, xlim=xlim, ylim=ylim)
    raise NotImplementedError("only stars and galaxies have photon scattering images")


def compute_model_patch(src, fits_image, u=None, xlim=None, ylim=None):
    """sources.py:390-395."""
    patch, ylim, xlim = compute_scatter_on_pixels(src, fits_image, u=u, xlim=xlim, ylim=ylim)
    band_flux = (src.flux_dict[fits_image.band] / fits_image.calib) * fits_image.kappa
    if patch is None:
        return None, None, None
    return band_flux * patch, ylim, xlim


def get_bounding_box(params, img):
    """sources.py:81-94 (floor/ceil footprint used for MH bounding boxes)."""
    if params.is_star():
        bound = img.R
    else:
        bound = gen_galaxy_psf_image_bound(params.shape, params.u, img)
    px, py = img.equa2pixel(params.u)
    xlim = (np.max([0, np.floor(px - bound)]), np.min([img.nelec.shape[1], np.ceil(px + bound)]))
    ylim = (np.max([0, np.floor(py - bound)]), np.min([img.nelec.shape[0], np.ceil(py + bound)]))
    return xlim, ylim


def make_bbox(params, img, pixel_radius):
    """sources.py:434-455 (image_bbox)."""
    img_ymax, img_xmax = img.nelec.shape
    px, py = img.equa2pixel(params.u)
    xlim = (np.max([0, int(np.floor(px - pixel_radius))]),
            np.min([img_xmax, int(np.ceil(px + pixel_radius))]))
    ylim = (np.max([0, int(np.floor(py - pixel_radius))]),
            np.min([img_ymax, int(np.ceil(py + pixel_radius))]))
    return xlim, ylim


def source_log_likelihood(src, sample_image_list, u=None, fluxes=None, shape=None):
    """Source.log_likelihood, sources.py:134-183.  sample_image_list holds
    (z_patch, (y0, y1), (x0, x1), image) tuples (the NativePatch fields)."""
    u = src.u if u is None else u
    fluxes = src.fluxes if fluxes is None else fluxes
    shape = src.shape if shape is None else shape
    ll = 0
    for z, ylim, xlim, img in sample_image_list:
        psf_ns, _, _ = compute_scatter_on_pixels(src, img, u=u, shape=shape, xlim=xlim, ylim=ylim)
        band_flux = (fluxes[BANDS.index(img.band)] / img.calib) * img.kappa
        if psf_ns is None:
            ll += -band_flux * np.sum(img.weights)
            continue
        model_patch = band_flux * psf_ns
        mask = (model_patch > 0.)
        ll += np.sum(np.log(model_patch[mask]) * np.asarray(z)[mask]) - band_flux * np.sum(img.weights)
    return ll


def source_log_likelihood_isolated(src, boxes, images, u=None, fluxes=None, shape=None):
    """Source.log_likelihood_isolated, sources.py:188-237."""
    u = src.u if u is None else u
    fluxes = src.fluxes if fluxes is None else fluxes
    shape = src.shape if shape is None else shape
    ll = 0
    for (ylim, xlim), img in zip(boxes, images):
        data = img.nelec[int(ylim[0]):int(ylim[1]), int(xlim[0]):int(xlim[1])]
        psf_ns, _, _ = compute_scatter_on_pixels(src, img, u=u, shape=shape, xlim=xlim, ylim=ylim)
        band_flux = (fluxes[BANDS.index(img.band)] / img.calib) * img.kappa
        model_patch = band_flux * psf_ns + img.epsilon
        ll += np.sum(np.log(model_patch) * data) - np.sum(model_patch)
    return ll


def image_like(src, img, xlim, ylim, background_img):
    """The per-image closure of Source.calculate_acceptance_logprob,
    sources.py:278-292."""
    x0, x1, y0, y1 = int(xlim[0]), int(xlim[1]), int(ylim[0]), int(ylim[1])
    data_img = img.nelec[y0:y1, x0:x1]
    mask_img = img.invvar[y0:y1, x0:x1]
    model_img, _, _ = compute_model_patch(src, img, xlim=xlim, ylim=ylim)
    return poisson_loglike(data=data_img, model_img=background_img + model_img, mask=mask_img)


def calculate_acceptance_logprob(cur, proposal, images, boxes, backgrounds,
                                 logprior_cur=0., logprior_prop=0.,
                                 logprob_proposal=0., logprob_reverse=0., logdet=0.):
    """Source.calculate_acceptance_logprob, sources.py:275-306 (priors and
    proposal densities passed in as numbers: they are host-side model terms)."""
    curr_like = np.sum([image_like(cur, img, b[0], b[1], bg)
                        for img, b, bg in zip(images, boxes, backgrounds)])
    prop_like = np.sum([image_like(proposal, img, b[0], b[1], bg)
                        for img, b, bg in zip(images, boxes, backgrounds)])
    return (prop_like + logprior_prop) - (curr_like + logprior_cur) + \
        (logprob_reverse - logprob_proposal) + logdet


def generate_background_patch(source, others, image, xlim, ylim):
    """sources.py:476-483 with the active set passed in."""
    bg = np.zeros((int(ylim[1] - ylim[0]), int(xlim[1] - xlim[0])))
    for s in others:
        p, _, _ = compute_model_patch(s, image, xlim=xlim, ylim=ylim)
        if p is not None:
            bg = bg + p
    return bg + image.epsilon


# --------------------------------------------------------------------------
# planck.py
# --------------------------------------------------------------------------
h = 6.6260693e-34
k = 1.3806488e-23
c = 299792458
hc_k = (h * c) / k
hpicc2 = 2 * np.pi * h * c * c
sigma = 2 * np.pi ** 5 * k ** 4 / (15 * c ** 2 * h ** 3)
sun_wattage = 3.846 * 1e26
m_per_ly = c * 31556952.
lens_area = .75 * np.pi * 1.25 ** 2
exposure_duration = 54.


def load_filter_curves(filter_file):
    """planck.py:31-55.  ``filter_file`` is the reference's own data file
    (CelestePy/filter_curves) or the copy exported by the product package."""
    wl = {b: [] for b in BANDS}
    sens = {b: [] for b in BANDS}
    with open(filter_file) as f:
        for line in f:
            band, wavelength, sensitivity = line.strip().split('\t')
            wl[band].append(float(wavelength) * 1e-4 * 1e-6)
            sens[band].append(float(sensitivity))
    h_vecs = {}
    for b in BANDS:
        wl[b] = np.array(wl[b])
        sens[b] = np.array(sens[b])
        order = wl[b].argsort()
        wl[b] = wl[b][order]
        sens[b] = sens[b][order]
        dLam = np.abs(wl[b][2] - wl[b][1])
        h_vecs[b] = sens[b] / (h * c / wl[b]) * dLam
    return wl, sens, h_vecs


_FILTERS = None


def set_filter_curves(filter_file):
    global _FILTERS
    _FILTERS = load_filter_curves(filter_file)
    return _FILTERS


def _filters():
    if _FILTERS is None:
        raise RuntimeError("call set_filter_curves(path) first")
    return _FILTERS


def spec_density(T, wavelengths):
    """planck.py:82-92."""
    w5 = wavelengths * wavelengths * wavelengths * wavelengths * wavelengths
    T4 = T * T * T * T
    with np.errstate(over='ignore'):
        radiances = hpicc2 / (w5 * (np.exp(hc_k / (wavelengths * T)) - 1))
    return radiances / (sigma * T4)


def photons_per_joule(T, band):
    """planck.py:94-142 (without the one-slot cache, which changes no value)."""
    wl, _, h_vecs = _filters()
    if T <= 0:
        return 0
    return spec_density(T, wl[band]).dot(h_vecs[band])


def photons_expected_brightness(T, b, band):
    """planck.py:155-158."""
    lens_watts = lens_area * b * sun_wattage / (m_per_ly ** 2)
    return photons_per_joule(T, band) * lens_watts * exposure_duration


# --------------------------------------------------------------------------
# experiments/redshift
# --------------------------------------------------------------------------
flux_constant = 275422870333.81744384765625   # redshift_utils.py:303


def project_to_bands(spectra, wavelengths):
    """experiments/redshift/redshift_utils.py:304-323."""
    wl, sens, _ = _filters()
    fluxes = np.zeros(5)
    for i, band in enumerate(BANDS):
        sensitivity = np.interp(wavelengths, wl[band] * (10 ** 10), sens[band])
        norm = np.sum(sensitivity)
        if norm <= 0.:
            fluxes[i] = 0.
        else:
            flambda2fnu = wavelengths ** 2 / 2.99792e18
            fthru = np.sum(sensitivity * spectra * flambda2fnu) / norm
            fluxes[i] = fthru * flux_constant
    return fluxes


def pixel_likelihood(z, w, m, fluxes, fluxes_ivar, lam0, B):
    """experiments/redshift/quasar_infer_photometry.py:31-46."""
    if np.isinf(m):
        return -np.inf
    lam_obs = lam0 * (1. + z)
    spec = np.dot(w, B)
    mu = project_to_bands(spec, lam_obs) * m / (1. + z)
    return -0.5 * np.sum(fluxes_ivar * (fluxes - mu) * (fluxes - mu))


def softmax(x):
    """experiments/redshift/redshift_utils.py:394-397."""
    x_tilde = np.exp(x - x.max())
    return x_tilde / x_tilde.sum()


def set_filter_tables_npz(npz_path):
    """Same tables as load_filter_curves (planck.py:31-55) from the packed copy
    of the reference's data file (desi_mcmc_b200/data/filter_curves.npz, made by
    tools/make_filter_table.py); /root/reference is absent on the GPU box."""
    global _FILTERS
    z = np.load(npz_path)
    wl, sens, h_vecs = {}, {}, {}
    for b in BANDS:
        wl[b] = z["wl_" + b] * 1e-4 * 1e-6
        sens[b] = z["sens_" + b]
        order = wl[b].argsort()
        wl[b] = wl[b][order]
        sens[b] = sens[b][order]
        dLam = np.abs(wl[b][2] - wl[b][1])
        h_vecs[b] = sens[b] / (h * c / wl[b]) * dLam
    _FILTERS = (wl, sens, h_vecs)
    return _FILTERS


# --------------------------------------------------------------------------
# celeste_sample_sources.pyx (N1).  rk_binomial of the vendored randomkit is
# the generator behind numpy.random.RandomState.binomial (verified stream-equal
# in tests/test_cpu_oracle.py), so the legacy RandomState stands in for the
# rk_state the Cython code reaches into.
# --------------------------------------------------------------------------
class NativePatch(object):
    """celeste_sample_sources.pyx:31-41."""

    def __init__(self, data, ylim, xlim):
        self.data = data
        self.x0, self.x1 = int(xlim[0]), int(xlim[1])
        self.y0, self.y1 = int(ylim[0]), int(ylim[1])


def sample_multinomial(n, probs, random_state):
    """celeste_sample_sources.pyx:127-156 (conditional binomials)."""
    S = probs.shape[0]
    samples = np.zeros(S, dtype=np.int64)
    unsampled = int(n)
    sum_probs = 0.
    for s in range(S):
        sum_probs += probs[s]
    for s in range(S - 1):
        curr_prob = probs[s]
        curr_sample = int(random_state.binomial(unsampled, curr_prob / sum_probs))
        samples[s] += curr_sample
        unsampled -= curr_sample
        if unsampled <= 0:
            break
        sum_probs -= curr_prob
    samples[S - 1] += unsampled
    return samples


def sample_source_counts(src_imgs, samp_imgs, nelec, src_boxes, epsilon, random_state=None):
    """celeste_sample_sources.pyx:61-122, including the strict box test of
    get_bounding_boxes_idx (:53-54)."""
    random_state = random_state or np.random.RandomState()
    dim0, dim1 = nelec.shape
    noise_sum = 0.
    for y in range(dim0):
        for x in range(dim1):
            num_photons_xy = nelec[y, x]
            possible = [s for s in range(src_boxes.shape[0])
                        if x > src_boxes[s, 0] and x < src_boxes[s, 1] and
                        y > src_boxes[s, 2] and y < src_boxes[s, 3]]
            if len(possible) == 0:
                noise_sum += num_photons_xy
                continue
            S = len(possible) + 1
            model_fluxes = np.zeros(S)
            for i, pi in enumerate(possible):
                p = src_imgs[pi]
                model_fluxes[i] = p.data[y - p.y0, x - p.x0]
            model_fluxes[S - 1] = epsilon
            src_photons = sample_multinomial(int(num_photons_xy), model_fluxes, random_state)
            for i, pi in enumerate(possible):
                p = samp_imgs[pi]
                p.data[y - p.y0, x - p.x0] = src_photons[i]
            noise_sum += src_photons[S - 1]
    return samp_imgs, noise_sum


def galaxy_source_like_grad(th, Z_s, images):
    """celeste_galaxy_conditionals.py:44-88: analytic in theta and the fluxes,
    central differences (1e-5) in sigma, phi, rho, u."""
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
        grad_theta_s += np.sum((Z_s[n] / f_nms - bs[img.band]) * (f_exp - f_dev))
        grad_bs[img.band] += 1. / bs[img.band] * np.sum(Z_s[n]) - np.sum(f_nms)
    grad_RU = np.zeros(5)
    for i, th_i in enumerate([1, 2, 3, 4, 5]):
        de = np.zeros(th.shape)
        de[th_i] = 1e-5
        dell = galaxy_source_like(th + de, Z_s, images) - galaxy_source_like(th - de, Z_s, images)
        grad_RU[i] = dell / (2. * de[th_i])
    return np.concatenate([[grad_theta_s], grad_RU, np.array([grad_bs[b] for b in BANDS])])


# --------------------------------------------------------------------------
# util/dists/flux_prior.py:4-93 and the prior refit's sufficient statistics
# (experiments/empirical_priors/make_source_priors.py:76-110)
# --------------------------------------------------------------------------
FLUX_COLOR_A = np.zeros((5, 5))
FLUX_COLOR_A[0, 0] = FLUX_COLOR_A[1, 1] = FLUX_COLOR_A[3, 2] = FLUX_COLOR_A[4, 3] = 1.
FLUX_COLOR_A[2, :] = -1
FLUX_COLOR_A[2, -1] = 1.


def mog_params(covs):
    """MixtureOfGaussians.update_params, util/dists/mog.py:55-56."""
    return (np.array([np.linalg.inv(c) for c in covs]), np.array([np.linalg.det(c) for c in covs]))


def flux_color_logpdf(log_fluxes, means, covs, pis):
    """FluxColorMoG.logpdf, flux_prior.py:39-40."""
    icovs, dets = mog_params(covs)
    return mog_loglike(np.dot(log_fluxes, FLUX_COLOR_A), means, icovs, dets, pis)


def gal_shape_unconstrained(x):
    """GalShapeMoG.to_unconstrained, flux_prior.py:49-54."""
    sigma, ab, phi = x[0], x[1], x[2]
    return np.array([np.log(sigma), np.log(ab) - np.log(1. - ab), np.log(phi) - np.log(np.pi - phi)])


def mog_responsibilities(x, means, covs, pis):
    icovs, dets = mog_params(covs)
    xx = np.atleast_2d(x)
    centered = xx[:, :, np.newaxis] - means.T[np.newaxis, :, :]
    solved = np.einsum('ijk,lji->lki', icovs, centered)
    logprobs = (-0.5 * np.sum(solved * centered, axis=1) - np.log(2 * np.pi)
                - 0.5 * np.log(dets) + np.log(pis))
    return np.exp(logprobs - logsumexp(logprobs, axis=1)[:, None])


def prior_sufficient_stats(resp, colors):
    """[sum r_k | sum r_k c | sum r_k c c^T (upper triangle, np.triu_indices order)]."""
    D = colors.shape[1]
    iu = np.triu_indices(D)
    s0 = resp.sum(axis=0)
    s1 = resp.T @ colors
    s2 = resp.T @ (colors[:, iu[0]] * colors[:, iu[1]])
    return np.concatenate([s0, s1.ravel(), s2.ravel()])



# --------------------------------------------------------------------------
# fits_image.py:48-155 (file route) and util/data/get_data.py:38-42,89-136
# --------------------------------------------------------------------------
def fits_image_from_header(band, img, header):
    """FitsImage.__init__ for a frame/stamp whose header carries CALIB, SKY, GAIN,
    CRPIX/CRVAL/CD and PSF_P0..17: dn = img / CALIB + SKY, nelec = round(dn * GAIN)
    (:85-87), rho_n = CRPIX - 1 (:99), psf vector layout (:126-138)."""
    dn = np.asarray(img, dtype=np.float64) / header["CALIB"] + header["SKY"]
    nelec = np.round(dn * header["GAIN"])
    psfvec = np.array([header["PSF_P%d" % i] for i in range(18)])
    cv = psfvec[9:].reshape(3, 3)
    covars = np.zeros((3, 2, 2))
    for i in range(3):
        covars[i] = np.array([[cv[i, 0], cv[i, 2]], [cv[i, 2], cv[i, 1]]])
    return SyntheticImage(band=band, nelec=nelec, calib=header["CALIB"], gain=header["GAIN"],
                          sky=header["SKY"],
                          rho_n=np.array([header["CRPIX1"], header["CRPIX2"]]) - 1,
                          phi_n=np.array([header["CRVAL1"], header["CRVAL2"]]),
                          Ups_n=np.array([[header["CD1_1"], header["CD1_2"]],
                                          [header["CD2_1"], header["CD2_2"]]]),
                          psf_weights=psfvec[0:3], psf_means=psfvec[3:9].reshape(3, 2), psf_covars=covars)


def mags2nanomaggies(mags):
    return np.power(10., (mags - 22.5) / -2.5)
