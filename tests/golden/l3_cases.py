"""Seeded L3 cases run through two backends with the same keys:

  * ``run(reference_backend())`` -- the reference's OWN Python (celeste.py,
    celeste_galaxy_conditionals.py, sources.py, fits_image.py, planck.py, util/...),
    imported in place from /root/reference by oracle/ref_import.py;
  * ``run(oracle_backend())``    -- oracle/ref_cpu.py, the restatement the GPU parity
    tests check the CUDA path against.

tests/golden/make_golden_l3.py stores the first as tests/golden/reference_l3.npz (it cannot
travel to the GPU box otherwise); tests/test_cpu_oracle_l3.py asserts the second equals it.
Test infrastructure only.
"""
from __future__ import annotations

import types

import numpy as np

from desi_mcmc_b200 import synth


# ---------------------------------------------------------------------------
# backends
# ---------------------------------------------------------------------------
def oracle_backend():
    import os
    from desi_mcmc_b200 import _ffi
    from oracle import ref_cpu as ref
    ref.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    b = types.SimpleNamespace(name="oracle")
    b.make_image = lambda meta, nelec, invvar=None: ref.SyntheticImage(nelec=nelec, invvar=invvar,
                                                                       **synth.image_kwargs(meta))
    b.make_src = lambda s: ref.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"],
                                         phi=s["phi"], sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"])
    b.calc_bounding_radius = ref.calc_bounding_radius
    b.mog_loglike = ref.mog_loglike
    b.MixtureOfGaussians = ref.MixtureOfGaussians
    b.gen_galaxy_ra_dec_basis = ref.gen_galaxy_ra_dec_basis
    b.gen_galaxy_transformation = ref.gen_galaxy_transformation
    b.gen_point_source_psf_image = ref.gen_point_source_psf_image
    b.gen_galaxy_psf_image = ref.gen_galaxy_psf_image
    b.gen_galaxy_psf_image_bound = lambda src, img: ref.gen_galaxy_psf_image_bound(
        [src.theta, src.sigma, src.phi, src.rho], src.u, img)
    b.gen_src_image_with_fluxes = ref.gen_src_image_with_fluxes
    b.gen_model_image = ref.gen_model_image
    b.celeste_likelihood = ref.celeste_likelihood
    b.celeste_likelihood_multi_image = ref.celeste_likelihood_multi_image
    b.poisson_loglike = ref.poisson_loglike
    b.photons_per_joule = ref.photons_per_joule
    b.photons_expected_brightness = ref.photons_expected_brightness
    b.exp_mixture = lambda: (ref.exp_amp, ref.exp_var)
    b.dev_mixture = lambda: (ref.dev_amp, ref.dev_var)

    def source_lls(src, sample_image_list, boxes, images, u=None):
        lst = [(z.data, (z.y0, z.y1), (z.x0, z.x1), img) for z, img in sample_image_list]
        bx = [((y0, y1), (x0, x1)) for (x0, x1, y0, y1) in boxes]
        return (ref.source_log_likelihood(src, lst, u=u),
                ref.source_log_likelihood_isolated(src, bx, images, u=u))
    b.source_lls = source_lls

    def accept_logprob(cur, prop, images, boxes, backgrounds):
        bx = [((x0, x1), (y0, y1)) for (x0, x1, y0, y1) in boxes]
        return ref.calculate_acceptance_logprob(cur, prop, images, bx, backgrounds)
    b.accept_logprob = accept_logprob
    b.generate_background_patch = ref.generate_background_patch
    return b


def reference_backend():
    """The reference's own modules (needs /root/reference and oracle/_ref)."""
    from oracle import ref_import as ri
    ri.install()
    fits_image = ri.load("fits_image")
    celeste = ri.load("celeste")
    gal = ri.load("celeste_galaxy_conditionals")
    sources = ri.load("sources")
    planck = ri.load("planck")
    mog = ri.load("util.dists.mog")
    bb = ri.load("util.bound.bounding_box")
    celeste_src = ri.load("celeste_src")
    mp = ri.load("mixture_profiles")

    b = types.SimpleNamespace(name="reference")

    def make_image(meta, nelec, invvar=None):
        """A FitsImage with the attributes its __init__ derives from a FITS header
        (fits_image.py:85-155), set from the same numbers instead of a file."""
        im = fits_image.FitsImage.__new__(fits_image.FitsImage)
        im.band = meta["band"]
        im.nelec = np.array(nelec, dtype=np.float64)
        im.shape = im.nelec.shape
        im.invvar = np.ones(im.shape) if invvar is None else np.array(invvar, dtype=np.float64)
        im.rho_n = np.array(meta["rho_n"], dtype=np.float64)
        im.phi_n = np.array(meta["phi_n"], dtype=np.float64)
        im.Ups_n = np.array(meta["Ups_n"], dtype=np.float64)
        im.Ups_n_inv = np.linalg.inv(im.Ups_n)                         # :102
        im.use_wcs = False
        im.kappa = meta["gain"]                                        # :118-122
        im.epsilon = meta["sky"] * im.kappa
        im.epsilon0 = im.epsilon
        im.calib = meta["calib"]
        im.weights = np.array(meta["psf_weights"], dtype=np.float64)
        im.means = np.array(meta["psf_means"], dtype=np.float64)
        im.covars = np.array(meta["psf_covars"], dtype=np.float64)
        im.invcovars = np.array([np.linalg.inv(c) for c in im.covars])      # :141
        im.logdets = np.array([np.linalg.slogdet(c)[1] for c in im.covars])  # :144-145
        im.psf_mog = mog.MixtureOfGaussians(means=im.means, covs=im.covars, pis=im.weights)
        im.R = bb.calc_bounding_radius(im.weights, im.means, im.covars, 0.001)   # :151-155
        im.pixel_grid = im.make_pixel_grid()
        return im
    b.make_image = make_image
    class _Fluxes(np.ndarray):
        """The reference reads SrcParams.fluxes both as a ugriz sequence (flux_dict,
        celeste_src.py:141-143) and as a band-keyed dict (celeste.py:41,50): an array that also
        answers to the band letter serves both call sites."""

        def __getitem__(self, k):
            if isinstance(k, str):
                k = "ugriz".index(k)
            return np.ndarray.__getitem__(self, k)

    b.make_src = lambda s: celeste_src.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"],
                                                 phi=s["phi"], sigma=s["sigma"], rho=s["rho"],
                                                 fluxes=np.asarray(s["fluxes"], dtype=np.float64).view(_Fluxes))
    b.calc_bounding_radius = bb.calc_bounding_radius
    b.mog_loglike = mog.mog_loglike
    b.MixtureOfGaussians = mog.MixtureOfGaussians
    b.gen_galaxy_ra_dec_basis = gal.gen_galaxy_ra_dec_basis
    b.gen_galaxy_transformation = gal.gen_galaxy_transformation
    b.gen_point_source_psf_image = celeste.gen_point_source_psf_image
    b.gen_galaxy_psf_image = lambda th, u, img, xlim=None, ylim=None: gal.gen_galaxy_psf_image(
        th, u, img, xlim=xlim, ylim=ylim, check_overlap=True, unconstrained=False, return_patch=True)
    b.gen_galaxy_psf_image_bound = gal.gen_galaxy_psf_image_bound
    b.gen_src_image_with_fluxes = celeste.gen_src_image_with_fluxes
    b.gen_model_image = celeste.gen_model_image
    b.celeste_likelihood = celeste.celeste_likelihood
    b.celeste_likelihood_multi_image = celeste.celeste_likelihood_multi_image
    b.poisson_loglike = sources.poisson_loglike
    b.photons_per_joule = planck.photons_per_joule
    b.photons_expected_brightness = planck.photons_expected_brightness
    b.exp_mixture = lambda: (mp.get_exp_mixture().amp, mp.get_exp_mixture().var[:, 0, 0])
    b.dev_mixture = lambda: (mp.get_dev_mixture().amp, mp.get_dev_mixture().var[:, 0, 0])

    def _source(params, images=None, boxes=None, backgrounds=None):
        model = types.SimpleNamespace(logprior=lambda p: 0., _source_type=None)
        model._source_type = lambda p, m: _bind(sources.Source(p, m))

        def _bind(s):
            if images is not None:
                s.bounding_boxes = {im: ((bx[0], bx[1]), (bx[2], bx[3])) for im, bx in zip(images, boxes)}
                s.background_image_dict = {im: bg for im, bg in zip(images, backgrounds)}
            return s
        return _bind(sources.Source(params, model))

    def source_lls(src, sample_image_list, boxes, images, u=None):
        """Source.log_likelihood (sources.py:134-183) and log_likelihood_isolated (:188-237)."""
        s = _source(src)
        # (NativePatch-like, FitsImage, pixel_grid) tuples as Field.resample_photons builds them
        lst = []
        for z, img in sample_image_list:
            y_grid = np.arange(z.y0, z.y1, dtype=float)
            x_grid = np.arange(z.x0, z.x1, dtype=float)
            xx, yy = np.meshgrid(x_grid, y_grid, indexing='xy')
            lst.append((z, img, np.column_stack((xx.ravel(order='C'), yy.ravel(order='C')))))
        s.sample_image_list = lst
        return s.log_likelihood(u=u), s.log_likelihood_isolated(u=u)
    b.source_lls = source_lls

    def accept_logprob(cur, prop, images, boxes, backgrounds):
        """Source.calculate_acceptance_logprob (sources.py:275-306), flat priors."""
        s = _source(cur, images, boxes, backgrounds)
        s.model._source_type = lambda p, m: _source(p, images, boxes, backgrounds)
        return s.calculate_acceptance_logprob(prop, 0., 0., 0., images)
    b.accept_logprob = accept_logprob

    def generate_background_patch(source, others, image, xlim, ylim):
        """sources.py:476-483 with the active set given (every other source intersects here)."""
        bbs = {image: (xlim, ylim)}
        me = _source(source)
        me.bounding_boxes = bbs
        lst = []
        for o in others:
            so = _source(o)
            so.bounding_boxes = bbs
            lst.append(so)
        return sources.generate_background_patch(me, lst + [me], image)
    b.generate_background_patch = generate_background_patch
    return b


# ---------------------------------------------------------------------------
# cases
# ---------------------------------------------------------------------------
class _Patch(object):
    """what NativePatch exposes (celeste_sample_sources.pyx:31-41)"""

    def __init__(self, data, ylim, xlim):
        self.data = np.array(data, dtype=np.float64)
        self.y0, self.y1 = int(ylim[0]), int(ylim[1])
        self.x0, self.x1 = int(xlim[0]), int(xlim[1])


def _put(out, key, val):
    out[key] = np.asarray(val, dtype=np.float64)


def run(b) -> dict:
    """All cases on backend b -> {key: float64 array}."""
    out = {}
    rng = np.random.RandomState(20260)

    # ---- A12 / A3 / A4: mixtures -------------------------------------------------
    field = synth.make_stamp(1001, size=25)                     # BASELINE config 1
    metas = field["images"]
    ea, ev = b.exp_mixture()
    da, dv = b.dev_mixture()
    _put(out, "prof_exp_amp", ea); _put(out, "prof_exp_var", ev)
    _put(out, "prof_dev_amp", da); _put(out, "prof_dev_var", dv)
    for i, m in enumerate(metas):
        w, mu, cv = m["psf_weights"], m["psf_means"], m["psf_covars"]
        _put(out, "bound_psf_%d" % i, b.calc_bounding_radius(w, mu, cv, 0.001))
        _put(out, "bound_psf5_%d" % i, b.calc_bounding_radius(w, mu + 3.5, cv, 1e-5, center=np.array([3., 4.])))
    K = 7
    means = rng.normal(0., 3., size=(K, 2))
    A = rng.normal(size=(K, 2, 2))
    covs = np.einsum("kij,klj->kil", A, A) + .5 * np.eye(2)
    pis = rng.dirichlet(np.ones(K))
    x = rng.normal(0., 4., size=(300, 2))
    icovs = np.array([np.linalg.inv(c) for c in covs])
    dets = np.array([np.linalg.det(c) for c in covs])
    _put(out, "mog_loglike", b.mog_loglike(x, means, icovs, dets, pis))
    mg = b.MixtureOfGaussians(means, covs, pis)
    psf = b.MixtureOfGaussians(metas[2]["psf_means"], metas[2]["psf_covars"], metas[2]["psf_weights"])
    _put(out, "mog_logpdf", mg.logpdf(x))
    cv_ = mg.convolve(psf)
    _put(out, "mog_conv_means", cv_.means); _put(out, "mog_conv_covs", cv_.covs); _put(out, "mog_conv_pis", cv_.pis)
    Aff = np.array([[1.3, -.4], [.2, .8]])
    af = mg.apply_affine(Aff, np.array([2., -1.]))
    _put(out, "mog_aff_means", af.means); _put(out, "mog_aff_covs", af.covs)
    cc = b.MixtureOfGaussians.convex_combine([mg, psf], [.3, .7])
    _put(out, "mog_cc_pis", cc.pis); _put(out, "mog_cc_means", cc.means)
    _put(out, "mog_grid", cv_.evaluate_grid((-6, 9), (-5, 7)))

    # ---- A9: galaxy transformation --------------------------------------------------
    nelecs = [rng.poisson(400., size=m["shape"]).astype(np.float64) for m in metas]
    imgs = [b.make_image(m, n) for m, n in zip(metas, nelecs)]
    star, galx = field["sources"]
    _put(out, "img_R", [im.R for im in imgs])
    _put(out, "img_eps", [im.epsilon for im in imgs])
    _put(out, "equa2pixel", [im.equa2pixel(galx["u"]) for im in imgs])
    _put(out, "pixel2equa", [im.pixel2equa(np.array([3.25, 17.5])) for im in imgs])
    _put(out, "cd_at_pixel", [im.cd_at_pixel(11.3, 7.9) for im in imgs])
    _put(out, "gal_basis", b.gen_galaxy_ra_dec_basis(galx["sigma"], galx["rho"], galx["phi"]))
    _put(out, "gal_transform", [b.gen_galaxy_transformation(galx["sigma"], galx["rho"], galx["phi"],
                                                             im.cd_at_pixel(11.3, 7.9)) for im in imgs])

    # ---- A5 / A10 / A11: unit-flux patches on the C1 stamp ----------------------------
    th = [galx["theta"], galx["sigma"], galx["phi"], galx["rho"]]
    for i, im in enumerate(imgs):
        p, yl, xl = b.gen_point_source_psf_image(star["u"], im)
        _put(out, "star_patch_%d" % i, p); _put(out, "star_box_%d" % i, [xl[0], xl[1], yl[0], yl[1]])
        p, yl, xl = b.gen_galaxy_psf_image(th, galx["u"], im)
        _put(out, "gal_patch_%d" % i, p); _put(out, "gal_box_%d" % i, [xl[0], xl[1], yl[0], yl[1]])
        p, yl, xl = b.gen_galaxy_psf_image(th, galx["u"], im, xlim=(3, 21), ylim=(2, 19))
        _put(out, "gal_patch_lim_%d" % i, p)
    p = b.gen_point_source_psf_image(imgs[0].pixel2equa(np.array([-80., 5.])), imgs[0])   # celeste.py:130-134
    _put(out, "star_off_image", [0. if p[0] is None else 1.])
    _put(out, "gal_bound", [b.gen_galaxy_psf_image_bound(b.make_src(galx), im) for im in imgs])

    # a larger frame: natural footprints, truncation vs floor/ceil quirks, image edges
    fld = synth.make_field(77, H=120, W=150, n_src=10, star_frac=.5)
    fnel = [rng.poisson(430., size=m["shape"]).astype(np.float64) for m in fld["images"]]
    mask = (rng.uniform(size=fld["images"][1]["shape"]) > .1).astype(np.float64)
    fimgs = [b.make_image(m, n, mask if k == 1 else None) for k, (m, n) in enumerate(zip(fld["images"], fnel))]
    fsrcs = [b.make_src(s) for s in fld["sources"]]
    for k, (sd, s) in enumerate(zip(fld["sources"], fsrcs)):
        for i in (0, 2, 4):
            p, yl, xl = b.gen_src_image_with_fluxes(s, fimgs[i])               # A6
            if p is None:
                _put(out, "rate_box_%d_%d" % (k, i), [-1, -1, -1, -1])
                continue
            _put(out, "rate_patch_%d_%d" % (k, i), p)
            _put(out, "rate_box_%d_%d" % (k, i), [xl[0], xl[1], yl[0], yl[1]])
    # (gen_galaxy_prof_psf_image, celeste_galaxy_conditionals.py:134-182, cannot be executed at the
    #  reference's HEAD: galaxy_prof_dict (:129-131) holds util.dists.mog mixtures, which have no
    #  .amp / .var.  Its pieces ARE pinned: A1 tables and A2 by the compiled Cython, A12 above.)
    # A7 / A8.  celeste.gen_model_image (celeste.py:203-219) asks gen_galaxy_psf_image for a full frame
    # (return_patch=False) but celeste_galaxy_conditionals.py:185-214 ignores that flag, so at the
    # reference's HEAD it only runs for star-only lists: pinned on the stars of this field.  Galaxies
    # enter the model image through their rate patches (pinned above) added on their boxes, which
    # is what models.py:88-102 does.
    stars_only = [s for s, sd in zip(fsrcs, fld["sources"]) if sd["a"] == 0]
    _put(out, "model_image_2", b.gen_model_image(stars_only, fimgs[2]))
    _put(out, "loglik", [b.celeste_likelihood(stars_only, im) for im in fimgs])
    _put(out, "loglik_multi", b.celeste_likelihood_multi_image(stars_only, fimgs))
    lam = np.zeros(fimgs[1].nelec.shape) + fimgs[1].epsilon
    for s in fsrcs:
        p, yl, xl = b.gen_src_image_with_fluxes(s, fimgs[1])
        if p is not None:
            lam[int(yl[0]):int(yl[1]), int(xl[0]):int(xl[1])] += p
    _put(out, "poisson_masked", b.poisson_loglike(fimgs[1].nelec, lam, mask))      # sources.py:6-12

    # ---- A11 conditional galaxy likelihood, A13 / A14 / A15 ------------------------------
    gi = [k for k, s in enumerate(fld["sources"]) if s["a"] == 1][0]
    g = fld["sources"][gi]
    Zs, samp_list, boxes = [], [], []
    for im in fimgs:
        p, yl, xl = b.gen_src_image_with_fluxes(fsrcs[gi], im)
        z = rng.poisson(p).astype(np.float64)
        Zs.append(z)
        samp_list.append((_Patch(z, yl, xl), im))
        boxes.append((int(xl[0]), int(xl[1]), int(yl[0]), int(yl[1])))
    # (galaxy_source_like, celeste_galaxy_conditionals.py:15-42, cannot be executed as written: it
    #  multiplies the (patch, ylim, xlim) TUPLE gen_galaxy_prof_psf_image returns by theta.  Its two
    #  per-profile renders are pinned above; the sum is the oracle's reading of the intent.)
    si = [k for k, s in enumerate(fld["sources"]) if s["a"] == 0][0]
    for tag, k in (("gal", gi), ("star", si)):
        sl, bx = [], []
        for im in fimgs:
            p, yl, xl = b.gen_src_image_with_fluxes(fsrcs[k], im)
            sl.append((_Patch(rng.poisson(p).astype(np.float64), yl, xl), im))
            bx.append((int(xl[0]), int(xl[1]), int(yl[0]), int(yl[1])))
        u2 = fld["sources"][k]["u"] + np.array([1.5e-5, -2e-5])
        _put(out, "src_ll_%s" % tag, b.source_lls(fsrcs[k], sl, bx, fimgs))
        _put(out, "src_ll_moved_%s" % tag, b.source_lls(fsrcs[k], sl, bx, fimgs, u=u2))
    # MH score over cached boxes / backgrounds (sources.py:275-306)
    others = [s for k, s in enumerate(fsrcs) if k != gi]
    bgs = []
    for im, bx in zip(fimgs, boxes):
        bgs.append(b.generate_background_patch(fsrcs[gi], others, im, (bx[0], bx[1]), (bx[2], bx[3])))
    _put(out, "background_2", bgs[2])
    props, _ = synth.make_proposals(5, g, 6)
    _put(out, "mh_delta", [b.accept_logprob(fsrcs[gi], b.make_src(p), fimgs, boxes, bgs) for p in props])

    # ---- A16 ---------------------------------------------------------------------------------
    Ts = [3000., 5777., 8395., 10000.]
    _put(out, "planck_ppj", [[b.photons_per_joule(T, band) for band in "ugriz"] for T in Ts])
    _put(out, "planck_expected", [[b.photons_expected_brightness(T, 2.5e-11, band) for band in "ugriz"] for T in Ts])
    return out


# ---------------------------------------------------------------------------
# A17 and the host samplers
# ---------------------------------------------------------------------------
def run_a17(project_to_bands, softmax) -> dict:
    """experiments/redshift/redshift_utils.py:304-323, :394-397 on a synthetic positive basis
    (K = 4, V = 1364, the shape of cache/basis_locked/*K-4_V-1364*)."""
    out = {}
    rng = np.random.RandomState(1004)
    K, V = 4, 1364
    lam0 = np.exp(np.linspace(np.log(700.), np.log(9000.), V))
    B = np.abs(rng.normal(size=(K, V))) + .05
    B = B / B.sum(axis=1, keepdims=True)
    for j, z in enumerate((0., .35, 2.2, 5.5)):
        w = softmax(rng.normal(size=K))
        _put(out, "a17_w_%d" % j, w)
        _put(out, "a17_flux_%d" % j, project_to_bands(np.dot(w, B), lam0 * (1. + z)))
    return out


def sampler_traces(slicesample, hmc) -> dict:
    """util/infer/slicesample.py:89-247 and util/infer/hmc.py:16-98 driven by closed-form
    densities from the global numpy.random stream, as the reference's callers do."""
    out = {}
    mu, sd = np.array([.3, -1., 2.]), np.array([.2, 1.5, .7])
    f = lambda x: float(-.5 * np.sum(((x - mu) / sd) ** 2))
    for tag, kw in (("default", dict()),
                    ("nostep", dict(step_out=False, upper_bound=np.full(3, 9.), lower_bound=np.full(3, -9.),
                                    step=np.full(3, .4))),
                    ("dirs", dict(compwise=False, numdir=3, sigma=.7))):
        np.random.seed(2026)
        x, tr = np.zeros(3), []
        for _ in range(25):
            x, l = slicesample(x, f, **kw)
            tr.append(np.concatenate([x, [l]]))
        _put(out, "trace_slice_" + tag, tr)
    prec = np.array([4., .25])
    llh = lambda x: float(-.5 * np.sum(prec * x * x))
    grad = lambda x: -prec * x
    np.random.seed(11)
    x, p, tr = np.array([.5, -.5]), None, []
    for _ in range(40):
        x, p, eps, _ = hmc(x, llh, grad, eps=.15, num_steps=12, mass=np.ones(2), p_curr=p, refresh_alpha=.3)
        tr.append(np.concatenate([x, p]))
    _put(out, "trace_hmc", tr)
    return out
