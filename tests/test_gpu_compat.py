"""The reference's call surface (desi_mcmc_b200.compat) on the CUDA path against
the CPU oracle, and the device photon allocation (N1)."""
import numpy as np
import pytest

from desi_mcmc_b200 import synth
from tests import helpers as hp

pytestmark = pytest.mark.gpu

RATE_RTOL = 1e-5
LL_RTOL = 1e-6
FLOOR = 1e-12


def _stamp(eng, ref, seed=1001, size=25, nseed=3):
    field = synth.make_stamp(seed, size=size)
    nelec = hp.simulate_nelec(ref, field, nseed)
    return field, hp.gpu_images(eng, field, nelec), hp.oracle_images(ref, field, nelec), \
        hp.gpu_sources(eng, field), hp.oracle_sources(ref, field)


def test_compat_celeste_star_and_model(engine, oracle):
    from desi_mcmc_b200.compat import celeste
    field, gimgs, oimgs, gs, os_ = _stamp(engine, oracle)
    star_g, star_o = gs[0], os_[0]
    for gi, oi in zip(gimgs, oimgs):
        got, yl, xl = celeste.gen_point_source_psf_image(star_g.u, gi)
        want, wyl, wxl = oracle.gen_point_source_psf_image(star_o.u, oi)
        assert (tuple(yl), tuple(xl)) == (tuple(wyl), tuple(wxl)) and got.shape == want.shape
        assert hp.rel_err(got, want, FLOOR * want.max()) < RATE_RTOL
        # caller-supplied limits + full-frame variant
        got2, _, _ = celeste.gen_point_source_psf_image(star_g.u, gi, xlim=(3, 20), ylim=(2, 17))
        want2, _, _ = oracle.gen_point_source_psf_image(star_o.u, oi, xlim=(3, 20), ylim=(2, 17))
        assert hp.rel_err(got2, want2, FLOOR * want2.max()) < RATE_RTOL
        full, fy, fx = celeste.gen_point_source_psf_image(star_g.u, gi, return_patch=False)
        assert full.shape == gi.shape and fy == (0, gi.shape[0])
        for s_g, s_o in zip(gs, os_):
            a, ayl, axl = celeste.gen_src_image_with_fluxes(s_g, gi)
            b, byl, bxl = oracle.gen_src_image_with_fluxes(s_o, oi)
            assert (int(ayl[0]), int(ayl[1]), int(axl[0]), int(axl[1])) == \
                (int(byl[0]), int(byl[1]), int(bxl[0]), int(bxl[1]))
            assert hp.rel_err(a, b, FLOOR * b.max()) < RATE_RTOL
        lam = celeste.gen_model_image(gs, gi)
        assert hp.rel_err(lam, oracle.gen_model_image(os_, oi), 1.0) < RATE_RTOL
        want_ll = oracle.celeste_likelihood(os_, oi)
        assert abs(celeste.celeste_likelihood(gs, gi) - want_ll) <= LL_RTOL * abs(want_ll)
    total = oracle.celeste_likelihood_multi_image(os_, oimgs)
    assert abs(celeste.celeste_likelihood_multi_image(gs, gimgs) - total) <= LL_RTOL * abs(total)
    # off-image source -> (None, None, None)   (celeste.py:133-134)
    far = gimgs[0].pixel2equa(np.array([-400., 5.]))
    assert celeste.gen_point_source_psf_image(far, gimgs[0]) == (None, None, None)


def test_compat_galaxy_routes_and_conditional(engine, oracle):
    from desi_mcmc_b200.compat import celeste_galaxy_conditionals as gal
    field, gimgs, oimgs, gs, os_ = _stamp(engine, oracle, seed=1001, size=25)
    g_o = os_[1]
    th4 = g_o.shape
    for gi, oi in zip(gimgs, oimgs):
        got, yl, xl = gal.gen_galaxy_psf_image(th4, g_o.u, gi)
        want, wyl, wxl = oracle.gen_galaxy_psf_image(th4, g_o.u, oi)
        assert got.shape == want.shape and (yl, xl) == (wyl, wxl)
        assert hp.rel_err(got, want, FLOOR * want.max()) < RATE_RTOL
        R = oracle.gen_galaxy_transformation(th4[1], th4[3], th4[2], oi.Ups_n)
        assert np.allclose(R, gal.gen_galaxy_transformation(th4[1], th4[3], th4[2], gi.Ups_n))
        for prof in ('exp', 'dev'):
            a, ayl, axl = gal.gen_galaxy_prof_psf_image(prof, R, g_o.u, gi)
            b, byl, bxl = oracle.gen_galaxy_prof_psf_image(prof, R, g_o.u, oi)
            assert (ayl, axl) == (byl, bxl)
            assert hp.rel_err(a, b, FLOOR * b.max()) < RATE_RTOL
        assert abs(gal.gen_galaxy_psf_image_bound(gs[1], gi) -
                   oracle.gen_galaxy_psf_image_bound(th4, g_o.u, oi)) < 1e-9
    # conditional likelihood + gradient (celeste_galaxy_conditionals.py:15-88)
    rng = np.random.RandomState(5)
    th = np.concatenate([th4, g_o.u, g_o.fluxes])
    Z = []
    for oi in oimgs:
        R = oracle.gen_galaxy_transformation(th4[1], th4[3], th4[2], oi.Ups_n)
        f, _, _ = oracle.gen_galaxy_prof_psf_image('dev', R, g_o.u, oi)
        Z.append(rng.poisson(2000. * f + .2).astype(np.float64))
    want = oracle.galaxy_source_like(th, Z, oimgs)
    got = gal.galaxy_source_like(th, Z, gimgs)
    assert abs(got - want) <= LL_RTOL * abs(want)
    gwant = oracle.galaxy_source_like_grad(th, Z, oimgs)
    ggot = gal.galaxy_source_like_grad(th, Z, gimgs)
    # analytic entries (theta, fluxes) to rate accuracy; finite-difference entries
    # (sigma, phi, rho, u: differences of 1e5-sized numbers over 2e-5) to the
    # accuracy of the fp32 rates
    assert np.allclose(ggot[[0, 6, 7, 8, 9, 10]], gwant[[0, 6, 7, 8, 9, 10]], rtol=1e-4)
    scale = np.abs(want) * 2e-7 / 2e-5
    assert np.all(np.abs(ggot[1:6] - gwant[1:6]) <= scale + 1e-3 * np.abs(gwant[1:6]))


class _Model(object):
    def logprior(self, params):
        return 0.


def test_compat_source_class(engine, oracle):
    """Source.log_likelihood / log_likelihood_isolated / acceptance and a
    seeded resample_location chain against the oracle."""
    from desi_mcmc_b200.compat import sources as S
    from desi_mcmc_b200.compat.celeste_sample_sources import NativePatch
    from desi_mcmc_b200.compat.slicesample import slicesample
    field, gimgs, oimgs, gs, os_ = _stamp(engine, oracle, seed=1001, size=25, nseed=8)
    rng = np.random.RandomState(21)
    for si in (0, 1):
        src = S.Source(gs[si], _Model())
        yl, xl = (2, 23), (3, 24)
        zs = [rng.poisson(3., size=(yl[1] - yl[0], xl[1] - xl[0])).astype(np.float64) for _ in gimgs]
        src.sample_image_list = [(NativePatch(z, yl, xl), gi, None) for z, gi in zip(zs, gimgs)]
        olist = [(z, yl, xl, oi) for z, oi in zip(zs, oimgs)]
        want = oracle.source_log_likelihood(os_[si], olist)
        assert abs(src.log_likelihood() - want) <= LL_RTOL * abs(want)
        u2 = os_[si].u + np.array([2e-5, -1e-5])
        want = oracle.source_log_likelihood(os_[si], olist, u=u2)
        assert abs(src.location_likelihood(u2) - want) <= LL_RTOL * abs(want)
        want = oracle.source_log_likelihood_isolated(os_[si], [(yl, xl)] * 5, oimgs)
        assert abs(src.log_likelihood_isolated() - want) <= LL_RTOL * abs(want)
        # seeded slice-sampling chain: same uniforms, same decisions, same trace
        np.random.seed(77)
        u_gpu = src.resample_location()
        np.random.seed(77)
        u_cpu, _ = slicesample(os_[si].u.copy(), lambda u: oracle.source_log_likelihood(os_[si], olist, u=u),
                               step=None, step_out=False, upper_bound=os_[si].u + .0025,
                               lower_bound=os_[si].u - .0025)
        assert np.allclose(u_gpu, u_cpu, rtol=0, atol=1e-12)
    # MH acceptance log-probability (sources.py:275-306); fresh parameters, the
    # chains above moved gs[*].u
    gs = hp.gpu_sources(engine, field)
    cur, other = S.Source(gs[1], _Model()), S.Source(gs[0], _Model())
    for s in (cur, other):
        s.bounding_boxes = S.make_bbox_dict(s.params, gimgs, pixel_radius=9)
    cur.background_image_dict = {gi: S.generate_background_patch(cur, [cur, other], gi) for gi in gimgs}
    props, _ = synth.make_proposals(3, field["sources"][1], 1)
    prop_g = hp.gpu_sources(engine, props)[0]
    prop_o = hp.oracle_sources(oracle, props)[0]
    boxes = [oracle.make_bbox(os_[1], oi, 9) for oi in oimgs]
    bgs = [oracle.generate_background_patch(os_[1], [os_[0]], oi, b[0], b[1]) for oi, b in zip(oimgs, boxes)]
    for gi, bg in zip(gimgs, bgs):
        assert hp.rel_err(cur.background_image_dict[gi], bg, 1.0) < RATE_RTOL
    want = oracle.calculate_acceptance_logprob(os_[1], prop_o, oimgs, boxes, bgs, logdet=.25)
    got = cur.calculate_acceptance_logprob(prop_g, 0., 0., .25, gimgs)
    assert abs(got - want) < 2e-3


def test_compat_mog_poisson_planck(engine, oracle):
    import os
    from desi_mcmc_b200 import _ffi
    from desi_mcmc_b200.compat import mog, planck, sources as S, bounding_box as bb
    rng = np.random.RandomState(2)
    K = 5
    means = rng.normal(size=(K, 2)) * 3
    A = rng.normal(size=(K, 2, 2))
    covs = np.einsum('kij,klj->kil', A, A) + np.eye(2)[None] * .5
    pis = rng.dirichlet(np.ones(K))
    x = rng.normal(size=(300, 2)) * 4
    m = mog.MixtureOfGaussians(means, covs, pis)
    om = oracle.MixtureOfGaussians(means, covs, pis)
    # densities within the 1e-5 rate bar (log space: 1e-5 absolute = 1e-5 relative on the density)
    assert np.allclose(m.logpdf(x), om.logpdf(x), rtol=1e-12, atol=1e-12)
    # the fp32 density evaluator behind evaluate_grid / the render kernels: the 1e-5 rate bar
    g2 = mog._gmm_device(x, pis, means, covs)
    assert np.allclose(g2, np.exp(om.logpdf(x)))          # the reference KAT's criterion (test_gmm.py:98)
    # mog_loglike is the float64 log-space evaluator: 1e-12, and finite in the far tail like the
    # reference's logsumexp (mog.py:16-17) where the density itself underflows
    assert np.allclose(mog.mog_loglike(x, means, om.icovs, om.dets, pis), om.logpdf(x), rtol=1e-12, atol=1e-12)
    far = np.array([[400., -350.], [1e3, 1e3]])
    got_far = mog.mog_loglike(far, means, om.icovs, om.dets, pis)
    assert np.all(np.isfinite(got_far)) and np.allclose(got_far, om.logpdf(far), rtol=1e-12)
    psf = mog.MixtureOfGaussians(np.zeros((2, 2)), np.array([np.eye(2), 3 * np.eye(2)]), np.array([.7, .3]))
    opsf = oracle.MixtureOfGaussians(psf.means, psf.covs, psf.pis)
    c, oc = m.apply_affine(A[0], np.array([1., 2.])).convolve(psf), om.apply_affine(A[0], np.array([1., 2.])).convolve(opsf)
    assert np.allclose(c.covs, oc.covs) and np.allclose(c.means, oc.means) and np.allclose(c.pis, oc.pis)
    g, og = c.evaluate_grid((-4, 9), (-3, 7)), oc.evaluate_grid((-4, 9), (-3, 7))
    assert g.shape == og.shape and hp.rel_err(g, og, FLOOR * og.max()) < RATE_RTOL
    assert abs(bb.calc_bounding_radius(c.pis, c.means, c.covs, 1e-5, center=np.array([1., 2.])) -
               oracle.calc_bounding_radius(oc.pis, oc.means, oc.covs, 1e-5, center=np.array([1., 2.]))) < 1e-9
    data = rng.poisson(50., size=(40, 33)).astype(np.float64)
    model = rng.uniform(-2., 90., size=(40, 33))
    mask = (rng.uniform(size=(40, 33)) > .1).astype(np.float64)
    want = oracle.poisson_loglike(data, model, mask)
    assert abs(S.poisson_loglike(data, model, mask) - want) <= 1e-12 * abs(want)
    oracle.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    for T in (3500., 8395.):
        for band in "ugriz":
            want = oracle.photons_expected_brightness(T, 2e-10, band)
            assert abs(planck.photons_expected_brightness(T, 2e-10, band) - want) <= 1e-12 * abs(want)
            want = oracle.photons_per_joule(T, band)
            assert abs(planck.photons_per_joule(T, band) - want) <= 1e-11 * abs(want)


# ---------------------------------------------------------------------------
# N1: photon allocation
# ---------------------------------------------------------------------------
def test_binomial_sampler_distribution(engine):
    """Device binomial draws against the exact pmf (chi-square) in both regimes
    of the sampler (inversion from zero / from the mode) and at the reference's
    regression case (CelestePy/test/test_celeste_sample_sources.py:4-8)."""
    from scipy import stats
    assert np.all(engine.binomial_draws(914, 6.29379e-16, 0, 1000) == 0)
    assert np.all(engine.binomial_draws(50, 1.0, 1, 10) == 50)
    N = 200000
    for n, p, seed in [(40, .1, 3), (1000, .3, 4), (5000, .97, 5), (430, .004, 6), (100000, .45, 7)]:
        d = engine.binomial_draws(n, p, seed, N)
        assert d.min() >= 0 and d.max() <= n and np.all(d == np.round(d))
        sd = np.sqrt(n * p * (1 - p))
        assert abs(d.mean() - n * p) < 5 * sd / np.sqrt(N)
        assert abs(d.var() - sd ** 2) < 6 * sd ** 2 * np.sqrt(2. / N) + 6 * sd / np.sqrt(N)
        lo, hi = int(max(0, np.floor(n * p - 4 * sd))), int(min(n, np.ceil(n * p + 4 * sd)))
        edges = np.unique(np.linspace(lo, hi + 1, 41).astype(int))
        obs = np.histogram(d, bins=np.concatenate([[-.5], edges - .5, [n + .5]]))[0]
        cdf = stats.binom.cdf(np.concatenate([[-1], edges - 1, [n]]), n, p)
        exp = N * np.diff(cdf)
        keep = exp > 5
        chi2 = np.sum((obs[keep] - exp[keep]) ** 2 / exp[keep])
        assert chi2 < stats.chi2.ppf(1 - 1e-6, keep.sum()), (n, p, chi2)


def test_sample_source_counts_conservation_and_moments(engine, oracle):
    """Every photon goes to exactly one owner; first row/column of a patch get
    none (the reference's strict box test); per-pixel means match the
    multinomial expectation and the reference's own sampler."""
    field, gimgs, oimgs, gs, os_ = _stamp(engine, oracle, seed=1001, size=25, nseed=9)
    reps = 60
    acc = None
    for rep in range(reps):
        al = engine.resample_photons(gimgs, gs, seed=1000 + rep)
        if acc is None:
            acc = [[np.zeros_like(al.patch(s, b)) for b in range(5)] for s in range(2)]
        for b, gi in enumerate(gimgs):
            nel = gi.nelec.cpu().numpy().astype(np.float64)
            tot = al.noise_sums[b]
            for s in range(2):
                z = al.patch(s, b)
                assert np.all(z >= 0) and np.all(z == np.round(z))
                assert np.all(z[0, :] == 0) and np.all(z[:, 0] == 0)
                tot += z.sum()
                acc[s][b] += z
            assert tot == nel.sum()
    al2 = engine.resample_photons(gimgs, gs, seed=1000 + reps - 1)
    assert np.array_equal(al2.samp.cpu().numpy(), al.samp.cpu().numpy())      # seeded -> reproducible
    b = 2
    oi = oimgs[b]
    rates = [oracle.gen_src_image_with_fluxes(s, oi) for s in os_]
    boxes = np.array([[r[2][0], r[2][1], r[1][0], r[1][1]] for r in rates], dtype=np.int64)
    H, W = oi.nelec.shape
    lam = np.full((H, W), oi.epsilon)
    own = np.zeros((2, H, W), dtype=bool)
    for s, (patch, yl, xl) in enumerate(rates):
        yl, xl = (int(yl[0]), int(yl[1])), (int(xl[0]), int(xl[1]))
        own[s, yl[0] + 1:yl[1], xl[0] + 1:xl[1]] = True
        full = np.zeros((H, W)); full[yl[0]:yl[1], xl[0]:xl[1]] = patch
        lam += np.where(own[s], full, 0.)
    for s, (patch, yl, xl) in enumerate(rates):
        yl, xl = (int(yl[0]), int(yl[1])), (int(xl[0]), int(xl[1]))
        nel = oi.nelec[yl[0]:yl[1], xl[0]:xl[1]]
        share = patch / lam[yl[0]:yl[1], xl[0]:xl[1]]
        share[0, :] = 0; share[:, 0] = 0
        mean, var = nel * share, nel * share * (1 - share)
        got = acc[s][b] / reps
        z = (got - mean) / np.sqrt(var / reps + 1e-9)
        assert np.abs(z[var > 1.]).max() < 6.
        assert abs(got.sum() - mean.sum()) < 6 * np.sqrt(var.sum() / reps)
    # the reference's own sampler (legacy MT19937 stream) on the same inputs: same moments
    rs = np.random.RandomState(0)
    src_imgs = [oracle.NativePatch(r[0], r[1], r[2]) for r in rates]
    osum = [np.zeros_like(r[0]) for r in rates]
    for rep in range(12):
        samp = [oracle.NativePatch(np.zeros_like(r[0]), r[1], r[2]) for r in rates]
        samp, noise = oracle.sample_source_counts(src_imgs, samp, oi.nelec, boxes, oi.epsilon, rs)
        assert noise + sum(p.data.sum() for p in samp) == oi.nelec.sum()
        for s in range(2):
            osum[s] += samp[s].data
    for s in range(2):
        tot_ref, tot_gpu = osum[s].sum() / 12, acc[s][b].sum() / reps
        assert abs(tot_ref - tot_gpu) < 6 * np.sqrt(tot_gpu / 12 + tot_gpu / reps)


def test_compat_sample_source_counts(engine, oracle):
    from desi_mcmc_b200.compat import celeste_sample_sources as css
    rng = np.random.RandomState(4)
    nelec = rng.poisson(300., size=(20, 24)).astype(np.float64)
    boxes = np.array([[2, 14, 3, 15], [8, 22, 6, 18]])
    src = [css.NativePatch(rng.uniform(5., 60., size=(12, 12)), (3, 15), (2, 14)),
           css.NativePatch(rng.uniform(5., 60., size=(12, 14)), (6, 18), (8, 22))]
    samp = [css.NativePatch(np.zeros((12, 12)), (3, 15), (2, 14)), css.NativePatch(np.zeros((12, 14)), (6, 18), (8, 22))]
    samp, noise = css.sample_source_counts(src, samp, nelec, boxes, 40., np.random.RandomState(1))
    assert noise + samp[0].data.sum() + samp[1].data.sum() == nelec.sum()
    assert samp[0].data[1:, 1:].sum() > 0 and np.all(samp[0].data[0] == 0) and np.all(samp[1].data[:, 0] == 0)
    assert css.sample_binomial(914, 6.29379e-16, np.random.RandomState(0)) == 0


def test_fits_image_from_real_stamp(engine, oracle):
    """N4: compat.fits_image.FitsImage on the stamp the reference ships (fixture
    tests/golden/stamp_r_253.npz): same fields as the oracle's restatement of
    fits_image.py:85-155, and a star rendered with the stamp's REAL PSF."""
    import os
    from desi_mcmc_b200.compat import celeste, fits_image as fi
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "stamp_r_253.npz"))
    img, hdr = z["img"], {str(k): float(v) for k, v in zip(z["keys"], z["values"])}
    g = fi.FitsImage("r", img=img, header=hdr)
    o = oracle.fits_image_from_header("r", img, hdr)
    assert np.array_equal(g.nelec.cpu().numpy().astype(np.float64), o.nelec)
    assert g.epsilon == o.epsilon == g.epsilon0 and g.kappa == o.kappa and g.calib == o.calib
    assert np.array_equal(g.rho_n, o.rho_n) and np.array_equal(g.Ups_n, o.Ups_n)
    assert np.array_equal(g.weights, o.weights) and np.array_equal(g.covars, o.covars)
    assert abs(g.R - o.R) <= 1e-12 * o.R
    u = o.pixel2equa(np.array([24.3, 26.8]))
    patch, ylim, xlim = celeste.gen_point_source_psf_image(u, g)
    want, wyl, wxl = oracle.gen_point_source_psf_image(u, o)
    assert (tuple(ylim), tuple(xlim)) == (tuple(wyl), tuple(wxl)) and patch.shape == want.shape
    assert hp.rel_err(patch, want, 1e-7 * want.max()) < 1e-5
    src = engine.SrcParams(u, a=0, fluxes=[8., 9., 10., 11., 12.])
    osrc = oracle.SrcParams(u=u, a=0, fluxes=[8., 9., 10., 11., 12.])
    got = celeste.celeste_likelihood([src], g)
    ref_ll = oracle.celeste_likelihood([osrc], o)
    assert abs(got - ref_ll) <= 1e-6 * abs(ref_ll)
