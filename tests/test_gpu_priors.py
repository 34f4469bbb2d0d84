"""N3: empirical-prior mixtures (util/dists/mog.py:5-21,
flux_prior.py:4-93) and their sufficient statistics on the device, float64,
against the oracle restatement."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from desi_mcmc_b200 import engine as eng
    eng.init(0)
    return eng


@pytest.fixture(scope="module")
def oracle():
    from oracle import ref_cpu
    return ref_cpu


def _mog(rng, K, D):
    means = rng.normal(0, 1.5, size=(K, D))
    A = rng.normal(0, 1, size=(K, D, D))
    covs = np.einsum('kij,klj->kil', A, A) + .3 * np.eye(D)[None]
    pis = rng.dirichlet(np.ones(K) * 2)
    return means, covs, pis


@pytest.mark.parametrize("K,D", [(8, 5), (50, 5), (4, 3), (3, 1), (40, 2)])
def test_mog_logpdf_matches_oracle(engine, oracle, K, D):
    rng = np.random.RandomState(100 + K + D)
    means, covs, pis = _mog(rng, K, D)
    icovs, dets = oracle.mog_params(covs)
    x = rng.normal(0, 2.5, size=(1500, D))
    x[0] = 60.                                               # far tail: every component underflows together
    want = oracle.mog_loglike(x, means, icovs, dets, pis)
    got, resp = engine.mog_logpdf(x, means, icovs, dets, pis, return_resp=True)
    assert np.max(np.abs(got - want) / np.maximum(1., np.abs(want))) < 1e-13
    r = resp.cpu().numpy()
    assert np.allclose(r.sum(axis=1), 1., atol=1e-12)
    assert np.max(np.abs(r - oracle.mog_responsibilities(x, means, covs, pis))) < 1e-12
    assert engine.mog_logpdf(x[3], means, icovs, dets, pis) == got[3]      # 1-D input -> scalar (:18-19)


def test_flux_prior_classes(engine, oracle):
    from desi_mcmc_b200.compat import flux_prior as fp
    rng = np.random.RandomState(7)
    means, covs, pis = _mog(rng, 8, 5)
    prior = fp.FluxColorMoG(means, covs, pis)
    lf = np.log(rng.lognormal(2., .7, size=(300, 5)))
    want = oracle.flux_color_logpdf(lf, means, covs, pis)
    assert np.max(np.abs(prior.logpdf(lf) - want)) < 1e-11
    assert np.allclose(fp.FluxColorMoG.to_fluxes(fp.FluxColorMoG.to_colors(lf)), lf)
    c = fp.FluxColorMoG.to_colors(lf[:1])[0]
    assert np.allclose(c, [lf[0, 0] - lf[0, 2], lf[0, 1] - lf[0, 2], lf[0, 3] - lf[0, 2], lf[0, 4] - lf[0, 2],
                           lf[0, 2]])                         # [lu-lr, lg-lr, li-lr, lz-lr, lr] (:17-19)
    m3, c3, p3 = _mog(rng, 4, 3)
    shape = fp.GalShapeMoG(m3, c3, p3)
    xs = np.array([2.3, .6, 1.1])
    ic, dt = oracle.mog_params(c3)
    assert abs(shape.logpdf(xs) - oracle.mog_loglike(oracle.gal_shape_unconstrained(xs), m3, ic, dt, p3)) < 1e-12
    m1, c1, p1 = _mog(rng, 3, 1)
    ic, dt = oracle.mog_params(c1)
    rad, ab = fp.GalRadiusMoG(m1, c1, p1), fp.GalAbMoG(m1, c1, p1)
    assert abs(rad.logpdf(1.7) - oracle.mog_loglike(np.array([np.log(1.7)]), m1, ic, dt, p1)) < 1e-12
    assert abs(ab.logpdf(.4) - oracle.mog_loglike(np.array([np.log(.4) - np.log(.6)]), m1, ic, dt, p1)) < 1e-12


def test_prior_sufficient_stats_and_sweep_vector(engine, oracle):
    from desi_mcmc_b200 import dist
    from desi_mcmc_b200.compat import flux_prior as fp
    rng = np.random.RandomState(9)
    K = 8
    means, covs, pis = _mog(rng, K, 5)
    prior = fp.FluxColorMoG(means, covs, pis)
    lf = np.log(rng.lognormal(2., .7, size=(777, 5)))
    colors = fp.FluxColorMoG.to_colors(lf)
    logp, resp = prior.responsibilities(colors)
    block = engine.prior_sufficient_stats(resp, colors)
    want = oracle.prior_sufficient_stats(oracle.mog_responsibilities(colors, means, covs, pis), colors)
    got = block.cpu().numpy()
    assert got.shape == (K * 21,) and np.max(np.abs(got - want) / np.maximum(1., np.abs(want))) < 1e-12
    assert np.array_equal(engine.prior_sufficient_stats(resp, colors).cpu().numpy(), got)    # fixed-shape tree
    st = dist.SweepStats(n_prior_comp=K, device=engine.device())
    st.add_prior_stats_device(1, block)
    st.add_prior_stats(0, resp.cpu().numpy(), colors)         # the host route fills the same layout
    s0, s1, s2 = st.prior_stats(1)
    h0, h1, h2 = st.prior_stats(0)
    assert np.allclose(s0, h0, rtol=1e-12) and np.allclose(s1, h1, rtol=1e-12) and np.allclose(s2, h2, rtol=1e-12)
    assert abs(s0.sum() - 777.) < 1e-9
