"""The live Gibbs sweep through the reference's call surface:
models.Celeste / Field.resample_photons / Source.resample and the
celeste_mcmc conditionals, on the CUDA path, checked against the oracle."""
import numpy as np
import pytest

from desi_mcmc_b200 import synth
from tests import helpers as hp

pytestmark = pytest.mark.gpu

LL_RTOL = 1e-6      # per-image log-likelihood (north_star)
RATE_RTOL = 1e-5    # rendered rates


@pytest.fixture(scope="module")
def engine():
    from desi_mcmc_b200 import engine as eng
    eng.init(0)
    return eng


@pytest.fixture(scope="module")
def oracle():
    from oracle import ref_cpu
    return ref_cpu


def _model(engine, oracle, seed=1001, size=25):
    from desi_mcmc_b200.compat import models
    field = synth.make_stamp(seed, size=size)          # BASELINE config 1: 1 star + 1 galaxy
    nelec = hp.simulate_nelec(oracle, field, 7)
    gimgs = hp.gpu_images(engine, field, nelec)
    m = models.Celeste()
    m.add_field({im.band: im for im in gimgs})
    m.initialize_sources(init_src_params=hp.gpu_sources(engine, field))
    return m, field, nelec, gimgs


def test_model_image_and_loglik_match_oracle(engine, oracle):
    m, field, nelec, gimgs = _model(engine, oracle)
    oimgs = hp.oracle_images(oracle, field, nelec)
    osrcs = hp.oracle_sources(oracle, field)
    for gi, oi in zip(gimgs, oimgs):
        oi.epsilon = gi.epsilon                        # Field.__init__ set epsilon = median(nelec) (:117-118)
        assert gi.epsilon == float(np.median(oi.nelec))
        lam = m.render_model_image(gi)
        want = oracle.gen_model_image(osrcs, oi)
        assert hp.rel_err(lam, want, 1.0) < RATE_RTOL
        ll = m.img_log_likelihood(gi)
        want_ll = oracle.celeste_likelihood(osrcs, oi)
        assert abs(ll - want_ll) <= LL_RTOL * abs(want_ll)
        ll2 = m.img_log_likelihood(gi, mod_img=lam)    # the reference's second route (:104-108)
        assert abs(ll2 - want_ll) <= 1e-5 * abs(want_ll)
        crop = m.render_model_image(gi, xlim=(3, 17), ylim=(2, 11))
        assert crop.shape == (9, 14) and np.array_equal(crop, lam[2:11, 3:17])


def test_gibbs_sweep_conserves_photons_and_moves_the_chain(engine, oracle):
    m, field, nelec, gimgs = _model(engine, oracle)
    np.random.seed(11)
    rs = np.random.RandomState(5)
    noise = m.field_list[0].resample_photons(m.srcs, random_state=rs)
    for b, im in m.field_list[0].img_dict.items():
        total = float(im.nelec.sum().item())
        got = noise[b]
        for s in m.srcs:
            for samp_img, fimg, grid in s.sample_image_list:
                if fimg is im:
                    assert grid.shape == (samp_img.data.size, 2)
                    got += float(np.sum(samp_img.data))
        assert got == total                            # every photon is allocated exactly once (:83-122)
        assert im.epsilon > 0
    before = [(s.params.u.copy(), np.array(s.params.fluxes, dtype=float).copy()) for s in m.srcs]
    ll0 = [s.log_likelihood() for s in m.srcs]
    m.resample_sources()
    for s, (u0, f0), l0 in zip(m.srcs, before, ll0):
        assert np.all(np.isfinite(s.params.u)) and np.all(np.isfinite(s.params.fluxes))
        assert np.all(np.asarray(s.params.fluxes) > 0)
        assert np.all(np.abs(s.params.u - u0) <= .0025 + 1e-12)      # slice bounds u_lower/u_upper (:30-32)
        assert not np.array_equal(np.asarray(s.params.fluxes, dtype=float), f0)
        assert np.isfinite(s.log_likelihood())
    m.resample_model()                                 # a second full sweep runs from the new state


def test_celeste_mcmc_conditionals_match_oracle(engine, oracle):
    from desi_mcmc_b200.compat import celeste_mcmc as cm, planck
    m, field, nelec, gimgs = _model(engine, oracle)
    oimgs = hp.oracle_images(oracle, field, nelec)
    for gi, oi in zip(gimgs, oimgs):
        oi.epsilon = gi.epsilon
    osrcs = hp.oracle_sources(oracle, field)
    # photon allocation of one image through the reference's entry point (:98-150)
    samp_imgs, noise_sum = cm.sample_source_photons_single_image_cython(
        gimgs[2], [s.params for s in m.srcs], random_state=np.random.RandomState(3))
    assert noise_sum + sum(float(np.sum(p.data)) for p in samp_imgs) == float(nelec[2].sum())
    for p, os_ in zip(samp_imgs, osrcs):
        patch, ylim, xlim = oracle.gen_src_image_with_fluxes(os_, oimgs[2])
        assert (p.y0, p.y1, p.x0, p.x1) == (ylim[0], ylim[1], xlim[0], xlim[1])
    # location conditional of the star given its photons (:393-404)
    star = [s for s in m.srcs if s.is_star()][0].params
    ostar = [s for s in osrcs if s.a == 0][0]
    samp_all, _ = cm.sample_source_photons_field(gimgs, [s.params for s in m.srcs],
                                                 random_state=np.random.RandomState(4))
    s_idx = [i for i, s in enumerate(m.srcs) if s.is_star()][0]
    z = [samp_all[n][s_idx] for n in range(len(gimgs))]
    for du in (0., 2e-5):
        u = np.array(star.u) + du
        got = cm.loc_like(u, star, gimgs, z)
        want = 0.
        ostar.u = u
        for n, oi in enumerate(oimgs):
            xlim, ylim = (z[n].x0, z[n].x1), (z[n].y0, z[n].y1)
            f, _, _ = oracle.gen_src_image_with_fluxes(ostar, oi, xlim=xlim, ylim=ylim)
            mask = (z[n].data > 0) & (f > 0)
            want += np.sum(np.log(f[mask]) * z[n].data[mask]) - f.sum()
        assert abs(got - want) <= 1e-6 * abs(want)
    # black-body conditional (:374-390)
    import os
    from desi_mcmc_b200 import _ffi
    oracle.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    fs = np.array([.9, .95, 1., .97, .8])
    sums = np.array([300., 800., 1200., 900., 400.])
    for th in ([5500., .3], [9000., .05]):
        got = cm.temp_bright_like(np.array(th), fs, sums, gimgs)
        want = 0.
        for i, im in enumerate(gimgs):
            e = fs[i] * oracle.photons_expected_brightness(th[0], th[1], im.band)
            if e > 0:
                want += sums[i] * np.log(e) - e
        assert abs(got - want) <= 1e-9 * abs(want)
    assert cm.temp_bright_like(np.array([10., .3]), fs, sums, gimgs) == -np.inf
    assert cm.unif_lnpdf(.5, 0., 1.) == 0. and cm.unif_lnpdf(1., 0., 1.) == -np.inf


def test_hmc_galaxy_step_runs_on_device_gradients(engine, oracle):
    """One short HMC trajectory over (shape, location, fluxes) of the galaxy
    (celeste_mcmc.py:273-319); the finite-difference gradient batch agrees with
    the oracle's gradient at the start point."""
    from desi_mcmc_b200.compat import celeste_mcmc as cm, celeste_galaxy_conditionals as gal
    m, field, nelec, gimgs = _model(engine, oracle, size=31)
    oimgs = hp.oracle_images(oracle, field, nelec)
    g = [s for s in m.srcs if s.is_galaxy()][0].params
    th = np.concatenate(([g.theta, g.sigma, g.phi, g.rho], g.u, np.asarray(g.fluxes, dtype=float)))
    # photon images shaped like the Cython-route footprints
    Z = []
    rng = np.random.RandomState(2)
    for gi, oi in zip(gimgs, oimgs):
        R = oracle.gen_galaxy_transformation(g.sigma, g.rho, g.phi, oi.Ups_n)
        f, _, _ = oracle.gen_galaxy_prof_psf_image('exp', R, g.u, oi)
        Z.append(rng.poisson(50. * f / f.max()).astype(np.float64))
    grad = gal.galaxy_source_like_grad(th, Z, gimgs)
    want = oracle.galaxy_source_like_grad(th, Z, oimgs)
    scale = np.maximum(np.abs(want), 1e-3 * np.max(np.abs(want)))
    assert np.all(np.abs(grad - want) / scale < 1e-4)       # float64 on both sides; fp32 would be off by +-500
    np.random.seed(3)
    ll0 = gal.galaxy_source_like(th, Z, gimgs)
    p = cm.hmc_sample_galaxy_params(g, Z, gimgs, step_sz=1e-9, steps_per_sample=3)   # u has mass 1e-3 and d ll/du ~ 4e7
    assert p.shape == (11,) and np.all(np.isfinite(p))
    th1 = np.concatenate(([g.theta, g.sigma, g.phi, g.rho], g.u, [g.fluxes[b] for b in "ugriz"]))
    assert np.all(np.isfinite(th1)) and np.isfinite(gal.galaxy_source_like(th1, Z, gimgs)) and np.isfinite(ll0)


def test_batched_resample_equals_per_source_resample(engine, oracle):
    """All sources advanced in lock-step (one launch per slice round) end in the
    state the reference's per-source loop reaches with the same per-source RNG."""
    from desi_mcmc_b200.compat import models
    from desi_mcmc_b200.compat.slicesample import run_lockstep

    def build():
        field = synth.make_field(91, H=96, W=128, n_src=12, star_frac=.5, margin=12)
        nelec = hp.simulate_nelec(oracle, field, 9)
        gimgs = hp.gpu_images(engine, field, nelec)
        m = models.Celeste()
        m.add_field({im.band: im for im in gimgs})
        m.initialize_sources(init_src_params=hp.gpu_sources(engine, field))
        m.field_list[0].resample_photons(m.srcs, random_state=np.random.RandomState(21))
        return m

    a, b = build(), build()
    for s, src in enumerate(a.srcs):
        src.resample(rng=np.random.RandomState(100 + s))
    engine.launch_count(reset=True)
    rounds = b.resample_sources_batched(seed=100)
    launches = engine.launch_count()
    for sa, sb in zip(a.srcs, b.srcs):
        assert np.allclose(sa.params.fluxes, sb.params.fluxes, rtol=1e-12, atol=0)
        assert np.allclose(sa.params.u, sb.params.u, rtol=0, atol=1e-13)
        assert not np.array_equal(sb.params.u, sb.u_lower + .0025)     # the chain moved
    # 12 chains shared their rounds: far fewer launches than 12 separate samplers need
    assert rounds == run_lockstep.last_rounds and launches <= 4 * (rounds + 2)
    print("lock-step rounds", rounds, "kernel launches", launches)
