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


def test_on_device_hmc_equals_oracle_driven_hmc(engine, oracle):
    """N2: cg_hmc_galaxy_params (the whole trajectory of hmc_sample_galaxy_params, celeste_mcmc.py:281-323, in
    one launch) against util/infer/hmc.py's control flow driven by the ORACLE's float64 galaxy_source_like /
    galaxy_source_like_grad (celeste_galaxy_conditionals.py:15-88) from the same numpy.random stream: same
    energies, same decision, same end state and momentum."""
    from desi_mcmc_b200.compat import celeste_mcmc as cm
    from desi_mcmc_b200.compat.hmc import hmc
    m, field, nelec, gimgs = _model(engine, oracle, size=31)
    oimgs = hp.oracle_images(oracle, field, nelec)
    g = [s for s in m.srcs if s.is_galaxy()][0].params
    th0 = np.concatenate(([g.theta, g.sigma, g.phi, g.rho], g.u, np.asarray(g.fluxes, dtype=float)))
    Z, rng = [], np.random.RandomState(2)
    for oi in oimgs:
        R = oracle.gen_galaxy_transformation(g.sigma, g.rho, g.phi, oi.Ups_n)
        f, _, _ = oracle.gen_galaxy_prof_psf_image('exp', R, g.u, oi)
        Z.append(rng.poisson(50. * f / f.max()).astype(np.float64))
    mass, step, n_steps = engine.HMC_GALAXY_MASS, 2e-9, 4
    # the oracle-driven chain: initial momentum, refresh normals and the acceptance uniform from numpy.random
    np.random.seed(11)
    p0 = np.sqrt(mass) * np.random.randn(11)
    x_o, p_o, _, _ = hmc(x_curr=th0.copy(), llhfunc=lambda t: oracle.galaxy_source_like(t, Z, oimgs),
                         grad_llhfunc=lambda t: oracle.galaxy_source_like_grad(t, Z, oimgs), eps=step,
                         num_steps=n_steps, mass=mass, p_curr=p0, refresh_alpha=.9)
    np.random.seed(11)
    p_d, acc = cm.hmc_sample_galaxy_params_on_device([g], [Z], gimgs, step_sz=step, steps_per_sample=n_steps)
    x_d = np.concatenate(([g.theta, g.sigma, g.phi, g.rho], g.u, [g.fluxes[b] for b in "ugriz"]))
    h = cm.hmc_sample_galaxy_params_on_device.last["h"][0]
    want_h0 = oracle.galaxy_source_like(th0, Z, oimgs) - .5 * np.sum(p0 * p0 / mass)
    assert abs(h[0] - want_h0) <= 1e-9 * abs(want_h0)
    assert bool(acc[0]) == (not np.array_equal(x_o, th0))
    moved_o, moved_d = x_o - th0, x_d - th0
    scale = np.maximum(np.abs(moved_o), 1e-6 * np.max(np.abs(moved_o)))
    assert np.any(moved_o != 0.) and np.all(np.abs(moved_d - moved_o) / scale < 1e-4), (moved_d, moved_o)
    pscale = np.maximum(np.abs(p_o), 1e-6 * np.max(np.abs(p_o)))
    assert np.all(np.abs(p_d[0] - p_o) / pscale < 1e-4), (p_d[0], p_o)
    print("device HMC: %.2f ms for %d leapfrog steps, accepted %s" %
          (cm.hmc_sample_galaxy_params_on_device.last["kernel_ms"], n_steps, bool(acc[0])))


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


def _oracle_chain_step(oracle, osrc, olist, oimgs, rng):
    """One Source.resample() (sources.py:242-245) driven entirely by the oracle's float64 likelihoods:
    resample_fluxes (:327-350, unit patches on the natural footprints as compat does) then
    resample_location (:308-319) with the reference sampler's control flow."""
    from desi_mcmc_b200.compat.slicesample import slicesample
    a_0, b_0 = 5., .005
    counts = np.zeros(5)
    psf = np.zeros(5)
    for (z, yl, xl, oi) in olist:
        b = "ugriz".index(oi.band)
        counts[b] += np.sum(z)
        p, _, _ = oracle.compute_scatter_on_pixels(osrc, oi)
        psf[b] += np.sum(p) * oi.kappa / oi.calib
    osrc.fluxes = rng.gamma(a_0 + counts, 1. / (b_0 + psf))
    n_eval = [0]

    def ll(u):
        n_eval[0] += 1
        return oracle.source_log_likelihood(osrc, olist, u=u)
    u, _ = slicesample(osrc.u.copy(), ll, step=None, step_out=False, upper_bound=osrc.u + 1.,
                       lower_bound=osrc.u - 1., rng=rng)
    osrc.u = u
    return n_eval[0]


def test_multi_sweep_trace_equals_oracle_driven_chain(engine, oracle):
    """north_star: identical chain traces for the stated seeds.  Config 1 (star + galaxy, 25x25x5), photon
    patches held fixed (the allocation is distributional, N1), 12 sweeps of resample_fluxes +
    resample_location per source from RandomState(1001 + source): the device chain (per source, and all
    sources in lock-step) against a chain driven by the oracle's float64 likelihoods.  Every slice-sampler
    decision agrees BY CONSTRUCTION (near-ties are re-decided in float64 on the device, compat.slicesample
    guard band), so the location traces -- functions of the uniforms alone once the decisions agree --
    are equal to rounding, and the gamma-conditional fluxes to 1e-6."""
    from desi_mcmc_b200.compat import sources as S
    from desi_mcmc_b200.compat.celeste_sample_sources import NativePatch
    from tests.test_gpu_compat import _Model, _stamp
    field, gimgs, oimgs, gs, os_ = _stamp(engine, oracle, seed=1001, size=25, nseed=8)
    prng = np.random.RandomState(5)
    n_sweeps = 12
    olists, chains = [], {"dev": [], "lock": []}
    for si in (0, 1):
        olist = []
        for oi in oimgs:
            p, yl, xl = oracle.gen_src_image_with_fluxes(os_[si], oi)
            olist.append((prng.poisson(p).astype(np.float64), (int(yl[0]), int(yl[1])), (int(xl[0]), int(xl[1])), oi))
        olists.append(olist)
    for name in chains:
        for si in (0, 1):
            src = S.Source(hp.gpu_sources(engine, field)[si], _Model())
            src.sample_image_list = [(NativePatch(z, yl, xl), gi, None) for (z, yl, xl, _), gi in zip(olists[si], gimgs)]
            chains[name].append(src)
    osrcs = hp.oracle_sources(oracle, field)
    rng_o = [np.random.RandomState(1001 + si) for si in (0, 1)]
    rng_d = [np.random.RandomState(1001 + si) for si in (0, 1)]
    trace_o, trace_d, n_eval, n_ref = [], [], 0, 0
    for k in range(n_sweeps):
        for si in (0, 1):
            n_eval += _oracle_chain_step(oracle, osrcs[si], olists[si], oimgs, rng_o[si])
            chains["dev"][si].resample(rng=rng_d[si])
            n_ref += chains["dev"][si].n_refined[0]
        trace_o.append(np.concatenate([np.concatenate([o.u, o.fluxes]) for o in osrcs]))
        trace_d.append(np.concatenate([np.concatenate([s.params.u, s.params.fluxes]) for s in chains["dev"]]))
    trace_o, trace_d = np.array(trace_o), np.array(trace_d)
    cols_u = [0, 1, 7, 8]
    cols_f = [c for c in range(14) if c not in cols_u]
    assert np.allclose(trace_d[:, cols_u], trace_o[:, cols_u], rtol=0, atol=1e-12)
    assert np.allclose(trace_d[:, cols_f], trace_o[:, cols_f], rtol=1e-6, atol=0)
    assert not np.allclose(trace_o[0, cols_u], trace_o[-1, cols_u], rtol=0, atol=1e-9)      # the chains moved
    assert n_eval > 400                        # hundreds of slice decisions, each one agreed
    # ... and the lock-step driver (one launch per round for all sources) walks the same trace
    trace_l = []
    for k in range(n_sweeps):
        # resample_sources_batched seeds RandomState(seed + s) afresh: give it the chain's streams instead
        from desi_mcmc_b200.compat.slicesample import run_lockstep, slicesample_gen
        bc = S.BatchedConditional(chains["lock"])
        sums = bc.unit_patch_sums()
        if k == 0:
            rng_l = [np.random.RandomState(1001 + si) for si in (0, 1)]
        for src, rng, ps in zip(chains["lock"], rng_l, sums):
            src.resample_fluxes(rng=rng, psf_sums_by_image=ps)
        bc.refresh()
        gens = [slicesample_gen(src.params.u.copy(), **src._location_slice_args(rng)) for src, rng in zip(chains["lock"], rng_l)]
        for src, (u, ll) in zip(chains["lock"], run_lockstep(gens, bc.loglik)):
            src.params.u = u
        trace_l.append(np.concatenate([np.concatenate([s.params.u, s.params.fluxes]) for s in chains["lock"]]))
    trace_l = np.array(trace_l)
    assert np.allclose(trace_l[:, cols_u], trace_d[:, cols_u], rtol=0, atol=1e-13)
    assert np.allclose(trace_l[:, cols_f], trace_d[:, cols_f], rtol=1e-12, atol=0)
    print("oracle evaluations", n_eval, "float64 refinements on the device chain", n_ref)


def test_on_device_slice_sampler_equals_lockstep(engine, oracle):
    """N2: cg_slice_sample_locations runs every source's slice sampler in ONE launch (the sampler's
    control flow and its likelihood evaluations inside the kernel, host-supplied uniforms in the
    reference's draw order).  Three sweeps with persistent per-source generators: locations, fluxes
    and the generators' states equal those of the lock-step host driver bit for bit; the location
    phase costs one launch however many shrink rounds the chains need."""
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
    rng_a = [np.random.RandomState(100 + s) for s in range(len(a.srcs))]
    rng_b = [np.random.RandomState(100 + s) for s in range(len(b.srcs))]
    from desi_mcmc_b200.compat import sources as S
    from desi_mcmc_b200.compat.slicesample import slicesample_gen
    tot_dev, tot_evals = 0, 0
    for sweep in range(3):
        # host lock-step driver with the persistent generators (resample_sources_batched's body)
        bc = S.BatchedConditional(a.srcs)
        for src, rng, ps in zip(a.srcs, rng_a, bc.unit_patch_sums()):
            src.resample_fluxes(rng=rng, psf_sums_by_image=ps)
        bc.refresh()
        gens = [slicesample_gen(src.params.u.copy(), **src._location_slice_args(rng)) for src, rng in zip(a.srcs, rng_a)]
        for src, (u, ll) in zip(a.srcs, run_lockstep(gens, bc.loglik)):
            src.params.u = u
        rounds = run_lockstep.last_rounds
        engine.launch_count(reset=True)
        st = b.resample_sources_on_device(rngs=rng_b)
        launches = engine.launch_count()
        tot_dev += st["device"]
        tot_evals += st["evals"]
        for sa, sb, ra, rb in zip(a.srcs, b.srcs, rng_a, rng_b):
            assert np.array_equal(sa.params.u, sb.params.u)
            assert np.array_equal(sa.params.fluxes, sb.params.fluxes)
            assert ra.random_sample() == rb.random_sample()          # the streams continue in step
        # unit-patch render (2 launches) + ONE sampler launch + the lock-step rounds of any fallback
        assert st["fallback"] > 0 or launches <= 4
        assert rounds >= 10
    assert tot_dev >= 30 and tot_evals > 300
    print("device-sampled", tot_dev, "of", 3 * len(a.srcs), "likelihood evaluations in the kernel", tot_evals)
