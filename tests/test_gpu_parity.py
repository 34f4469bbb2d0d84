"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the
same seeded inputs.  Tolerances are BASELINE.json's: rendered rates 1e-5
relative, per-image log-likelihood 1e-6 relative, accept/reject identical."""
import os

import numpy as np
import pytest

from desi_mcmc_b200 import synth
from tests import helpers as hp

pytestmark = pytest.mark.gpu

RATE_RTOL = 1e-5
LL_RTOL = 1e-6
# a rate below RATE_FLOOR x (patch peak) is compared absolutely: fp32 flushes
# below 2^-126 and the float64 oracle does not
RATE_FLOOR = 1e-12


def kat_inputs(K=42, n=200):
    """CelestePy/test/test_gmm.py:63-85, same seed and draw order."""
    np.random.seed(41)
    means = 2 * np.random.randn(K * 2).reshape(K, 2)
    covs = np.zeros((K, 2, 2))
    for i in range(K):
        covs[i] = np.random.randn(2, 2)
        covs[i] = covs[i].dot(covs[i].T)
        np.linalg.inv(covs[i])
        np.linalg.slogdet(covs[i])
    ws = np.random.rand(K)
    ws /= np.sum(ws)
    xx, yy = np.meshgrid(np.linspace(-3, 3, n), np.linspace(-5, 5, n))
    X = np.column_stack((xx.ravel(), yy.ravel()))
    return X, ws, means, covs


def test_gmm_like_2d_kat(engine, oracle):
    from desi_mcmc_b200.compat import gmm_like_fast
    X, ws, means, covs = kat_inputs()
    want = oracle.gmm_like_2d(X, ws, means, covs)
    got = np.zeros(X.shape[0])
    assert gmm_like_fast.gmm_like_2d(got, X, ws, means, covs) is None
    assert np.allclose(got, want)                       # the reference test's own criterion
    assert hp.rel_err(got, want, RATE_FLOOR * want.max()) < RATE_RTOL


def test_gmm_like_2d_errors(engine):
    from desi_mcmc_b200.compat import gmm_like_fast
    X, ws, means, covs = kat_inputs(K=3, n=4)
    with pytest.raises(ValueError):
        gmm_like_fast.gmm_like_2d(np.zeros(16), X, ws[:2], means, covs)
    with pytest.raises(ValueError):
        gmm_like_fast.gmm_like_2d(np.zeros(16), X.astype(np.float32), ws, means, covs)
    p = np.ones(0)
    gmm_like_fast.gmm_like_2d(p, X[:0], ws, means, covs)    # empty input is a no-op


def test_mvn2d_logpdf(engine, oracle):
    from desi_mcmc_b200.compat import gmm_like_fast
    X, ws, means, covs = kat_inputs(K=3, n=100)
    ci = np.linalg.inv(covs[2])
    _, logdet = np.linalg.slogdet(covs[2])
    want = oracle.multivariate_normal_2d_covinv_logdet_logpdf(np.zeros(X.shape[0]), X, means[2], ci, logdet)
    got = gmm_like_fast.multivariate_normal_2d_covinv_logdet_logpdf(np.zeros(X.shape[0]), X, means[2], ci, logdet)
    assert np.allclose(got, want, rtol=1e-12, atol=1e-9)


def _check_patches(eng, ref, field, flags, unit):
    H, W = field["images"][0]["shape"]
    zeros = [np.zeros((H, W))] * 5
    gimgs = hp.gpu_images(eng, field, zeros)
    oimgs = hp.oracle_images(ref, field, zeros)
    gs, os_ = hp.gpu_sources(eng, field), hp.oracle_sources(ref, field)
    patches, boxes = eng.render_patches(gimgs, gs, default_flags=flags)
    patches = patches.cpu().numpy()
    worst = 0.0
    for si, s in enumerate(os_):
        for bi, img in enumerate(oimgs):
            it = si * 5 + bi
            if s.a == 0:
                want, ylim, xlim = ref.gen_point_source_psf_image(s.u, img)
            else:
                want, ylim, xlim = ref.gen_galaxy_psf_image(s.shape, s.u, img)
            if not unit:
                want = want * img.nmgy2counts(s.flux_dict[img.band])
            b = boxes[it]
            assert (b["x0"], b["x1"], b["y0"], b["y1"]) == (int(xlim[0]), int(xlim[1]), int(ylim[0]), int(ylim[1]))
            h, w = b["y1"] - b["y0"], b["x1"] - b["x0"]
            got = patches[it, :h * w].reshape(h, w)
            assert got.shape == want.shape
            worst = max(worst, hp.rel_err(got, want, RATE_FLOOR * want.max()))
    return worst


def test_unit_patches_stamp_c1(engine, oracle):
    from desi_mcmc_b200 import _ffi
    field = synth.make_stamp(1001)
    assert _check_patches(engine, oracle, field, _ffi.CG_ITEM_FLUX_UNIT, True) < RATE_RTOL


def test_rate_patches_small_field(engine, oracle):
    from desi_mcmc_b200 import _ffi
    field = synth.make_field(77, H=160, W=192, n_src=14, star_frac=.5)
    assert _check_patches(engine, oracle, field, _ffi.CG_ITEM_FLUX_BAND, False) < RATE_RTOL


def test_field_loglik_small(engine, oracle):
    field = synth.make_field(78, H=150, W=170, n_src=10, star_frac=.5)
    nelec = hp.simulate_nelec(oracle, field, 5)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    ll, models = engine.field_loglik(gimgs, hp.gpu_sources(engine, field), return_model=True)
    ll = ll.cpu().numpy()
    osrcs = hp.oracle_sources(oracle, field)
    for i, img in enumerate(oimgs):
        lam = oracle.gen_model_image(osrcs, img)
        assert hp.rel_err(models[i].cpu().numpy(), lam, 1.0) < RATE_RTOL
        want = oracle.celeste_likelihood(osrcs, img)
        assert abs(ll[i] - want) <= LL_RTOL * abs(want)
    total = oracle.celeste_likelihood_multi_image(osrcs, oimgs)
    assert abs(ll.sum() - total) <= LL_RTOL * abs(total)


def test_field_loglik_no_sources(engine, oracle):
    field = synth.make_field(79, H=64, W=100, n_src=0)
    rng = np.random.RandomState(3)
    nelec = [rng.poisson(400., size=(64, 100)).astype(np.float64) for _ in range(5)]
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    from desi_mcmc_b200 import _ffi
    ll = engine.field_loglik(gimgs, np.zeros(0, dtype=_ffi.SRC_DTYPE)).cpu().numpy()
    for i, img in enumerate(oimgs):
        want = oracle.celeste_likelihood([], img)
        assert abs(ll[i] - want) <= LL_RTOL * abs(want)


def _stamp_problem(eng, ref, seed, size, P, with_mask=True):
    rng = np.random.RandomState(seed)
    field = synth.make_stamp(seed, size=size, n_star=1, n_gal=1)
    nelec = hp.simulate_nelec(ref, field, seed + 1)
    invvar = None
    if with_mask:
        invvar = [(rng.uniform(size=(size, size)) > .03).astype(np.float64) for _ in range(5)]
    gimgs = hp.gpu_images(eng, field, nelec, invvar)
    oimgs = hp.oracle_images(ref, field, nelec, invvar)
    cur = field["sources"][1]            # the galaxy
    other = field["sources"][0]
    props, log_u = synth.make_proposals(seed + 2, cur, P)
    return field, gimgs, oimgs, cur, other, props, log_u


def test_score_proposals_parity(engine, oracle):
    """A15 (sources.py:275-306, :258): delta and accept decisions for the seed."""
    P, size = 96, 32
    field, gimgs, oimgs, cur, other, props, log_u = _stamp_problem(engine, oracle, 1003, size, P)
    ocur = hp.oracle_sources(oracle, [cur])[0]
    oother = hp.oracle_sources(oracle, [other])
    oprops = hp.oracle_sources(oracle, props)
    boxes, bgs, gboxes = [], [], []
    for img in oimgs:
        xlim, ylim = oracle.make_bbox(ocur, img, pixel_radius=12)
        boxes.append((xlim, ylim))
        bgs.append(oracle.generate_background_patch(ocur, oother, img, xlim, ylim))
        gboxes.append((xlim[0], xlim[1], ylim[0], ylim[1]))
    log_terms = np.random.RandomState(9).normal(0, .1, size=P)
    want = np.array([oracle.calculate_acceptance_logprob(ocur, p, oimgs, boxes, bgs, logdet=lt)
                     for p, lt in zip(oprops, log_terms)])
    scorer = engine.ProposalScorer(gimgs, gboxes, bgs, max_proposals=P, guard=1e-2)
    packed = engine.pack_sources(hp.gpu_sources(engine, [cur] + props))
    delta, accept, n_guard = scorer.score(packed, log_u, log_terms)
    print("max |delta - oracle| = %.3g, guard re-scored %d of %d" % (np.max(np.abs(delta - want)), n_guard, P))
    assert np.array_equal(accept, log_u < want)
    # fp32 rates: |delta error| is bounded well inside the guard band
    assert np.max(np.abs(delta - want)) < 2e-3
    # guard-band proposals were re-scored in float64
    near = np.abs(log_u - want) < 5e-3
    assert n_guard >= near.sum()
    if near.any():
        assert np.max(np.abs(delta[near] - want[near])) < 1e-5


def test_items_loglik_modes(engine, oracle):
    """A13 (conditional, sources.py:134-183) and A14 (isolated, :188-237)."""
    from desi_mcmc_b200 import _ffi
    field = synth.make_stamp(1001, size=25)
    nelec = hp.simulate_nelec(oracle, field, 4)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    osrcs = hp.oracle_sources(oracle, field)
    gsrcs = engine.pack_sources(hp.gpu_sources(engine, field))
    rng = np.random.RandomState(12)
    import torch
    for si, s in enumerate(osrcs):
        xl, yl = (3, 24), (2, 23)
        # isolated: data = nelec crop, background epsilon, no mask
        items = engine.make_items([si] * 5, np.arange(5), flags=_ffi.CG_ITEM_NO_MASK,
                                  box=[(xl[0], xl[1], yl[0], yl[1])] * 5)
        ll, _ = engine.items_loglik(gimgs, gsrcs, items, mode=_ffi.CG_LL_POISSON_BG)
        want = oracle.source_log_likelihood_isolated(s, [(yl, xl)] * 5, oimgs)
        assert abs(ll.sum().item() - want) <= LL_RTOL * abs(want)
        # conditional on allocated photons z
        zs = [rng.poisson(3., size=(yl[1] - yl[0], xl[1] - xl[0])).astype(np.float64) for _ in range(5)]
        want = oracle.source_log_likelihood(s, [(z, yl, xl, img) for z, img in zip(zs, oimgs)])
        aux = torch.from_numpy(np.concatenate([z.ravel() for z in zs]).astype(np.float32)).to(engine.device())
        offs = np.arange(5, dtype=np.int64) * zs[0].size
        items["aux"] = np.arange(5)
        ll, _ = engine.items_loglik(gimgs, gsrcs, items, mode=_ffi.CG_LL_CONDITIONAL, aux=aux, aux_offsets=offs)
        assert abs(ll.sum().item() - want) <= LL_RTOL * abs(want)


def test_planck_and_eigen_fluxes(engine, oracle):
    from desi_mcmc_b200 import _ffi
    oracle.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    T = np.array([3000., 5777., 8395., 12000., 0.])
    b = np.array([1e-10, 2e-10, 1., 3e-11, 1.])
    got = engine.planck_fluxes(T, b)
    for i in range(T.size):
        for j, band in enumerate(oracle.BANDS):
            want = oracle.photons_expected_brightness(T[i], b[i], band)
            assert abs(got[i, j] - want) <= 1e-12 * abs(want)
    rng = np.random.RandomState(6)
    K, V, Cn = 4, 1364, 7
    lam0 = np.linspace(600., 12000., V)
    B = np.abs(rng.normal(size=(K, V))) + .1
    B /= B.sum(axis=1, keepdims=True)
    z = rng.uniform(.5, 5., size=Cn)
    w = rng.dirichlet(np.ones(K), size=Cn)
    m = np.exp(rng.normal(10., 1., size=Cn))
    fl = np.abs(rng.normal(5., 2., size=(Cn, 5)))
    iv = rng.uniform(.5, 2., size=(Cn, 5))
    mu, ll = engine.eigen_fluxes(z, w, m, B, lam0, fl, iv)
    for c in range(Cn):
        spec = np.dot(w[c], B)
        want_mu = oracle.project_to_bands(spec, lam0 * (1 + z[c])) * m[c] / (1 + z[c])
        assert np.allclose(mu[c], want_mu, rtol=1e-11, atol=0)
        want_ll = oracle.pixel_likelihood(z[c], w[c], m[c], fl[c], iv[c], lam0, B)
        assert abs(ll[c] - want_ll) <= 1e-10 * abs(want_ll)


def test_planck_star_render(engine, oracle):
    """celeste.py:38-45: a star with a temperature takes the black-body flux."""
    from desi_mcmc_b200 import _ffi
    oracle.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    field = synth.make_field(80, H=96, W=96, n_src=6, star_frac=1.0, planck_frac=1.0, margin=20)
    nelec = hp.simulate_nelec(oracle, field, 2)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    ll = engine.field_loglik(gimgs, hp.gpu_sources(engine, field)).cpu().numpy()
    osrcs = hp.oracle_sources(oracle, field)
    for i, img in enumerate(oimgs):
        want = oracle.celeste_likelihood(osrcs, img)
        assert abs(ll[i] - want) <= LL_RTOL * abs(want)


def test_fields_batch_matches_single_calls(engine, oracle):
    """cg_fields_loglik over three unequal fields == three cg_field_loglik calls
    (bitwise: the gather sums sources in source order) and == the oracle."""
    specs = [(81, 150, 170, 9), (82, 96, 200, 0), (83, 130, 131, 14)]
    fields = [synth.make_field(sd, H=h, W=w, n_src=n, star_frac=.5) for sd, h, w, n in specs]
    gimgs, packed, single = [], [], []
    for fld in fields:
        nelec = hp.simulate_nelec(oracle, fld, fld["seed"])
        gi = hp.gpu_images(engine, fld, nelec)
        pk = engine.pack_sources(hp.gpu_sources(engine, fld))
        gimgs.append(gi)
        packed.append(pk)
        single.append(engine.field_loglik(gi, pk).cpu().numpy())
    fs = engine.FieldSet(gimgs, [max(1, p.shape[0]) for p in packed])
    got = fs.loglik(packed).copy()
    assert np.array_equal(got, np.concatenate(single))
    fld = fields[2]
    oimgs = hp.oracle_images(oracle, fld, [im.nelec.cpu().numpy().astype(np.float64) for im in gimgs[2]])
    osrcs = hp.oracle_sources(oracle, fld)
    for b, img in enumerate(oimgs):
        want = oracle.celeste_likelihood(osrcs, img)
        assert abs(got[10 + b] - want) <= LL_RTOL * abs(want)
    # the same through the pinned views + CUDA-graph replay (FieldSweep): a sampler
    # that keeps its state in the staging buffer, one driver call per step
    fs2 = engine.FieldSet(gimgs, [max(1, p.shape[0]) for p in packed])
    for f, pk in enumerate(packed):
        fs2.pinned_sources(f, pk.shape[0])[:] = pk
    fs2.upload()
    sweep = engine.FieldSweep([fs2], n_streams=1).capture()
    sweep.run()
    assert np.array_equal(fs2.ll.cpu().numpy(), got)
    v = fs2.pinned_sources(2, packed[2].shape[0])
    v["flux"] *= 1.5                                   # new parameter values, same counts: replay is valid
    fs2.upload()
    sweep.run()
    moved = fs2.ll.cpu().numpy()
    assert np.array_equal(moved[:10], got[:10]) and np.all(moved[10:] != got[10:])
    pk2 = packed[2].copy()
    pk2["flux"] *= 1.5
    assert np.array_equal(moved[10:], engine.field_loglik(gimgs[2], pk2).cpu().numpy())


def test_component_culling_error_bound(engine):
    """Per-warp component culling changes a rate by at most K * 2^-27 relative."""
    from desi_mcmc_b200 import _ffi
    field = synth.make_field(84, H=220, W=260, n_src=12, star_frac=.3)
    zeros = [np.zeros((220, 260))] * 5
    gimgs = hp.gpu_images(engine, field, zeros)
    gs = hp.gpu_sources(engine, field)
    try:
        engine.set_option(_ffi.CG_OPT_ROW_RECURRENCE, 0)      # culling alone: one exp2 per pixel on both sides
        engine.set_option(_ffi.CG_OPT_CULL, 0)
        full, boxes = engine.render_patches(gimgs, gs)
        issued_full = engine.last_issued_count()
        engine.set_option(_ffi.CG_OPT_CULL, 1)
        culled, boxes2 = engine.render_patches(gimgs, gs)
        issued_cull = engine.last_issued_count()
    finally:
        engine.set_option(_ffi.CG_OPT_CULL, 1)
        engine.set_option(_ffi.CG_OPT_ROW_RECURRENCE, 1)
    assert np.array_equal(boxes, boxes2)
    full, culled = full.cpu().numpy().astype(np.float64), culled.cpu().numpy().astype(np.float64)
    assert np.all(culled <= full * (1 + 1e-6))
    nz = full > 0
    rel = np.max((full[nz] - culled[nz]) / full[nz])
    print("culling: max relative deficit %.3g, issued %.3g -> %.3g lane-evals" % (rel, issued_full, issued_cull))
    assert rel <= 42 * 2.0 ** -27 + 2.5e-7          # + fp32 summation-order slack
    assert issued_cull < issued_full


def test_row_recurrence_error_bound(engine):
    """CG_OPT_ROW_RECURRENCE: 2 exp2 + the geometric recurrence along 4 pixels instead
    of 4 exp2.  Against the one-exp2-per-pixel path of the same kernels a rate moves
    by a few fp32 roundings: below 5e-6 relative wherever the pixel matters, below the
    north star's 1e-5 down to 1e-12 of the patch peak (the error grows with the log2
    ratio of neighbouring pixels, i.e. in the far tail of narrow components), in
    render_kernel and in the fused field kernel alike."""
    from desi_mcmc_b200 import _ffi
    field = synth.make_field(84, H=220, W=260, n_src=12, star_frac=.3)
    gimgs = hp.gpu_images(engine, field, [np.zeros((220, 260))] * 5)
    gs = hp.gpu_sources(engine, field)
    try:
        engine.set_option(_ffi.CG_OPT_ROW_RECURRENCE, 0)
        exact, boxes = engine.render_patches(gimgs, gs)
        ll0, m0 = engine.field_loglik(gimgs, gs, return_model=True)
        engine.set_option(_ffi.CG_OPT_ROW_RECURRENCE, 1)
        recur, boxes2 = engine.render_patches(gimgs, gs)
        ll1, m1 = engine.field_loglik(gimgs, gs, return_model=True)
    finally:
        engine.set_option(_ffi.CG_OPT_ROW_RECURRENCE, 1)
    assert np.array_equal(boxes, boxes2)
    exact, recur = exact.cpu().numpy().astype(np.float64), recur.cpu().numpy().astype(np.float64)
    assert np.all(np.isfinite(recur))
    assert not np.array_equal(exact, recur)               # the recurrence path did run
    # per patch, like the oracle parity tests: values below RATE_FLOOR x peak compared absolutely
    peak = exact.max(axis=1, keepdims=True)
    err = np.abs(recur - exact)
    rel_p = np.max(err / np.maximum(exact, RATE_FLOOR * np.maximum(peak, 1e-300)))
    # where a pixel holds more than 2^-27 of its patch's peak (the culling tolerance)
    rel_core = np.max(np.where(exact > peak * 2.0 ** -27, err / np.maximum(exact, 1e-300), 0.0))
    rel_m = max(hp.rel_err(b.cpu().numpy(), a.cpu().numpy(), 0.0) for a, b in zip(m0, m1))
    print("row recurrence: max relative change: patches %.3g (above 2^-27 of the peak %.3g), model image %.3g"
          % (rel_p, rel_core, rel_m))
    assert rel_p < RATE_RTOL and rel_core < 5e-6 and rel_m < 4e-6
    assert np.all(np.abs((ll1 - ll0).cpu().numpy()) <= 1e-7 * np.abs(ll0.cpu().numpy()))


def test_row_recurrence_falls_back_on_steep_components(engine, oracle):
    """A PSF far narrower than a pixel makes neighbouring pixels differ by more than
    2^96: the recurrence must not be used there (no inf/NaN), and the result still
    matches the oracle."""
    field = synth.make_stamp(87, size=48, n_star=2, n_gal=1)
    for m in field["images"]:
        m["psf_covars"] = np.asarray(m["psf_covars"]) * 0.004          # sigma ~ 0.06-0.1 px
    nelec = hp.simulate_nelec(oracle, field, 8)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    ll, models = engine.field_loglik(gimgs, hp.gpu_sources(engine, field), return_model=True)
    osrcs = hp.oracle_sources(oracle, field)
    for i, img in enumerate(oimgs):
        got = models[i].cpu().numpy()
        assert np.all(np.isfinite(got))
        assert hp.rel_err(got, oracle.gen_model_image(osrcs, img), 1.0) < RATE_RTOL
    patches, _ = engine.render_patches(gimgs, hp.gpu_sources(engine, field))
    assert np.all(np.isfinite(patches.cpu().numpy()))


@pytest.mark.parametrize("scale", [0.5, 0.15, 0.05])
def test_row_recurrence_narrow_psf_parity(engine, oracle, scale):
    """PSFs down to a fraction of a pixel wide: neighbouring pixels differ by large
    factors, the regime where the recurrence's error is largest (and, for the
    narrowest, where some warps fall back).  Rates still match the oracle to 1e-5."""
    from desi_mcmc_b200 import _ffi
    field = synth.make_stamp(88, size=40, n_star=2, n_gal=2)
    for m in field["images"]:
        m["psf_covars"] = np.asarray(m["psf_covars"]) * scale
    assert _check_patches(engine, oracle, field, _ffi.CG_ITEM_FLUX_BAND, False) < RATE_RTOL
    nelec = hp.simulate_nelec(oracle, field, 9)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    ll, models = engine.field_loglik(gimgs, hp.gpu_sources(engine, field), return_model=True)
    ll = ll.cpu().numpy()
    osrcs = hp.oracle_sources(oracle, field)
    for i, img in enumerate(oimgs):
        assert hp.rel_err(models[i].cpu().numpy(), oracle.gen_model_image(osrcs, img), 1.0) < RATE_RTOL
        want = oracle.celeste_likelihood(osrcs, img)
        assert abs(ll[i] - want) <= LL_RTOL * abs(want)


def test_field_loglik_crowded_tile_overflow(engine, oracle):
    """More sources over one 64x16 image tile than a tile list holds (48): the
    gather falls back to scanning all sources; result still matches the oracle."""
    field = synth.make_field(85, H=72, W=96, n_src=70, star_frac=.9, margin=30)
    nelec = hp.simulate_nelec(oracle, field, 6)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    ll, models = engine.field_loglik(gimgs, hp.gpu_sources(engine, field), return_model=True)
    ll = ll.cpu().numpy()
    osrcs = hp.oracle_sources(oracle, field)
    for i, img in enumerate(oimgs):
        lam = oracle.gen_model_image(osrcs, img)
        assert hp.rel_err(models[i].cpu().numpy(), lam, 1.0) < RATE_RTOL
        want = oracle.celeste_likelihood(osrcs, img)
        assert abs(ll[i] - want) <= LL_RTOL * abs(want)


def test_field_loglik_scan_rounds(engine, oracle):
    """150 sources over one tile: the overflow scan of the fused tile kernel
    takes five rounds of 32 sources; two runs are bitwise identical (sources are added in source order)."""
    from desi_mcmc_b200 import _ffi
    field = synth.make_field(86, H=48, W=80, n_src=150, star_frac=.95, margin=20)
    nelec = hp.simulate_nelec(oracle, field, 8)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    srcs = hp.gpu_sources(engine, field)
    osrcs = hp.oracle_sources(oracle, field)
    ll, models = engine.field_loglik(gimgs, srcs, return_model=True)
    ll = ll.cpu().numpy()
    ll_again = engine.field_loglik(gimgs, srcs).cpu().numpy()
    assert np.array_equal(ll, ll_again)
    # the sum of separately rendered patches (render_kernel) is the same model image
    patches, boxes = engine.render_patches(gimgs, srcs)
    patches = patches.cpu().numpy()
    n_img = len(gimgs)
    for i, gi in enumerate(gimgs):
        lam = np.full(gi.shape, np.float32(gi.epsilon), dtype=np.float32)
        for s_ in range(len(field["sources"])):
            b = boxes[s_ * n_img + i]
            h, w = int(b["y1"] - b["y0"]), int(b["x1"] - b["x0"])
            if h > 0 and w > 0:
                lam[b["y0"]:b["y1"], b["x0"]:b["x1"]] += patches[s_ * n_img + i, :h * w].reshape(h, w)
        assert hp.rel_err(models[i].cpu().numpy(), lam, 1.0) < 1e-6
    for i, img in enumerate(oimgs):
        lam = oracle.gen_model_image(osrcs, img)
        assert hp.rel_err(models[i].cpu().numpy(), lam, 1.0) < RATE_RTOL
        want = oracle.celeste_likelihood(osrcs, img)
        assert abs(ll[i] - want) <= LL_RTOL * abs(want)


def test_score_proposals_bright_source_guard(engine, oracle):
    """A bright source (> 1e6 photo-electrons in its box): the fp32 error of a score grows with the
    counts it multiplies, so the automatic guard band widens with sqrt(sum nelec^2) and the decisions
    still equal the float64 oracle's; with the band disabled the raw fp32 scores are measurably off."""
    P, size = 64, 32
    field = synth.make_stamp(1003, size=size, n_star=1, n_gal=1)
    for s in field["sources"]:
        s["fluxes"] = np.asarray(s["fluxes"]) * 400.           # ~1e6 - 1e7 electrons per band
    nelec = hp.simulate_nelec(oracle, field, 5)
    assert sum(float(n.sum()) for n in nelec) > 5e6
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    cur, other = field["sources"][1], field["sources"][0]
    props, _ = synth.make_proposals(77, cur, P)
    ocur = hp.oracle_sources(oracle, [cur])[0]
    oother = hp.oracle_sources(oracle, [other])
    oprops = hp.oracle_sources(oracle, props)
    boxes, bgs, gboxes = [], [], []
    for img in oimgs:
        xlim, ylim = oracle.make_bbox(ocur, img, pixel_radius=12)
        boxes.append((xlim, ylim))
        bgs.append(oracle.generate_background_patch(ocur, oother, img, xlim, ylim))
        gboxes.append((xlim[0], xlim[1], ylim[0], ylim[1]))
    want = np.array([oracle.calculate_acceptance_logprob(ocur, p, oimgs, boxes, bgs) for p in oprops])
    # log-uniforms placed NEAR the scores: every decision is a close call
    rng = np.random.RandomState(4)
    log_u = want + rng.normal(0., .05, size=P)
    packed = engine.pack_sources(hp.gpu_sources(engine, [cur] + props))
    auto = engine.ProposalScorer(gimgs, gboxes, bgs, max_proposals=P)
    assert auto.guard > 1e-2                                   # widened by the counts
    delta, accept, n_guard = auto.score(packed, log_u)
    assert np.array_equal(accept, log_u < want)
    assert n_guard > 0
    raw = engine.ProposalScorer(gimgs, gboxes, bgs, max_proposals=P, guard=0.0)
    d_raw, _, _ = raw.score(packed, log_u)
    err = np.max(np.abs(d_raw - want))
    print("auto guard %.3g, re-scored %d of %d, raw fp32 |delta error| %.3g" % (auto.guard, n_guard, P, err))
    assert err < auto.guard                                     # the band covers the fp32 error


def test_planck_without_filter_curves_fails_loudly(engine):
    """A black-body star rendered before cg_set_filter_curves must not come out dark: the device
    flags it and the synchronising entry points raise (the engine loads the packaged curves itself
    whenever it packs or receives a CG_SRC_PLANCK source; this drives the C ABI past that)."""
    import torch
    from desi_mcmc_b200 import _ffi
    field = synth.make_field(80, H=64, W=64, n_src=2, star_frac=1.0, planck_frac=1.0, margin=20)
    gimgs = hp.gpu_images(engine, field, [np.full((64, 64), 400.)] * 5)
    packed = engine.pack_sources(hp.gpu_sources(engine, field))
    assert np.all(packed["flags"] & _ffi.CG_SRC_PLANCK)
    _ffi.check(engine.lib().cg_shutdown())              # drops the device filter table
    engine._state.update(device=None, filters=False)
    engine.init(0)
    gimgs = hp.gpu_images(engine, field, [np.full((64, 64), 400.)] * 5)
    d_srcs = engine.to_device(packed)                    # device bytes: the engine does not inspect them
    with pytest.raises(MemoryError):
        engine.field_loglik(gimgs, d_srcs)
    ll = engine.field_loglik(gimgs, packed).cpu().numpy()       # a host array: the curves are loaded first
    assert np.all(np.isfinite(ll))


def _four_component_psf(field):
    """every image's PSF gets a fourth, wide component (CG_MAX_PSF = 4: a galaxy then has
    (6 + 8) x 4 = 56 = CG_MAX_COMP components; the reference's PSFs have 3)"""
    for m in field["images"]:
        w = np.concatenate([np.asarray(m["psf_weights"], dtype=np.float64) * .9, [.1]])
        means = np.concatenate([np.asarray(m["psf_means"], dtype=np.float64), [[.3, -.2]]])
        covs = np.concatenate([np.asarray(m["psf_covars"], dtype=np.float64), [[[9., 1.], [1., 7.]]]])
        m["psf_weights"], m["psf_means"], m["psf_covars"] = w, means, covs
    return field


def test_galaxy_on_four_component_psf(engine, oracle):
    """K = 56: the component table, the per-warp shared tables and the bin-list packing (item * 64 + K) at
    their maximum -- patches, model image and log-likelihood against the oracle."""
    from desi_mcmc_b200 import _ffi
    field = _four_component_psf(synth.make_field(79, H=120, W=140, n_src=8, star_frac=.4))
    assert _check_patches(engine, oracle, field, _ffi.CG_ITEM_FLUX_BAND, False) < RATE_RTOL
    nelec = hp.simulate_nelec(oracle, field, 6)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    ll, models = engine.field_loglik(gimgs, hp.gpu_sources(engine, field), return_model=True)
    ll = ll.cpu().numpy()
    osrcs = hp.oracle_sources(oracle, field)
    for i, img in enumerate(oimgs):
        assert hp.rel_err(models[i].cpu().numpy(), oracle.gen_model_image(osrcs, img), 1.0) < RATE_RTOL
        want = oracle.celeste_likelihood(osrcs, img)
        assert abs(ll[i] - want) <= LL_RTOL * abs(want)


def test_score_proposals_more_near_ties_than_the_old_guard_list(engine, oracle):
    """2048 proposals ALL inside the guard band (round 1 capped the re-score list at 1024 and kept the
    fp32 decision beyond it): every one is re-scored in float64 and its decision is the float64 one."""
    P, size = 2048, 24
    field, gimgs, oimgs, cur, other, props, _ = _stamp_problem(engine, oracle, 1005, size, P, with_mask=False)
    ocur = hp.oracle_sources(oracle, [cur])[0]
    oother = hp.oracle_sources(oracle, [other])
    boxes, bgs, gboxes = [], [], []
    for img in oimgs:
        xlim, ylim = oracle.make_bbox(ocur, img, pixel_radius=9)
        boxes.append((xlim, ylim))
        bgs.append(oracle.generate_background_patch(ocur, oother, img, xlim, ylim))
        gboxes.append((xlim[0], xlim[1], ylim[0], ylim[1]))
    packed = engine.pack_sources(hp.gpu_sources(engine, [cur] + props))
    raw = engine.ProposalScorer(gimgs, gboxes, bgs, max_proposals=P, guard=0.0)
    d32, _, n0 = raw.score(packed, np.zeros(P))
    assert n0 == 0
    # thresholds a hair away from the fp32 scores: near-ties, every one of them
    log_u = d32 + np.random.RandomState(4).uniform(-2e-3, 2e-3, size=P)
    scorer = engine.ProposalScorer(gimgs, gboxes, bgs, max_proposals=P, guard=1e-2)
    delta, accept, n_guard = scorer.score(packed, log_u)
    assert n_guard == P
    assert np.array_equal(accept, log_u < delta)
    pick = np.random.RandomState(5).choice(P, 48, replace=False)      # the oracle takes ~10 ms per proposal
    oprops = hp.oracle_sources(oracle, [props[k] for k in pick])
    want = np.array([oracle.calculate_acceptance_logprob(ocur, p, oimgs, boxes, bgs) for p in oprops])
    # float64 tables / exp / log on the device; what is left is the float32 rounding of the background patches
    assert np.max(np.abs(delta[pick] - want)) < 5e-5
    clear = np.abs(log_u[pick] - want) > 5e-5
    assert np.array_equal(accept[pick][clear], (log_u[pick] < want)[clear])
    assert np.max(np.abs(d32 - delta)) > 1e-6                          # the fp32 scores did differ
