"""BASELINE config 4 -- the crowded-field Gibbs sweep on the device (engine.GibbsSweep) against
the oracle: eigen-spectrum / black-body fluxes feeding the render without a host hop, photon
allocation on the bin lists, gamma conditionals, batched proposals for every source scored by
the conditional (sources.py:134-183) and the full MH ratio (sources.py:275-306) with
backgrounds (sources.py:476-483).  Tolerances: rates 1e-5, log-likelihoods 1e-6 (north star)."""
import os

import numpy as np
import pytest

from desi_mcmc_b200 import synth
from tests import helpers as hp

pytestmark = pytest.mark.gpu
RATE_RTOL, LL_RTOL, FLOOR = 1e-5, 1e-6, 1e-6


def _oracle_fluxes(oracle, s, metas, B, lam0):
    """band fluxes (nanomaggies) the oracle gives a source of the crowded field"""
    if s.get("eigen"):
        e = s["eigen"]
        return oracle.project_to_bands(np.dot(e["w"], B), lam0 * (1. + e["z"])) * e["m"] / (1. + e["z"])
    if s["a"] == 0 and s["t"]:
        return np.array([oracle.photons_expected_brightness(s["t"], s["b"], m["band"]) * m["calib"] / m["gain"]
                         for m in metas])
    return np.asarray(s["fluxes"], dtype=np.float64)


def _setup(engine, oracle, seed=1004, H=200, W=240, n_src=90, nelec_level=None, **kw):
    from desi_mcmc_b200 import _ffi
    oracle.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    field = synth.make_crowded_field(seed, H=H, W=W, n_src=n_src)
    B, lam0 = synth.make_eigen_basis(seed)
    metas = field["images"]
    osrcs = []
    for s in field["sources"]:
        f = _oracle_fluxes(oracle, s, metas, B, lam0)
        osrcs.append(oracle.SrcParams(u=s["u"], a=s["a"], theta=s["theta"], phi=s["phi"], sigma=s["sigma"],
                                      rho=s["rho"], fluxes=f))
    rng = np.random.RandomState(seed + 1)
    if nelec_level is None:
        nelecs = []
        for m in metas:
            im = oracle.SyntheticImage(nelec=np.zeros((H, W)), **synth.image_kwargs(m))
            lam = np.full((H, W), im.epsilon)
            for o in osrcs:
                p, yl, xl = _capped_patch(oracle, o, im, kw.get("bound_cap", 25.))
                if p is not None:
                    lam[yl[0]:yl[1], xl[0]:xl[1]] += p
            nelecs.append(rng.poisson(lam).astype(np.float64))
    else:
        nelecs = [np.full((H, W), float(nelec_level)) for _ in metas]
    gimgs = hp.gpu_images(engine, field, nelecs)
    oimgs = hp.oracle_images(oracle, field, nelecs)
    gsrcs = []
    for s in field["sources"]:
        e = s.get("eigen")
        gsrcs.append(engine.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                                      sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"],
                                      flux_row=e["row"] if e else None))
    packed = engine.pack_sources(gsrcs)
    es = [s["eigen"] for s in field["sources"] if s.get("eigen")]
    eig = engine.EigenSpectra([e["z"] for e in es], [e["w"] for e in es], [e["m"] for e in es], B, lam0) if es else None
    sw = engine.GibbsSweep(gimgs, packed, eigen=eig, seed=seed, **kw)
    return field, gimgs, oimgs, osrcs, packed, sw


def _capped_box(oracle, src, img, cap):
    """the footprint prepare_kernel gives a source with bound_cap (celeste.py:139-140 /
    celeste_galaxy_conditionals.py:208-211 with the radius capped)"""
    H, W = img.nelec.shape
    px, py = img.equa2pixel(src.u)
    if src.a == 0:
        bound = min(img.R, cap)
        return (max(0, int(px - bound)), min(int(px + bound + 1), W), max(0, int(py - bound)), min(int(py + bound + 1), H))
    bound = min(oracle.gen_galaxy_psf_image_bound(src.shape, src.u, img), cap)
    return (int(max(0, np.floor(px - bound))), int(min(W, np.ceil(px + bound))),
            int(max(0, np.floor(py - bound))), int(min(H, np.ceil(py + bound))))


def _capped_patch(oracle, src, img, cap):
    x0, x1, y0, y1 = _capped_box(oracle, src, img, cap)
    if x1 <= x0 or y1 <= y0:
        return None, None, None
    p, _, _ = oracle.compute_model_patch(src, img, xlim=(x0, x1), ylim=(y0, y1))
    return p, (y0, y1), (x0, x1)


def test_eigen_and_planck_fluxes_feed_the_render_on_device(engine, oracle):
    """CG_SRC_EIGEN: the band fluxes projected by eigen_flux_kernel are read from HBM by
    prepare_kernel; CG_SRC_PLANCK stars integrate the black body inside it.  The rendered rate
    patches equal the oracle's with host-computed fluxes."""
    field, gimgs, oimgs, osrcs, packed, sw = _setup(engine, oracle, n_src=60)
    mu = sw.eigen.project().cpu().numpy()
    es = [(k, s["eigen"]) for k, s in enumerate(field["sources"]) if s.get("eigen")]
    assert len(es) >= 2
    for k, e in es:
        assert np.allclose(mu[e["row"]], osrcs[k].fluxes, rtol=1e-11, atol=0)
    patches, boxes = engine.render_patches(gimgs, packed, bound_cap=25.)
    n_checked = 0
    for k, (s, o) in enumerate(zip(field["sources"], osrcs)):
        if not (s.get("eigen") or (s["a"] == 0 and s["t"])):
            continue
        for b, oi in enumerate(oimgs):
            want, yl, xl = _capped_patch(oracle, o, oi, 25.)
            bx = boxes[k * 5 + b]
            if want is None:
                assert bx["x1"] <= bx["x0"] or bx["y1"] <= bx["y0"]
                continue
            assert (bx["x0"], bx["x1"], bx["y0"], bx["y1"]) == (xl[0], xl[1], yl[0], yl[1])
            got = patches[k * 5 + b, :want.size].cpu().numpy().reshape(want.shape)
            assert hp.rel_err(got, want, FLOOR * want.max()) < RATE_RTOL
            n_checked += 1
    assert n_checked >= 30
    # a row outside the table poisons the results instead of rendering garbage
    bad = packed.copy()
    bad["flux_row"][es[0][0]] = 10 ** 6
    with pytest.raises(MemoryError):
        engine.render_patches(gimgs, bad, bound_cap=25.)


def test_allocation_on_bin_lists_conserves_and_follows_the_multinomial(engine, oracle):
    """cg_field_sample_photons: every photon goes to exactly one owner (per image: sum of the
    sampled patches + sky == sum(int(nelec))); first row / column of a box get none; the
    per-item statistics equal the patch sums; with 4e6 photons per pixel the shares equal
    rate / (epsilon + sum of the owners' rates) to binomial accuracy -- the oracle's rates."""
    level = 4.0e6
    field, gimgs, oimgs, osrcs, packed, sw = _setup(engine, oracle, n_src=70, nelec_level=level)
    sw.eigen.project()
    sw.allocate(seed=5)
    boxes = sw.boxes()
    samp = sw.samp.cpu().numpy()
    counts, unit, noise = sw.counts.cpu().numpy(), sw.unit.cpu().numpy(), sw.noise.cpu().numpy()
    S, n_img = boxes.shape
    H, W = gimgs[0].shape
    for b in range(n_img):
        tot = noise[b]
        for s in range(S):
            bx = boxes[s, b]
            h, w = bx["y1"] - bx["y0"], bx["x1"] - bx["x0"]
            z = samp[s * n_img + b, :h * w].reshape(h, w)
            assert np.all(z >= 0) and np.all(z == np.round(z))
            if h and w:
                assert np.all(z[0, :] == 0) and np.all(z[:, 0] == 0)
            assert counts[s * n_img + b] == z.sum(dtype=np.float64)
            tot += z.sum(dtype=np.float64)
        assert tot == level * H * W
    # boxes and unit-patch sums against the oracle, shares on one band
    b = 2
    oi = oimgs[b]
    lam = np.full((H, W), oi.epsilon)
    rates = []
    for s, o in enumerate(osrcs):
        want = _capped_box(oracle, o, oi, 25.)
        bx = boxes[s, b]
        got = (bx["x0"], bx["x1"], bx["y0"], bx["y1"])
        if want[1] <= want[0] or want[3] <= want[2]:
            assert got[1] <= got[0] or got[3] <= got[2]
            rates.append(None)
            continue
        assert got == want
        p, yl, xl = _capped_patch(oracle, o, oi, 25.)
        band_flux = (o.fluxes[b] / oi.calib) * oi.kappa
        assert abs(unit[s * n_img + b] - p.sum() / band_flux) <= 5e-6 * (p.sum() / band_flux)
        rates.append((p, yl, xl))
        lam[yl[0] + 1:yl[1], xl[0] + 1:xl[1]] += p[1:, 1:]
    worst = 0.
    for s, r in enumerate(rates):
        if r is None:
            continue
        p, yl, xl = r
        share = p / lam[yl[0]:yl[1], xl[0]:xl[1]]
        share[0, :] = 0
        share[:, 0] = 0
        z = samp[s * n_img + b, :p.size].reshape(p.shape)
        sd = np.sqrt(level * share * (1 - share)) + 1.
        worst = max(worst, float((np.abs(z - level * share) / sd).max()))
    assert worst < 6.5
    # seeded: the same call reproduces bit for bit
    sw.allocate(seed=5)
    assert np.array_equal(sw.samp.cpu().numpy(), samp) and np.array_equal(sw.noise.cpu().numpy(), noise)
    assert np.array_equal(sw.unit.cpu().numpy(), unit)


def test_gamma_conditionals_on_device(engine, oracle):
    """cg_resample_fluxes / cg_resample_epsilon: flux ~ Gamma(a0 + counts, 1 / (b0 + psf kappa/calib))
    (sources.py:327-350), epsilon ~ Gamma(a0 + noise, 1 / (b0 + H W)) (models.py:156-160).  The
    standardised draws over all (source, band) are N(0, 1) to sampling accuracy; fluxes that are
    functions of other parameters are left alone."""
    field, gimgs, oimgs, osrcs, packed, sw = _setup(engine, oracle, n_src=150)
    sw.eigen.project()
    sw.allocate(seed=11)
    counts = sw.counts.cpu().numpy().reshape(-1, 5)
    unit = sw.unit.cpu().numpy().reshape(-1, 5)
    before = sw.sources()
    zs = []
    for rep in range(8):
        sw.resample_fluxes(seed=100 + rep)
        after = sw.sources()
        for k, s in enumerate(field["sources"]):
            fixed = bool(s.get("eigen")) or (s["a"] == 0 and bool(s["t"]))
            if fixed:
                assert np.array_equal(after["flux"][k], before["flux"][k])
                continue
            for b, im in enumerate(gimgs):
                a_n = 5. + counts[k, b]
                b_n = .005 + unit[k, b] * im.kappa / im.calib
                zs.append((after["flux"][k, b] * b_n - a_n) / np.sqrt(a_n))
    zs = np.array(zs)
    assert zs.size > 3000 and np.all(np.isfinite(zs))
    assert abs(zs.mean()) < 5 / np.sqrt(zs.size) and abs(zs.std() - 1) < .05
    noise = sw.noise.cpu().numpy()
    ez = []
    for rep in range(40):
        sw.resample_epsilon(seed=rep)
        e = sw.eps.cpu().numpy()
        for b, im in enumerate(gimgs):
            a_n, b_n = 5. + noise[b], .005 + im.shape[0] * im.shape[1]
            ez.append((e[b] * b_n - a_n) / np.sqrt(a_n))
    ez = np.array(ez)
    assert abs(ez.mean()) < 5 / np.sqrt(ez.size) and abs(ez.std() - 1) < .2
    sw.stats()
    assert gimgs[0].epsilon == sw.eps.cpu().numpy()[0]


def test_batched_proposals_for_every_source_match_the_oracle(engine, oracle):
    """cg_make_proposals + cg_score_proposals_multi in both modes + cg_select_proposals against the
    oracle's Source.log_likelihood / calculate_acceptance_logprob / generate_background_patch."""
    field, gimgs, oimgs, osrcs, packed, sw = _setup(engine, oracle, n_src=80, proposals=6)
    S, P, n_img = sw.n_src, sw.P, sw.n_img
    sw.eigen.project()
    sw.allocate(seed=3)
    sw.propose(seed=21)
    states = sw.states()
    cur = sw.sources()
    assert np.array_equal(states[:, 0].tobytes(), cur.tobytes())
    du = (states["u"][:, 1:] - cur["u"][:, None]) / sw.scales[0]
    assert abs(du.std() - 1) < .1 and abs(du.mean()) < .1
    for k, s in enumerate(field["sources"]):
        if s.get("eigen") or (s["a"] == 0 and s["t"]):
            assert np.array_equal(states["flux"][k, 1:], np.repeat(cur["flux"][k][None], P, 0))
        if s["a"] == 1:
            assert np.all((states["theta"][k] >= .01) & (states["theta"][k] <= .99))
    rng = np.random.RandomState(8)
    log_u = np.log(rng.uniform(size=S * P))

    def oracle_src(rec, k):
        o = osrcs[k]
        s = field["sources"][k]
        fixed = bool(s.get("eigen")) or (s["a"] == 0 and bool(s["t"]))
        return oracle.SrcParams(u=rec["u"].copy(), a=o.a, theta=float(rec["theta"]), phi=float(rec["phi"]),
                                sigma=float(rec["sigma"]), rho=float(rec["rho"]),
                                fluxes=o.fluxes if fixed else rec["flux"].copy())

    # ---- conditional on the allocated photons (A13) -----------------------------------------
    sw.score(0, log_u)
    delta = sw.delta[0].cpu().numpy().reshape(S, P)
    acc = sw.accept[0].cpu().numpy().reshape(S, P).astype(bool)
    boxes = sw.boxes()
    samp = sw.samp.cpu().numpy()
    sel = list(range(0, S, 3))
    worst = 0.
    for k in sel:
        lst = []
        for b, oi in enumerate(oimgs):
            bx = boxes[k, b]
            h, w = bx["y1"] - bx["y0"], bx["x1"] - bx["x0"]
            if h <= 0 or w <= 0:
                continue
            z = samp[k * n_img + b, :h * w].reshape(h, w).astype(np.float64)
            lst.append((z, (bx["y0"], bx["y1"]), (bx["x0"], bx["x1"]), oi))
        c = oracle.source_log_likelihood(oracle_src(states[k, 0], k), lst)
        for p in (1, P):
            want = oracle.source_log_likelihood(oracle_src(states[k, p], k), lst) - c
            err = abs(delta[k, p - 1] - want) / max(abs(c), 1.)
            worst = max(worst, err)
            if abs(log_u[k * P + p - 1] - want) > 1e-6 * max(abs(c), 1.):
                assert acc[k, p - 1] == (log_u[k * P + p - 1] < want)
    assert worst < LL_RTOL, worst
    # first accepted proposal becomes the state
    sw.select(0)
    new = sw.sources()
    n_acc = 0
    for k in range(S):
        hit = np.flatnonzero(acc[k])
        want = states[k, hit[0] + 1] if hit.size else states[k, 0]
        assert new[k].tobytes() == want.tobytes()
        n_acc += int(hit.size > 0)
    assert int(sw.n_acc.cpu().numpy()[0]) == n_acc

    # ---- backgrounds + full MH ratio over the cached boxes (A15) ----------------------------------
    offs, idx = sw.h_nbr
    st = sw.h_static
    for k in range(0, S, 7):                      # the neighbour lists are the reference's intersection test
        for b in range(n_img):
            a_ = st[k, b]
            want = [t for t in range(S) if t != k and
                    abs(int(a_["x0"]) - int(st[t, b]["x0"])) * 2 < (a_["x1"] - a_["x0"]) + (st[t, b]["x1"] - st[t, b]["x0"]) and
                    abs(int(a_["y0"]) - int(st[t, b]["y0"])) * 2 < (a_["y1"] - a_["y0"]) + (st[t, b]["y1"] - st[t, b]["y0"])]
            assert list(idx[offs[k * n_img + b]:offs[k * n_img + b + 1]]) == want
    sw.backgrounds()
    sw.propose(seed=22)
    states = sw.states()
    sw.score(1, log_u)
    delta = sw.delta[1].cpu().numpy().reshape(S, P)
    acc = sw.accept[1].cpu().numpy().reshape(S, P).astype(bool)
    bg = sw.bg.cpu().numpy()
    cur_o = [oracle_src(states[k, 0], k) for k in range(S)]
    worst_bg, worst = 0., 0.
    for k in range(0, S, 9):
        bxs, bgs = [], []
        for b, oi in enumerate(oimgs):
            bx = st[k, b]
            xlim, ylim = (int(bx["x0"]), int(bx["x1"])), (int(bx["y0"]), int(bx["y1"]))
            others = [cur_o[t] for t in idx[offs[k * n_img + b]:offs[k * n_img + b + 1]]]
            want = oracle.generate_background_patch(cur_o[k], others, oi, xlim, ylim)
            got = bg[k * n_img + b, :want.size].reshape(want.shape)
            worst_bg = max(worst_bg, hp.rel_err(got, want, FLOOR * want.max()))
            bxs.append((xlim, ylim))
            bgs.append(want)
        for p in (1, P):
            want = oracle.calculate_acceptance_logprob(cur_o[k], oracle_src(states[k, p], k), oimgs, bxs, bgs)
            c = sum(oracle.image_like(cur_o[k], oi, bx[0], bx[1], g) for oi, bx, g in zip(oimgs, bxs, bgs))
            worst = max(worst, abs(delta[k, p - 1] - want) / max(abs(c), 1.))
            if abs(log_u[k * P + p - 1] - want) > 1e-6 * max(abs(c), 1.):
                assert acc[k, p - 1] == (log_u[k * P + p - 1] < want)
    assert worst_bg < RATE_RTOL, worst_bg
    assert worst < LL_RTOL, worst


def test_hmc_of_every_galaxy_in_one_launch(engine, oracle):
    """GibbsSweep.hmc_galaxies (cg_hmc_galaxy_params on the sweep's boxes and allocated photons, live
    cd_at_pixel route): the start energy of every galaxy equals the oracle's conditional likelihood of its
    photons (celeste_galaxy_conditionals.py:41 on the capped footprints) minus the kinetic term; stars and
    rejected galaxies keep their state bit for bit, accepted ones move; seeded -> reproducible."""
    from desi_mcmc_b200.engine import HMC_GALAXY_MASS
    field, gimgs, oimgs, osrcs, packed, sw = _setup(engine, oracle, n_src=40, proposals=0, background_pass=False)
    sw.eigen.project()
    sw.allocate(seed=3)
    before = sw.sources()
    boxes, samp = sw.boxes(), sw.samp.cpu().numpy()
    n_img = len(gimgs)
    engine.launch_count(reset=True)
    p, acc = sw.hmc_galaxies(seed=9, step=2e-9, n_steps=3)
    assert engine.launch_count() == 1
    after = sw.sources()
    gal = np.nonzero(before["type"] == 1)[0]
    st = sw.hmc_stats
    assert gal.size > 20 and p.shape == (gal.size, 11) and np.all(st["status"] == 0) and np.all(np.isfinite(st["h"]))
    # the generator of hmc_galaxies, replayed: kinetic term of the momentum before the refresh
    rng = np.random.RandomState(9)
    ke = np.sum((np.sqrt(HMC_GALAXY_MASS) * rng.randn(gal.size, 11)) ** 2 / HMC_GALAXY_MASS, axis=1)
    for j in range(0, gal.size, 5):
        s = int(gal[j])
        ll = 0.
        for b, oi in enumerate(oimgs):
            bx = boxes[s, b]
            pch, yl, xl = _capped_patch(oracle, osrcs[s], oi, 25.)
            if pch is None:
                continue
            assert (xl[0], xl[1], yl[0], yl[1]) == (bx["x0"], bx["x1"], bx["y0"], bx["y1"])
            z = samp[s * n_img + b, :pch.size].reshape(pch.shape).astype(np.float64)
            ll += np.sum(z * np.log(pch) - pch)
        want = ll - .5 * ke[j]
        assert abs(st["h"][j, 0] - want) <= 1e-9 * max(abs(want), 1.), (s, st["h"][j, 0], want)
    stars = np.nonzero(before["type"] != 1)[0]
    assert before[stars].tobytes() == after[stars].tobytes()
    for j, s in enumerate(gal):
        same = before[s:s + 1].tobytes() == after[s:s + 1].tobytes()
        assert same == (not acc[j])
    assert acc.sum() > 0
    # same seed, same state -> same trajectory
    sw2 = _setup(engine, oracle, n_src=40, proposals=0, background_pass=False)[5]
    sw2.eigen.project()
    sw2.allocate(seed=3)
    p2, acc2 = sw2.hmc_galaxies(seed=9, step=2e-9, n_steps=3)
    assert np.array_equal(p, p2) and np.array_equal(acc, acc2) and sw2.sources().tobytes() == after.tobytes()


def test_full_sweeps_are_reproducible_and_consistent(engine, oracle):
    """Three whole sweeps with no host hop in between: seeded -> bitwise reproducible; the
    log-likelihood a sweep reports is the field likelihood of the state it ends in; photons are
    conserved in every sweep; the guard band re-scores near-ties in float64."""
    out = []
    for rep in range(2):
        field, gimgs, oimgs, osrcs, packed, sw = _setup(engine, oracle, n_src=60, proposals=8, guard=.5)
        for _ in range(3):
            sw.sweep()
        st = sw.stats()
        out.append((sw.sources().tobytes(), st["loglik"].copy(), st["epsilon"].copy()))
        tot = sw.noise.cpu().numpy() + sw.counts.cpu().numpy().reshape(-1, 5).sum(0)
        nel = np.array([np.floor(g.nelec.cpu().numpy().astype(np.float64)).sum() for g in gimgs])
        assert np.array_equal(tot, nel)
        assert st["accepted"].sum() > 0 and st["guard"].sum() > 0
        ll = engine.field_loglik(gimgs, sw.sources(), bound_cap=25., max_patch_px=sw.max_patch_px).cpu().numpy()
        assert np.array_equal(ll, st["loglik"])
    assert out[0][0] == out[1][0] and np.array_equal(out[0][1], out[1][1]) and np.array_equal(out[0][2], out[1][2])


def test_geweke_joint_distribution_of_allocation_and_flux_update(engine, oracle):
    """Geweke's joint-distribution test (the method of experiments/rjump/geweke_test.py:94-121) for the
    device transition "allocate photons, then redraw the fluxes": the successive-conditional simulator
    (image | flux from the model on the host, flux' | image from the DEVICE kernels) must leave the
    prior Gamma(a0, 1/b0) invariant.  First and second moments of the flux in the band that has data,
    with batch-means standard errors; the bands without data must be exact prior draws."""
    import torch
    from desi_mcmc_b200 import _ffi
    a0, b0 = 5., .005
    size = 25
    field = synth.make_stamp(1001, size=size, n_star=1, n_gal=0)
    meta = dict(field["images"][2])
    meta["calib"] = 40.                                    # weak data: ~100 photons at the prior mean
    img = engine.Image(nelec=np.zeros((size, size)), **synth.image_kwargs(meta))
    star = field["sources"][0]
    rec = engine.pack_sources([engine.SrcParams(u=star["u"], a=0, fluxes=np.full(5, a0 / b0))])
    sw = engine.GibbsSweep([img], rec, bound_cap=25., proposals=0, background_pass=False)
    unit, boxes = engine.render_patches([img], rec, default_flags=_ffi.CG_ITEM_FLUX_UNIT, bound_cap=25.)
    bx = boxes[0]
    h, w = int(bx["y1"] - bx["y0"]), int(bx["x1"] - bx["x0"])
    psi = unit[0, :h * w].cpu().numpy().astype(np.float64).reshape(h, w)
    c = img.kappa / img.calib
    rng = np.random.RandomState(2026)
    n_iter, burn = 2400, 200
    f = rng.gamma(a0, 1. / b0)
    trace = np.zeros((n_iter, 5))
    cur = rec.copy()
    for it in range(n_iter):
        lam = np.full((size, size), img.epsilon)
        lam[bx["y0"]:bx["y1"], bx["x0"]:bx["x1"]] += f * c * psi
        img.set_nelec(rng.poisson(lam).astype(np.float32))
        cur["flux"][0, :] = f
        sw.d_cur.copy_(torch.from_numpy(cur.view(np.uint8).reshape(-1)))
        sw.allocate(seed=10 * it + 1)
        sw.resample_fluxes(seed=10 * it + 2)
        trace[it] = sw.sources()["flux"][0]
        f = trace[it, 2]
    tr = trace[burn:]
    mean, second = a0 / b0, a0 * (a0 + 1) / b0 ** 2

    def batch_se(x, nb=20):
        m = x[:len(x) // nb * nb].reshape(nb, -1).mean(axis=1)
        return m.std(ddof=1) / np.sqrt(nb)
    z1 = (tr[:, 2].mean() - mean) / batch_se(tr[:, 2])
    z2 = ((tr[:, 2] ** 2).mean() - second) / batch_se(tr[:, 2] ** 2)
    print("Geweke z-scores (band with data): mean %.2f, second moment %.2f; lag-1 autocorrelation %.2f" %
          (z1, z2, np.corrcoef(tr[:-1, 2], tr[1:, 2])[0, 1]))
    assert abs(z1) < 4. and abs(z2) < 4.
    for b in (0, 1, 3, 4):                                 # no image in these bands: independent prior draws
        x = tr[:, b]
        assert abs(x.mean() - mean) < 4.5 * np.sqrt(a0) / b0 / np.sqrt(len(x))
        assert abs(x.var() / (a0 / b0 ** 2) - 1) < .15
    # the chain actually depends on the data: photons were allocated to the source
    assert sw.counts.cpu().numpy().sum() > 0
