"""BASELINE.json's full-size configurations, checked through size-independent
properties plus oracle spot checks the CPU finishes in seconds (the oracle
cannot render a 2048x1489x5 field with 500 sources in test time)."""
import numpy as np
import pytest
import torch

from desi_mcmc_b200 import synth
from tests import helpers as hp

pytestmark = pytest.mark.gpu

H, W, N_SRC = 1489, 2048, 500


@pytest.fixture(scope="module")
def engine():
    from desi_mcmc_b200 import engine as eng
    eng.init(0)
    return eng


@pytest.fixture(scope="module")
def oracle():
    from oracle import ref_cpu
    return ref_cpu


def test_c2_full_field_properties(engine, oracle):
    """Config 2: 2048x1489 x 5 bands, 500 sources, footprints capped at 101x101."""
    from desi_mcmc_b200 import _ffi
    dev = engine.device()
    fld = synth.make_field(1002, H, W, N_SRC, .6)
    imgs = [engine.Image(nelec=torch.zeros((H, W), dtype=torch.float32, device=dev), **synth.image_kwargs(m))
            for m in fld["images"]]
    srcs = engine.pack_sources(hp.gpu_sources(engine, fld))
    cap, mpp = 50.0, 103 * 103
    _, models = engine.field_loglik(imgs, srcs, bound_cap=cap, max_patch_px=mpp, return_model=True)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1002)
    for im, lam in zip(imgs, models):
        im.set_nelec(torch.poisson(lam, generator=gen))
    ll, models = engine.field_loglik(imgs, srcs, bound_cap=cap, max_patch_px=mpp, return_model=True)
    ll = ll.cpu().numpy()

    # (a) checksum of checksums: the fused kernel's sum equals an independent float64
    #     Poisson reduction (cg_poisson_loglike) of ITS OWN model image
    for i, im in enumerate(imgs):
        out = torch.zeros(1, dtype=torch.float64, device=dev)
        x64, m64 = im.nelec.double(), models[i].double()      # keep the temporaries alive across the call
        _ffi.check(engine.lib().cg_poisson_loglike(x64.data_ptr(), m64.data_ptr(), None, H * W,
                                                   out.data_ptr(), engine.stream_ptr()))
        assert abs(out.item() - ll[i]) <= 1e-9 * abs(ll[i])

    # (b) linearity: the model image is eps + the sum of the separately rendered
    #     patches (render_kernel, a different kernel and tiling)
    sub = np.arange(0, N_SRC, 7)
    patches, boxes = engine.render_patches(imgs, srcs[sub], bound_cap=cap, max_patch_px=mpp)
    _, m_sub = engine.field_loglik(imgs, srcs[sub], bound_cap=cap, max_patch_px=mpp, return_model=True)
    patches = patches.cpu().numpy()
    for i, im in enumerate(imgs):
        lam = np.full((H, W), np.float32(im.epsilon), dtype=np.float32)
        for k in range(len(sub)):
            b = boxes[k * 5 + i]
            h, w = int(b["y1"] - b["y0"]), int(b["x1"] - b["x0"])
            assert h <= 101 and w <= 101
            if h > 0 and w > 0:
                lam[b["y0"]:b["y1"], b["x0"]:b["x1"]] += patches[k * 5 + i, :h * w].reshape(h, w)
        assert hp.rel_err(m_sub[i].cpu().numpy(), lam, 1.0) < 1e-6

    # (c) permutation of the source list changes only the fp32 summation order
    perm = np.random.RandomState(0).permutation(N_SRC)
    ll_p = engine.field_loglik(imgs, srcs[perm], bound_cap=cap, max_patch_px=mpp).cpu().numpy()
    assert np.all(np.abs(ll_p - ll) <= 1e-9 * np.abs(ll))
    # (d) and the same list twice is bitwise reproducible
    ll_1 = engine.field_loglik(imgs, srcs, bound_cap=cap, max_patch_px=mpp).cpu().numpy()
    ll_2 = engine.field_loglik(imgs, srcs, bound_cap=cap, max_patch_px=mpp).cpu().numpy()
    assert np.array_equal(ll_1, ll_2)
    # (without a model output, sky-only bins take the one-log-per-bin shortcut: same sum to rounding)
    assert np.all(np.abs(ll_1 - ll) <= 1e-10 * np.abs(ll))

    # (e) oracle spot check: a 160x160 crop and every source whose footprint reaches it
    osrcs = hp.oracle_sources(oracle, fld)
    x0, y0, n = 700, 500, 160
    i = 2
    m = fld["images"][i]
    oimg = oracle.SyntheticImage(nelec=np.zeros((H, W)), **synth.image_kwargs(m))
    want = np.full((n, n), oimg.epsilon)
    hit = 0
    for s_, os_ in enumerate(osrcs):
        px, py = oimg.equa2pixel(os_.u)
        if px < x0 - 52 or px > x0 + n + 52 or py < y0 - 52 or py > y0 + n + 52:
            continue
        # footprint of the capped render (prepare_kernel: bound = min(bound, cap)), oracle patch on it
        pb = engine.render_patches([imgs[i]], srcs[s_:s_ + 1], bound_cap=cap, max_patch_px=mpp)[1][0]
        if pb["x1"] <= pb["x0"]:
            continue
        f, yl, xl = oracle.gen_src_image_with_fluxes(os_, oimg, xlim=(int(pb["x0"]), int(pb["x1"])),
                                                     ylim=(int(pb["y0"]), int(pb["y1"])))
        full = np.zeros((H, W))
        full[yl[0]:yl[1], xl[0]:xl[1]] = f
        want += full[y0:y0 + n, x0:x0 + n]
        hit += 1
    assert hit >= 3
    got = models[i].cpu().numpy()[y0:y0 + n, x0:x0 + n]
    assert hp.rel_err(got, want, 1.0) < 1e-5


def test_c3_full_batch_properties(engine, oracle):
    """Config 3: 4096 proposals of one galaxy on a 5-band 64x64 stamp (K = 42)."""
    P, size = 4096, 64
    field = synth.make_stamp(1003, size=size, n_star=1, n_gal=1)
    nelec = hp.simulate_nelec(oracle, field, 1004)
    gimgs = hp.gpu_images(engine, field, nelec)
    oimgs = hp.oracle_images(oracle, field, nelec)
    cur = field["sources"][1]
    props, log_u = synth.make_proposals(1005, cur, P)
    boxes = [(0, size, 0, size)] * 5
    bgs = [np.full((size, size), im.epsilon) for im in oimgs]
    packed = engine.pack_sources(hp.gpu_sources(engine, [cur] + props))
    sc = engine.ProposalScorer(gimgs, boxes, [b.astype(np.float32) for b in bgs], max_proposals=P, guard=1e-2)
    delta, accept, n_guard = sc.score(packed, log_u)
    assert delta.shape == (P,) and np.all(np.isfinite(delta))
    assert np.array_equal(accept.astype(bool), log_u < delta)            # the accept test (:258) on device
    # the current state proposed against itself scores exactly zero
    d0, a0, _ = sc.score(np.concatenate([packed[:1], packed[:1]]), np.array([-1e-9]))
    assert d0[0] == 0.0 and bool(a0[0])
    # batch independence: a proposal scores the same alone, in a small batch, or among 4096
    pick = np.array([0, 1, 17, 1000, 4095])
    d_small, _, _ = sc.score(np.concatenate([packed[:1], packed[1 + pick]]), log_u[pick])
    assert np.array_equal(d_small, delta[pick])
    # oracle: decisions and deltas of a 48-proposal subsample
    sub = np.random.RandomState(1).choice(P, 48, replace=False)
    ocur = hp.oracle_sources(oracle, [cur])[0]
    oprops = hp.oracle_sources(oracle, [props[k] for k in sub])
    ob = [((0, size), (0, size))] * 5
    want = np.array([oracle.calculate_acceptance_logprob(ocur, p, oimgs, ob, bgs) for p in oprops])
    assert np.array_equal(accept[sub].astype(bool), log_u[sub] < want)
    assert np.max(np.abs(delta[sub] - want)) < 5e-3


def test_c4_crowded_field_full_size(engine, oracle):
    """Config 4 at BASELINE's size: 5000 sources (10 % black-body stars, 5 % eigen-spectrum point
    sources, galaxies) with footprints forced to ~50 x 50 on a 2048 x 1489 x 5 field, one full Gibbs
    sweep with 16 proposals per source on the device.  Whole field: every photon of every band is
    conserved by the allocation, the sweep is finite and its reported log-likelihood is the field
    likelihood of the state it ends in.  A 200-source subsample against the oracle: footprints bit
    for bit, rendered rates <= 1e-5 (r band), conditional proposal scores <= 1e-6 of the source's
    log-likelihood with identical accept / reject; 12 sources x 2 proposals of the full MH ratio
    (backgrounds from ~16 active neighbours each) likewise."""
    import os
    import torch
    from desi_mcmc_b200 import _ffi
    from tests.test_gpu_crowded import _capped_box, _capped_patch, _oracle_fluxes
    oracle.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    H, W, S, P = 1489, 2048, 5000, 16
    field = synth.make_crowded_field(1004, H=H, W=W, n_src=S)
    B, lam0 = synth.make_eigen_basis(1004)
    imgs, packed, eig, sw = synth.instantiate_crowded(engine, field, B, lam0, simulate_seed=1004,
                                                      proposals=P, seed=1004)
    n_star = sum(1 for s in field["sources"] if s["a"] == 0 and s["t"])
    assert 400 < n_star < 600 and 200 < field["n_eigen"] < 300
    n_img = 5
    nelec_tot = np.array([float(torch.floor(im.nelec.double()).sum().item()) for im in imgs])

    # ---- the sweep, phase by phase, so the intermediate state can be checked ---------------
    eig.project()
    sw.allocate(seed=1)
    counts = sw.counts.cpu().numpy().reshape(S, n_img)
    assert np.array_equal(sw.noise.cpu().numpy() + counts.sum(0), nelec_tot)          # photon conservation
    boxes = sw.boxes()
    sw.resample_epsilon(seed=2)
    sw.resample_fluxes(seed=3)
    sw.propose(seed=4)
    states = sw.states()
    rng = np.random.RandomState(1004)
    log_u = np.log(rng.uniform(size=S * P))
    sw.score(0, log_u)
    delta = sw.delta[0].cpu().numpy().reshape(S, P)
    acc = sw.accept[0].cpu().numpy().reshape(S, P).astype(bool)
    assert np.all(np.isfinite(delta))
    assert np.array_equal(acc, log_u.reshape(S, P) < delta) or int(sw.n_guard.cpu().numpy()[0]) > 0

    metas = field["images"]
    oimgs = [oracle.SyntheticImage(nelec=im.nelec.cpu().numpy().astype(np.float64), **synth.image_kwargs(m))
             for im, m in zip(imgs, metas)]

    def oracle_src(rec, k):
        s = field["sources"][k]
        fixed = bool(s.get("eigen")) or (s["a"] == 0 and bool(s["t"]))
        f = _oracle_fluxes(oracle, s, metas, B, lam0) if fixed else rec["flux"].copy()
        return oracle.SrcParams(u=rec["u"].copy(), a=int(rec["type"]), theta=float(rec["theta"]),
                                phi=float(rec["phi"]), sigma=float(rec["sigma"]), rho=float(rec["rho"]), fluxes=f)

    sub = np.random.RandomState(7).choice(S, 200, replace=False)
    rate_items = engine.make_items(sub, [2] * sub.size)
    patches, pboxes = engine.render_patches(imgs, states[:, 0].copy(), items=rate_items, bound_cap=25.,
                                            max_patch_px=sw.max_patch_px)
    samp_rows = torch.from_numpy(np.concatenate([k * n_img + np.arange(n_img) for k in sub])).to(engine.device())
    samp = sw.samp[samp_rows].cpu().numpy().reshape(sub.size, n_img, -1)
    worst_rate, worst_ll, n_dec = 0., 0., 0
    for j, k in enumerate(sub):
        o = oracle_src(states[k, 0], k)
        for b in range(n_img):                                                          # footprints, every band
            bx = boxes[k, b]
            assert (bx["x0"], bx["x1"], bx["y0"], bx["y1"]) == _capped_box(oracle, o, oimgs[b], 25.)
        want, yl, xl = _capped_patch(oracle, o, oimgs[2], 25.)                          # rates, r band
        got = patches[j, :want.size].cpu().numpy().reshape(want.shape)
        worst_rate = max(worst_rate, hp.rel_err(got, want, 1e-6 * want.max()))
        lst = []
        for b in range(n_img):
            bx = boxes[k, b]
            h, w = bx["y1"] - bx["y0"], bx["x1"] - bx["x0"]
            lst.append((samp[j, b, :h * w].reshape(h, w).astype(np.float64), (bx["y0"], bx["y1"]),
                        (bx["x0"], bx["x1"]), oimgs[b]))
        c = oracle.source_log_likelihood(o, lst)
        p = 1 + (j % P)
        d_want = oracle.source_log_likelihood(oracle_src(states[k, p], k), lst) - c
        worst_ll = max(worst_ll, abs(delta[k, p - 1] - d_want) / max(abs(c), 1.))
        if abs(log_u[k * P + p - 1] - d_want) > 1e-6 * max(abs(c), 1.):
            assert acc[k, p - 1] == (log_u[k * P + p - 1] < d_want)
            n_dec += 1
    assert worst_rate < 1e-5, worst_rate
    assert worst_ll < 1e-6, worst_ll
    assert n_dec >= 190

    sw.select(0)
    sw.backgrounds()
    sw.propose(seed=5)
    states = sw.states()
    sw.score(1, log_u)
    delta = sw.delta[1].cpu().numpy().reshape(S, P)
    acc = sw.accept[1].cpu().numpy().reshape(S, P).astype(bool)
    offs, idx = sw.h_nbr
    st = sw.h_static
    worst_ll = 0.
    for k in sub[:12]:
        me = oracle_src(states[k, 0], k)
        bxs, bgs = [], []
        for b, oi in enumerate(oimgs):
            bx = st[k, b]
            xlim, ylim = (int(bx["x0"]), int(bx["x1"])), (int(bx["y0"]), int(bx["y1"]))
            others = [oracle_src(states[t, 0], t) for t in idx[offs[k * n_img + b]:offs[k * n_img + b + 1]]]
            bgs.append(oracle.generate_background_patch(me, others, oi, xlim, ylim))
            bxs.append((xlim, ylim))
        c = sum(oracle.image_like(me, oi, bx[0], bx[1], g) for oi, bx, g in zip(oimgs, bxs, bgs))
        for p in (1, P):
            want = oracle.calculate_acceptance_logprob(me, oracle_src(states[k, p], k), oimgs, bxs, bgs)
            worst_ll = max(worst_ll, abs(delta[k, p - 1] - want) / max(abs(c), 1.))
            if abs(log_u[k * P + p - 1] - want) > 1e-6 * max(abs(c), 1.):
                assert acc[k, p - 1] == (log_u[k * P + p - 1] < want)
    assert worst_ll < 1e-6, worst_ll
    sw.select(1)
    sw.loglik()
    res = sw.stats()
    assert np.all(np.isfinite(res["loglik"])) and res["accepted"].min() > 0
    ll = engine.field_loglik(imgs, sw.sources(), bound_cap=25., max_patch_px=sw.max_patch_px).cpu().numpy()
    assert np.array_equal(ll, res["loglik"])
    assert sw.workspace_bytes < 8e9          # ~3.4 GB of workspace + patches for the whole field
