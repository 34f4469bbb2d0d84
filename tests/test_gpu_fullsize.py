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
