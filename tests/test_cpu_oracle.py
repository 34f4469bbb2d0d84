"""CPU suite, part 1: the oracle against the reference's own known-answer
tests and against the compiled reference (oracle/_ref) when it is present."""
import os

import numpy as np
import pytest
from scipy.special import logsumexp
from scipy.stats import multivariate_normal

from oracle import build_ref, ref_cpu as ref
from tests.test_gpu_parity import kat_inputs

HAVE_REF = build_ref.built()


def test_kat_gmm_like_vs_scipy():
    """CelestePy/test/test_gmm.py:63-105: gmm_like_2d == sum_k scipy mvn."""
    X, ws, means, covs = kat_inputs()
    got = ref.gmm_like_2d(X, ws, means, covs)
    ll = np.zeros((X.shape[0], len(ws)))
    for k in range(len(ws)):
        ll[:, k] = multivariate_normal(mean=means[k], cov=covs[k]).logpdf(X) + np.log(ws[k])
    want = np.exp(logsumexp(ll, axis=1))
    assert np.allclose(got, want)
    # the live evaluator (util/dists/mog.py:5-21) is the same function in log space
    icovs = np.array([np.linalg.inv(c) for c in covs])
    dets = np.array([np.linalg.det(c) for c in covs])
    assert np.allclose(np.exp(ref.mog_loglike(X, means, icovs, dets, ws)), want)


def test_kat_mvn_logpdf_vs_scipy():
    """CelestePy/test/test_gmm.py:4-60, atol 1e-6 as there."""
    X, ws, means, covs = kat_inputs(K=3, n=100)
    ci = np.linalg.inv(covs[2])
    _, logdet = np.linalg.slogdet(covs[2])
    got = ref.multivariate_normal_2d_covinv_logdet_logpdf(np.zeros(X.shape[0]), X, means[2], ci, logdet)
    want = multivariate_normal(mean=means[2], cov=covs[2]).logpdf(X)
    assert np.allclose(got, want, atol=1e-6)


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built")
def test_restatement_vs_compiled_reference():
    gm, cf = build_ref.load()
    X, ws, means, covs = kat_inputs()
    got = np.zeros(X.shape[0])
    gm.gmm_like_2d(got, X, ws, means, covs)
    assert np.allclose(got, ref.gmm_like_2d(X, ws, means, covs), rtol=1e-12, atol=0)
    rng = np.random.RandomState(3)
    W = rng.normal(size=(2, 2))
    W = W.dot(W.T)
    v = rng.uniform(0, 50, size=2)
    iw = rng.dirichlet(np.ones(3))
    im = rng.normal(size=(3, 2))
    ic = np.array([np.eye(2) * s for s in (2., 7., 30.)])
    a = cf.gen_galaxy_prof_psf_mixture_params(W, v, iw, im, ic, ref.exp_amp, ref.exp_var)
    b = ref.gen_galaxy_prof_psf_mixture_params(W, v, iw, im, ic, ref.exp_amp, ref.exp_var)
    for x, y in zip(a, b):
        assert np.array_equal(np.asarray(x), y)
    th = np.array([.3, .7])
    a = cf.gen_galaxy_psf_mixture_params(th, W, v, iw, im, ic, ref.exp_amp, ref.exp_var, ref.dev_amp, ref.dev_var)
    b = ref.gen_galaxy_psf_mixture_params(th, W, v, iw, im, ic, ref.exp_amp, ref.exp_var, ref.dev_amp, ref.dev_var)
    for x, y in zip(a, b):
        assert np.array_equal(np.asarray(x), y)


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built")
def test_randomkit_matches_numpy_legacy_stream():
    """'identical random draws' = NumPy's legacy RandomState stream; the
    reference's vendored randomkit (deps/randomkit) produces the same one.
    First binomial draw is the reference's regression case
    (CelestePy/test/test_celeste_sample_sources.py:4-8)."""
    import ctypes as C
    lib = C.CDLL(os.path.join(build_ref.OUT, "librandomkit.so"))
    state = C.create_string_buffer(8192)
    lib.rk_seed.argtypes = [C.c_ulong, C.c_void_p]
    lib.rk_double.restype = C.c_double
    lib.rk_double.argtypes = [C.c_void_p]
    lib.rk_binomial.restype = C.c_long
    lib.rk_binomial.argtypes = [C.c_void_p, C.c_long, C.c_double]
    lib.rk_seed(89, state)
    rs = np.random.RandomState(89)
    assert [lib.rk_double(state) for _ in range(5)] == list(rs.random_sample(5))
    lib.rk_seed(0, state)
    rs = np.random.RandomState(0)
    args = [(914, 6.29379e-16), (1000, .3), (1000, .31), (1000, .29), (1000, .305)]
    assert [lib.rk_binomial(state, n, p) for n, p in args] == [rs.binomial(n, p) for n, p in args]
    assert rs is not None


def test_two_galaxy_routes_agree():
    """Cython route (celeste_galaxy_conditionals.py:134-182) and live MoG route
    (:185-214) are the same mathematics when the CD matrices coincide."""
    from desi_mcmc_b200 import synth
    from tests import helpers as hp
    field = synth.make_stamp(11, size=31)
    imgs = hp.oracle_images(ref, field, [np.zeros((31, 31))] * 5)
    g = hp.oracle_sources(ref, field)[1]
    for img in imgs:
        px, py = img.equa2pixel(g.u)
        cd = img.cd_at_pixel(px, py)
        R = ref.gen_galaxy_transformation(g.sigma, g.rho, g.phi, cd)
        fe, yl, xl = ref.gen_galaxy_prof_psf_image('exp', R, g.u, img, xlim=(0, 31), ylim=(0, 31))
        fd, _, _ = ref.gen_galaxy_prof_psf_image('dev', R, g.u, img, xlim=(0, 31), ylim=(0, 31))
        live, _, _ = ref.gen_galaxy_psf_image(g.shape, g.u, img, xlim=(0, 31), ylim=(0, 31))
        assert np.allclose(g.theta * fe + (1 - g.theta) * fd, live, rtol=1e-10, atol=0)


def test_wcs_round_trip():
    from desi_mcmc_b200 import synth
    from tests import helpers as hp
    field = synth.make_stamp(5, size=51)
    for img in hp.oracle_images(ref, field, [np.zeros((51, 51))] * 5):
        p = np.array([12.25, 40.5])
        assert np.allclose(img.equa2pixel(img.pixel2equa(p)), p, atol=1e-8)


def test_filter_table_matches_reference_file():
    from desi_mcmc_b200 import _ffi
    a = ref.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
    v_npz = ref.photons_per_joule(8395., 'r')
    assert abs(v_npz - 1.666005755452341e+17) < 1e-9 * v_npz       # SURVEY.md 8c restated value
    path = "/root/reference/CelestePy/filter_curves"
    if os.path.exists(path):
        b = ref.load_filter_curves(path)
        for band in ref.BANDS:
            for x, y in zip(a, b):
                assert np.array_equal(x[band], y[band])


GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")


def test_oracle_vs_golden_vectors():
    """Vectors produced by the compiled reference itself (tests/golden/make_golden.py)."""
    g = np.load(GOLDEN)
    X, ws, means, covs = kat_inputs()
    got = ref.gmm_like_2d(X, ws, means, covs)
    assert np.allclose(got[g["kat_idx"]], g["kat_probs"], rtol=1e-12, atol=0)
    X3, ws3, means3, covs3 = kat_inputs(K=3, n=100)
    ci = np.linalg.inv(covs3[2])
    _, logdet = np.linalg.slogdet(covs3[2])
    lls = ref.multivariate_normal_2d_covinv_logdet_logpdf(np.zeros(X3.shape[0]), X3, means3[2], ci, logdet)
    assert np.allclose(lls[::5], g["mvn_lls"], rtol=1e-12, atol=1e-12)
    pw, pm, pc = ref.gen_galaxy_prof_psf_mixture_params(g["mp_W"], g["mp_v"], g["mp_iw"], g["mp_im"],
                                                        g["mp_ic"], ref.dev_amp, ref.dev_var)
    assert np.array_equal(pw, g["prof_w"]) and np.array_equal(pm, g["prof_m"]) and np.array_equal(pc, g["prof_c"])
    jw, jm, jc = ref.gen_galaxy_psf_mixture_params(g["mp_th"], g["mp_W"], g["mp_v"], g["mp_iw"], g["mp_im"],
                                                   g["mp_ic"], ref.exp_amp, ref.exp_var, ref.dev_amp, ref.dev_var)
    assert np.array_equal(jw, g["joint_w"]) and np.array_equal(jm, g["joint_m"]) and np.array_equal(jc, g["joint_c"])
    assert np.array_equal(np.random.RandomState(89).random_sample(16), g["rk_doubles"])
    rs = np.random.RandomState(0)
    want = [rs.binomial(914, 6.29379e-16)] + [rs.binomial(1000, .3) for _ in range(7)]
    assert list(g["rk_binom"]) == want and g["rk_binom"][0] == 0


def test_oracle_sample_source_counts_conserves_photons():
    """celeste_sample_sources.pyx:61-156 restated over the legacy RandomState
    stream: every photon lands in one source patch or the sky sum, pixels on a
    patch's first row/column are never offered to it (:53-54)."""
    rng = np.random.RandomState(3)
    nelec = rng.poisson(200., size=(12, 15)).astype(np.float64)
    boxes = np.array([[1, 9, 2, 10], [5, 14, 4, 12]])
    src = [ref.NativePatch(rng.uniform(1., 50., size=(8, 8)), (2, 10), (1, 9)),
           ref.NativePatch(rng.uniform(1., 50., size=(8, 9)), (4, 12), (5, 14))]
    samp = [ref.NativePatch(np.zeros((8, 8)), (2, 10), (1, 9)), ref.NativePatch(np.zeros((8, 9)), (4, 12), (5, 14))]
    samp, noise = ref.sample_source_counts(src, samp, nelec, boxes, 30., np.random.RandomState(0))
    assert noise + samp[0].data.sum() + samp[1].data.sum() == nelec.sum()
    assert np.all(samp[0].data[0] == 0) and np.all(samp[0].data[:, 0] == 0) and samp[0].data.sum() > 0
    again = [ref.NativePatch(np.zeros((8, 8)), (2, 10), (1, 9)), ref.NativePatch(np.zeros((8, 9)), (4, 12), (5, 14))]
    again, noise2 = ref.sample_source_counts(src, again, nelec, boxes, 30., np.random.RandomState(0))
    assert noise2 == noise and np.array_equal(again[1].data, samp[1].data)
    # test_celeste_sample_sources.py:4-8
    assert ref.sample_multinomial(914, np.array([6.29379e-16, 1.]), np.random.RandomState(0))[0] == 0


def test_oracle_prior_restatement_vs_scipy():
    """util/dists/mog.py:5-21 for D = 5 against scipy's multivariate normal (the
    reference's -log(2 pi) does not scale with D: a constant offset of
    (D/2 - 1) log(2 pi)), and the sufficient-statistics layout."""
    from scipy.stats import multivariate_normal
    from scipy.special import logsumexp
    from oracle import ref_cpu as ref
    rng = np.random.RandomState(3)
    K, D = 6, 5
    means = rng.normal(size=(K, D))
    A = rng.normal(size=(K, D, D))
    covs = np.einsum('kij,klj->kil', A, A) + .5 * np.eye(D)
    pis = rng.dirichlet(np.ones(K))
    x = rng.normal(size=(50, D))
    icovs, dets = ref.mog_params(covs)
    got = ref.mog_loglike(x, means, icovs, dets, pis)
    want = logsumexp(np.stack([np.log(pis[k]) + multivariate_normal(means[k], covs[k]).logpdf(x)
                               for k in range(K)], 1), axis=1)
    assert np.allclose(got - want, (D / 2. - 1.) * np.log(2 * np.pi), atol=1e-10)
    r = ref.mog_responsibilities(x, means, covs, pis)
    st = ref.prior_sufficient_stats(r, x)
    assert st.shape == (K * (1 + D + 15),) and abs(st[:K].sum() - 50.) < 1e-10
    k = 2
    assert np.allclose(st[K + k * D:K + (k + 1) * D], (r[:, k:k + 1] * x).sum(0))
    outer = np.einsum('n,ni,nj->ij', r[:, k], x, x)
    assert np.allclose(st[K + K * D + k * 15:K + K * D + (k + 1) * 15], outer[np.triu_indices(D)])
    lf = rng.normal(2., .5, size=(4, 5))
    c = np.dot(lf, ref.FLUX_COLOR_A)
    assert np.allclose(c[:, 4], lf[:, 2]) and np.allclose(c[:, 0], lf[:, 0] - lf[:, 2])
