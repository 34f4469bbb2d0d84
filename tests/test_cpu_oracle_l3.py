"""CPU suite, part 3: the L3 oracle pinned on outputs of the REFERENCE'S OWN PYTHON.

tests/golden/reference_l3.npz was produced by running /root/reference/CelestePy/*.py and
experiments/redshift/redshift_utils.py in place (oracle/ref_import.py: syntax-only
Python-2 -> 3 rewrite of the text read, stand-ins for the absent third-party imports;
generator: tests/golden/make_golden_l3.py).  Here:
  * oracle/ref_cpu.py must reproduce every vector (it does so bit for bit; the bar is 1e-13);
  * the product's host-side samplers (compat/slicesample.py, compat/hmc.py) must reproduce the
    reference samplers' seeded traces exactly;
  * the A1 C-ABI entry points must reproduce the compiled reference's tables bit for bit;
  * where /root/reference is mounted, the vectors are regenerated live and compared, so the
    committed file cannot drift from what the reference computes.
"""
import os
import warnings

import numpy as np
import pytest

from oracle import build_ref, ref_cpu as ref, ref_import
from tests.golden import l3_cases as L

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "reference_l3.npz"))
TRACE_KEYS = [k for k in GOLD.files if k.startswith("trace_")]
A17_KEYS = [k for k in GOLD.files if k.startswith("a17_")]
L3_KEYS = [k for k in GOLD.files if k not in TRACE_KEYS and k not in A17_KEYS]


def _worst(got, keys):
    worst, where = 0., None
    for k in keys:
        a, b = GOLD[k], got[k]
        assert a.shape == b.shape, "%s: shape %s vs %s" % (k, a.shape, b.shape)
        if a.size == 0:
            continue
        d = float(np.max(np.abs(a - b) / np.maximum(np.abs(a), 1e-300)))
        if d > worst:
            worst, where = d, k
    return worst, where


def test_oracle_restatement_equals_reference_python():
    """A3-A10, A12-A16: mixtures, bounding radii, WCS, star / galaxy patches and boxes (incl.
    the truncation vs floor/ceil footprints and image edges), flux scaling, model image,
    Poisson log-likelihoods (plain, masked), conditional / isolated source likelihoods,
    background patches, MH acceptance log-probabilities, Planck fluxes."""
    assert len(L3_KEYS) >= 120
    got = L.run(L.oracle_backend())
    assert set(L3_KEYS) <= set(got)
    worst, where = _worst(got, L3_KEYS)
    assert worst <= 1e-13, (where, worst)


def test_oracle_eigen_flux_equals_reference_python():
    """A17: project_to_bands / softmax (redshift_utils.py:304-323, :394-397)."""
    ref.set_filter_tables_npz(os.path.join(os.path.dirname(HERE), "desi_mcmc_b200", "data", "filter_curves.npz"))
    got = L.run_a17(ref.project_to_bands, ref.softmax)
    worst, where = _worst(got, A17_KEYS)
    assert worst <= 1e-13, (where, worst)


def test_product_host_samplers_follow_reference_traces():
    """N2 host control flow: compat.slicesample / compat.hmc draw from numpy.random in the
    reference's order -- 25 slice updates x 3 configurations and 40 HMC iterations, seeded,
    equal to the traces of util/infer/slicesample.py and util/infer/hmc.py bit for bit."""
    from desi_mcmc_b200.compat.hmc import hmc
    from desi_mcmc_b200.compat.slicesample import slicesample
    got = L.sampler_traces(slicesample, hmc)
    assert len(TRACE_KEYS) == 4
    for k in TRACE_KEYS:
        assert np.array_equal(GOLD[k], got[k]), k


def test_a1_call_surface_bit_exact():
    """compat.celeste_fast.gen_galaxy_prof_psf_mixture_params / gen_galaxy_psf_mixture_params (the
    C-ABI host entry points) against the compiled reference's outputs committed in
    tests/golden/reference_vectors.npz (celeste_fast.pyx:100-140, :29-94): index / ordering work,
    so the bar is bit-exact."""
    from desi_mcmc_b200.compat import celeste_fast
    g = np.load(os.path.join(HERE, "golden", "reference_vectors.npz"))
    w, m, c = celeste_fast.gen_galaxy_prof_psf_mixture_params(
        g["mp_W"], g["mp_v"], g["mp_iw"], g["mp_im"], g["mp_ic"], ref.dev_amp, ref.dev_var)
    assert np.array_equal(w, g["prof_w"]) and np.array_equal(m, g["prof_m"]) and np.array_equal(c, g["prof_c"])
    assert w.shape == (24,) and m.shape == (24, 2) and c.shape == (24, 2, 2)
    w, m, c = celeste_fast.gen_galaxy_psf_mixture_params(
        g["mp_th"], g["mp_W"], g["mp_v"], g["mp_iw"], g["mp_im"], g["mp_ic"],
        ref.exp_amp, ref.exp_var, ref.dev_amp, ref.dev_var)
    assert np.array_equal(w, g["joint_w"]) and np.array_equal(m, g["joint_m"]) and np.array_equal(c, g["joint_c"])
    assert w.shape == (42,)
    # PSF-major ordering c = k * K_prof + j (celeste_fast.pyx:123-139): component 8 is PSF 1, profile 0
    w18, m18, _ = celeste_fast.gen_galaxy_prof_psf_mixture_params(
        g["mp_W"], g["mp_v"], g["mp_iw"], g["mp_im"], g["mp_ic"], ref.exp_amp, ref.exp_var)
    assert w18.shape == (18,) and w18[6] == g["mp_iw"][1] * ref.exp_amp[0]
    assert np.array_equal(m18[6], g["mp_v"] + g["mp_im"][1])


@pytest.mark.skipif(not (ref_import.available() and build_ref.built()),
                    reason="/root/reference is not mounted (GPU box): the committed vectors stand in")
def test_committed_vectors_are_what_the_reference_computes():
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        live = L.run(L.reference_backend())
        ru = ref_import.load_path("experiments/redshift/redshift_utils.py", "redshift_utils",
                                  dict(SPECTRA_LOC="", DR10QSO_LOC="", TRAIN_TEST_SPLIT_LOC="", EIGEN_FILE_LOC=""))
        live.update(L.run_a17(ru.project_to_bands, ru.softmax))
        live.update(L.sampler_traces(ref_import.load("util.infer.slicesample").slicesample,
                                     ref_import.load("util.infer.hmc").hmc))
    assert set(live) == set(GOLD.files)
    for k in GOLD.files:
        assert np.array_equal(GOLD[k], live[k]), k
    assert live and L.reference_backend().gen_point_source_psf_image.__code__.co_filename.startswith(
        ref_import.REF_ROOT)
