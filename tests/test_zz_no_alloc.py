"""cg_poisson_loglike_static: the Poisson reduction (sources.py:6-12) with its scratch in a file-scope
__device__ array -- an entry point that allocates nothing (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Called through the C-ABI and
compared with the existing entry point cg_poisson_loglike on two calls in a row (the second call meets
the first call's partials in the shared array).  Collected last on purpose
(test_zz_*)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(rng, n, masked):
    data = rng.poisson(50., size=n).astype(np.float64)
    model = rng.uniform(-2., 90., size=n)            # some pixels have model <= 0: excluded (sources.py:9)
    mask = (rng.uniform(size=n) > .1).astype(np.float64) if masked else None
    return data, model, mask


def test_static_scratch_entry_point_equals_existing_one(engine, oracle):
    import torch
    from desi_mcmc_b200 import _ffi
    L = engine.lib()
    dev = engine.device()
    rng = np.random.RandomState(5)
    out_new = torch.zeros(1, dtype=torch.float64, device=dev)
    out_old = torch.zeros(1, dtype=torch.float64, device=dev)
    # two calls in a row, the larger first: the second call reads fewer slots than the first one wrote
    for n, masked in ((2048 * 1489, True), (40 * 33, False), (300001, True)):
        data, model, mask = _case(rng, n, masked)
        d_d, d_m = torch.from_numpy(data).to(dev), torch.from_numpy(model).to(dev)
        d_k = torch.from_numpy(mask).to(dev) if masked else None
        kp = d_k.data_ptr() if masked else None
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        _ffi.check(L.cg_poisson_loglike_static(d_d.data_ptr(), d_m.data_ptr(), kp, n, out_new.data_ptr(),
                                               engine.stream_ptr()))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        _ffi.check(L.cg_poisson_loglike(d_d.data_ptr(), d_m.data_ptr(), kp, n, out_old.data_ptr(),
                                        engine.stream_ptr()))
        got, ref = float(out_new.cpu()[0]), float(out_old.cpu()[0])
        assert got == ref                                   # same kernels, same launch shape: same bits
        want = oracle.poisson_loglike(data, model, mask if masked else np.ones(n))
        assert abs(got - want) <= 1e-12 * abs(want)         # the bar of test_compat_mog_poisson_planck
    # n = 0: an empty sum; beyond the limit: a clear error, nothing launched
    _ffi.check(L.cg_poisson_loglike_static(d_d.data_ptr(), d_m.data_ptr(), None, 0, out_new.data_ptr(),
                                           engine.stream_ptr()))
    assert float(out_new.cpu()[0]) == 0.
    rc = L.cg_poisson_loglike_static(d_d.data_ptr(), d_m.data_ptr(), None, 1024 * 256 * 64 + 1, out_new.data_ptr(),
                                     engine.stream_ptr())
    assert rc != 0
    with pytest.raises(Exception, match="CG_POISSON_STATIC_MAX_N"):
        _ffi.check(rc)


def test_static_scratch_entry_point_under_graph_capture(engine):
    """What the entry point is for: it can be captured (the existing one may cudaMalloc its per-stream
    scratch on first use, which a capture forbids); two replays give the eager result."""
    import torch
    from desi_mcmc_b200 import _ffi
    L = engine.lib()
    dev = engine.device()
    rng = np.random.RandomState(6)
    data, model, mask = _case(rng, 123457, True)
    d_d, d_m, d_k = (torch.from_numpy(a).to(dev) for a in (data, model, mask))
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    eager = torch.zeros(1, dtype=torch.float64, device=dev)
    _ffi.check(L.cg_poisson_loglike_static(d_d.data_ptr(), d_m.data_ptr(), d_k.data_ptr(), data.size,
                                           eager.data_ptr(), engine.stream_ptr()))
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        _ffi.check(L.cg_poisson_loglike_static(d_d.data_ptr(), d_m.data_ptr(), d_k.data_ptr(), data.size,
                                               out.data_ptr(), engine.stream_ptr()))
    for _ in range(2):
        out.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert float(out.cpu()[0]) == float(eager.cpu()[0])
