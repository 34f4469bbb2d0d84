"""CPU suite, part 2: the C-ABI library loads and exports every declared
symbol (no compute without a GPU), host packing logic, fails-loudly behaviour,
and the N>1 path on gloo with world_size 2."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from desi_mcmc_b200 import _ffi, build
    build.build()
    return _ffi.load()


def test_every_declared_symbol_is_exported(lib):
    from desi_mcmc_b200 import _ffi
    hdr = open(os.path.join(ROOT, "include", "celeste_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(cg_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), "missing export " + name
    assert declared == set(_ffi.PROTOTYPES), declared ^ set(_ffi.PROTOTYPES)
    assert lib.cg_abi_version() == _ffi.CG_ABI_VERSION


def test_struct_layouts_match_header(lib):
    from desi_mcmc_b200 import _ffi
    assert C.sizeof(_ffi.CgImageDesc) == 312
    assert _ffi.SRC_DTYPE.itemsize == 120 and _ffi.SRC_DTYPE.fields["u"][1] == 8
    assert _ffi.ITEM_DTYPE.itemsize == 32 and _ffi.BOX_DTYPE.itemsize == 16


def test_host_entry_points_without_gpu(lib):
    """Pure-host ABI calls: bounding radius (util/bound/bounding_box.py:9-31)
    against the oracle; uninitialised engine calls fail with CG_ESTATE."""
    from desi_mcmc_b200 import _ffi
    from oracle import ref_cpu as ref
    rng = np.random.RandomState(2)
    K = 42
    means = rng.normal(size=(K, 2)) + 30.
    covs = np.zeros((K, 2, 2))
    for i in range(K):
        a = rng.normal(size=(2, 2))
        covs[i] = a.dot(a.T) + np.eye(2)
    center = np.array([30., 30.])
    for err in (1e-3, 1e-5):
        got = lib.cg_calc_bounding_radius(means.ctypes.data, covs.ctypes.data, K, err, center.ctypes.data)
        want = ref.calc_bounding_radius(np.ones(K), means, covs, err, center=center)
        assert abs(got - want) <= 1e-12 * want
    if not torch.cuda.is_available():
        out = C.c_double()
        rc = lib.cg_pipe_probe(0, C.byref(out), None)
        assert rc == _ffi.CG_ESTATE
        assert "cg_init" in _ffi.last_error()
        assert lib.cg_init(0) == _ffi.CG_ECUDA             # no device: loud failure, no fallback
        assert "no CPU fallback" in _ffi.last_error()


def test_engine_fails_loudly_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from desi_mcmc_b200 import engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        engine.init(0)


def test_pack_sources():
    from desi_mcmc_b200 import engine, _ffi
    s0 = engine.SrcParams(u=[1., 2.], a=0, fluxes=dict(u=1, g=2, r=3, i=4, z=5))
    s1 = engine.SrcParams(u=[3., 4.], a=1, theta=.3, sigma=2., phi=40., rho=.6, fluxes=np.arange(5.))
    s2 = engine.SrcParams(u=[5., 6.], a=0, t=5000., b=1e-10, fluxes=np.ones(5))
    arr = engine.pack_sources([s0, s1, s2])
    assert arr.dtype == _ffi.SRC_DTYPE and arr.shape == (3,)
    assert list(arr["type"]) == [0, 1, 0]
    assert list(arr["flux"][0]) == [1, 2, 3, 4, 5]
    assert arr["flags"][2] == _ffi.CG_SRC_PLANCK and arr["flags"][0] == 0
    assert (arr["theta"][1], arr["sigma"][1], arr["phi"][1], arr["rho"][1]) == (.3, 2., 40., .6)
    with pytest.raises(ValueError):
        engine.pack_sources([engine.SrcParams(u=[0, 0], a=2)])
    v = engine.pack_source_fields(np.array([1, 1]), np.zeros((2, 2)), np.ones((2, 5)), theta=[.1, .2],
                                  sigma=[1, 2], phi=[0, 1], rho=[.5, .6])
    assert list(v["theta"]) == [.1, .2]


def test_shard_fields_and_stats_layout():
    from desi_mcmc_b200 import dist as D
    owned = [D.shard_fields(64, r, 8) for r in range(8)]
    assert sorted(sum(owned, [])) == list(range(64)) and all(len(o) == 8 for o in owned)
    assert D.shard_fields(5, 1, 2) == [1, 3]
    st = D.SweepStats(n_prior_comp=3)
    rng = np.random.RandomState(0)
    resp = rng.dirichlet(np.ones(3), size=10)
    col = rng.normal(size=(10, 5))
    st.add_prior_stats(1, resp, col)
    s0, s1, s2 = st.prior_stats(1)
    assert np.allclose(s0, resp.sum(0)) and np.allclose(s1, resp.T @ col)
    iu = np.triu_indices(5)
    full = np.einsum("sk,si,sj->kij", resp, col, col)
    assert np.allclose(s2, full[:, iu[0], iu[1]])
    assert np.all(st.prior_stats(0)[0] == 0)


WORKER = r"""
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch
import torch.distributed as dist
from desi_mcmc_b200 import dist as D
D.init_process_group("gloo")
rank, world = D.rank_world()
assert world == 2
fields = D.shard_fields(7, rank, world)
st = D.SweepStats(n_prior_comp=2)
# per-field "log-likelihoods" stand in for the GPU results: value = field id + 1
st.add_loglik(torch.tensor([float(f + 1) for f in fields], dtype=torch.float64))
st.add_noise([float(rank + 1)] * 5)
st.add_counts(3 * len(fields), 2 * len(fields))
rng = np.random.RandomState(rank)
st.add_prior_stats(0, rng.dirichlet(np.ones(2), size=4), np.ones((4, 5)))
v = st.all_reduce()
assert v[0].item() == sum(range(1, 8)), v[0]
assert v[1].item() == 3.0 and v[6].item() == 21 and v[7].item() == 14
assert abs(st.prior_stats(0)[0].sum() - 8.0) < 1e-12
m = D.max_over_ranks(float(rank) + .5)
assert m == 1.5
dist.barrier()
dist.destroy_process_group()
open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ok_" + str(rank)), "w").write("ok")
"""


def test_gloo_world_size_2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29621", str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert (tmp_path / "ok_0").exists() and (tmp_path / "ok_1").exists()


def test_bench_reference_arm_runs():
    """`bench.py --impl reference` prints one JSON line with the contract keys."""
    import json
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    line = json.loads(res.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["unit"] == "evals/s"
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["e2e"]["h2d_bytes_per_step"] == 0


def test_static_boxes_and_active_neighbours_follow_the_reference():
    """Host logic of the crowded-field sweep: engine.static_boxes == make_bbox_dict (sources.py:434-455,
    oracle.make_bbox) and engine.active_neighbours == get_active_sources' intersection test
    (sources.py:457-472), brute force."""
    import types
    from desi_mcmc_b200 import engine, synth
    from oracle import ref_cpu as ref
    field = synth.make_crowded_field(5, H=300, W=340, n_src=150)
    imgs_o = [ref.SyntheticImage(nelec=np.zeros((300, 340)), **synth.image_kwargs(m)) for m in field["images"]]
    imgs = [types.SimpleNamespace(phi_n=o.phi_n, rho_n=o.rho_n, Ups_n_inv=o.Ups_n_inv, shape=(300, 340)) for o in imgs_o]
    u = np.array([s["u"] for s in field["sources"]])
    boxes = engine.static_boxes(imgs, u, 25.)
    for k in range(0, 150, 7):
        src = ref.SrcParams(u=u[k], a=0, fluxes=np.ones(5))
        for b, o in enumerate(imgs_o):
            xlim, ylim = ref.make_bbox(src, o, 25.)
            bx = boxes[k, b]
            assert (bx["x0"], bx["x1"], bx["y0"], bx["y1"]) == (xlim[0], xlim[1], ylim[0], ylim[1])
    offs, idx = engine.active_neighbours(boxes)
    assert offs.shape == (150 * 5 + 1,) and offs[-1] == idx.size and idx.size > 150
    for k in range(0, 150, 11):
        for b in range(5):
            a = boxes[k, b]
            want = [t for t in range(150) if t != k and
                    abs(int(a["x0"]) - int(boxes[t, b]["x0"])) * 2 < (a["x1"] - a["x0"]) + (boxes[t, b]["x1"] - boxes[t, b]["x0"]) and
                    abs(int(a["y0"]) - int(boxes[t, b]["y0"])) * 2 < (a["y1"] - a["y0"]) + (boxes[t, b]["y1"] - boxes[t, b]["y0"])]
            assert list(idx[offs[k * 5 + b]:offs[k * 5 + b + 1]]) == want
