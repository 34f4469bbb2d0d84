#!/usr/bin/env python
"""bench.py -- throughput of the B200 render + Poisson log-likelihood path.

Headline (the JSON line's `value`): BASELINE.json config 5 -- 64 synthetic Stripe-82 fields
(config 2 shape: 2048x1489 x 5 bands, 500 mixed sources, footprints capped at 101x101) sharded
round-robin over the ranks (STRONG scaling: 64 / 32 / 16 / 8 fields per GPU at 1 / 2 / 4 / 8
GPUs; `--fields-per-gpu F` switches to weak scaling).  A step = the full-field log-likelihood
of every field the rank owns (cg_fields_loglik, 8 fields per call) + the statistics the rank
contributes to the per-sweep all-reduce (sum of log-likelihoods, counts, sufficient statistics
of the star / galaxy colour priors: engine.SweepPayload), packed on the device straight into the
collective's send buffer.  The NCCL all-reduce of step k runs on a side stream under step k+1.

  value  component evaluations/s, source parameters already in HBM
  e2e    the same through the host API: packed source parameters copied pinned-host -> device
         every step, per-image log-likelihoods copied back and read on the host.  Image pixels
         stay resident (uploaded once, as FitsImage reads its frame once).

Further legs in the same line (rank 0): `config1` (25x25 stamp latency), `proposals` (config 3:
4096 proposals of one galaxy), `config4` (crowded-field Gibbs sweep), each with its own roofline
and CPU baseline.  `--impl reference` times the reference's own CPU route (compiled Cython
gmm_like_2d from oracle/_ref) with one thread AND all host threads on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W, N_SRC, STAR_FRAC, BOUND_CAP = 1489, 2048, 500, .6, 50.0
MAX_PATCH_PX = 103 * 103
METRIC = "source-pixel-component evals/s"
WORKLOAD = ("C5 = 64 x C2 full-field log-likelihood: synthetic Stripe-82 frames 2048x1489 x 5 bands, 500 mixed "
            "star/galaxy sources each (60% stars), footprints capped at 101x101")
PRIOR_K = 8


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region: NVML in-process
    (a sample every ~10 ms), nvidia-smi as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self._stop_ev = index, threading.Event()
        self.sm, self.mx, self.reasons, self.how = [], [], set(), "nvidia-smi"
        self.nv = self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].strip().isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv, self.how = pynvml, "nvml"
        except Exception:
            self.nv = None

    def _sample_nvml(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        if not self.mx:
            self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)))
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for nm, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                        ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(nm)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                              "--format=csv,noheader,nounits"], capture_output=True,
                             text=True, timeout=5).stdout.strip()
        r = [c.strip() for c in out.split(",")]
        if r and r[0].replace(".", "").isdigit():
            self.sm.append(float(r[0]))
        if len(r) > 1 and r[1].replace(".", "").isdigit():
            self.mx.append(float(r[1]))
        for nm, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[4:8]):
            if v.lower().startswith("active"):
                self.reasons.add(nm)

    def run(self):
        while not self._stop_ev.is_set():
            try:
                if self.nv is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop_ev.wait(0.01 if self.nv is not None else 0.1)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=6)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None,
                "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "how": self.how}


# ---------------------------------------------------------------------------
# CPU arms (the only places that execute oracle/)
# ---------------------------------------------------------------------------
def _oracle_field(field):
    from desi_mcmc_b200 import synth
    from oracle import ref_cpu as ref
    rng = np.random.RandomState(field["seed"])
    Hh, Ww = field["images"][0]["shape"]
    imgs = [ref.SyntheticImage(nelec=rng.poisson(430., size=(Hh, Ww)).astype(np.float64),
                               **synth.image_kwargs(m)) for m in field["images"]]
    srcs = [ref.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                          sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"]) for s in field["sources"]]
    return imgs, srcs


def cpu_baseline(field, budget_s):
    """The reference's fastest CPU route (compiled gmm_like_2d, OpenMP) on a bounded sample of field 0
    with ONE thread and with ALL host threads (B-fast at both counts); `value` is the
    all-thread rate."""
    from oracle import cpu_path
    be = cpu_path.backend()
    imgs, srcs = _oracle_field(field)
    n_all = cpu_path.host_threads()
    one = cpu_path.set_threads(1)
    r1 = cpu_path.timed_sample(imgs, srcs, BOUND_CAP, budget_s=budget_s / 3., be=be)
    used = cpu_path.set_threads(n_all) if be[0] == "reference" else 1
    r = cpu_path.timed_sample(imgs, srcs, BOUND_CAP, budget_s=budget_s * 2. / 3., be=be)
    how = "the compiled reference (oracle/_ref, OpenMP)" if be[0] == "reference" else "the NumPy port"
    return {"value": r["evals"] / r["seconds"], "unit": "evals/s", "cores": used,
            "host_cores": os.cpu_count(), "kind": be[0],
            "single_thread": {"value": r1["evals"] / r1["seconds"], "cores": one,
                              "sample": "first %d sources, %.3g evals in %.1f s" % (r1["n_src"], r1["evals"], r1["seconds"])},
            "sample": "field 0: first %d of %d sources x %d bands rendered with %s gmm_like_2d (float64) + "
                      "full-frame Poisson log-likelihood; %.3g evals in %.1f s on %d threads"
                      % (r["n_src"], len(srcs), r["n_img"], how, r["evals"], r["seconds"], used),
            "evals": r["evals"], "seconds": r["seconds"]}


def run_reference(args, rank):
    """--impl reference: CPU arm, rank 0 only."""
    if rank != 0:
        return
    from desi_mcmc_b200 import synth
    field = synth.make_field(1005, H, W, N_SRC, STAR_FRAC)
    per_step = max(3.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_baseline(field, per_step)
    ev, sec, ev1, sec1, last = 0, 0.0, 0.0, 0.0, None
    for _ in range(args.steps):
        last = cpu_baseline(field, per_step)
        ev += last["evals"]
        sec += last["seconds"]
    val = ev / sec
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "evals/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sec / max(1, args.steps),
            "higher_is_better": True, "scaling": "strong" if args.fields_per_gpu is None else "weak",
            "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": WORKLOAD, "sample_per_step": last["sample"]},
            "cpu_baseline": {"value": val, "unit": "evals/s", "cores": last["cores"], "kind": last["kind"],
                             "sample": last["sample"], "single_thread": last["single_thread"]},
            "e2e": {"value": val, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ---------------------------------------------------------------------------
# roofline of the render kernels, from live probes
# ---------------------------------------------------------------------------
def _peaks(engine):
    return {"xu": engine.pipe_probe(0), "fma": engine.pipe_probe(1), "fma2": engine.pipe_probe(7),
            "mix": engine.pipe_probe(6)}


def _roofline(name, evals, kernel_ms, launches, pk, fp32_per_eval, mufu_per_eval, slots_per_eval, unit_note, extra=None):
    """The lowest of three ceilings measured live on this GPU: FP32 pipe (operations/s / operations
    per eval), XU pipe (MUFU.EX2/s / MUFU per eval), instruction issue (FFMA instructions/s / issue
    slots per eval)."""
    bounds = {"fp32 pipe": max(pk["fma"], pk["fma2"]) / fp32_per_eval, "xu pipe": pk["xu"] / mufu_per_eval,
              "instruction issue": pk["fma"] / slots_per_eval}
    bname = min(bounds, key=bounds.get)
    achieved = evals / (kernel_ms * 1e-3) if kernel_ms > 0 else None
    r = {"kernel": name, "bound": bname, "achieved": achieved, "peak": bounds[bname],
         "unit": "evals/s (" + unit_note + ")", "frac": (achieved / bounds[bname]) if achieved else None,
         "bounds": bounds, "kernel_ms_total": kernel_ms, "kernel_launches": launches,
         "peak_source": "measured live on this GPU (cg_pipe_probe 0, 1, 7): MUFU.EX2 %.4g/s, FP32 FMA %.4g/s, as FFMA2 %.4g "
                        "operations/s" % (pk["xu"], pk["fma"], pk["fma2"])}
    if extra:
        r.update(extra)
    return r


# ---------------------------------------------------------------------------
# config 1: one 25x25 stamp (latency-bound; reported honestly as such)
# ---------------------------------------------------------------------------
def bench_stamp(engine, synth, with_cpu):
    import torch
    field = synth.make_stamp(1001, size=25)
    rng = np.random.RandomState(1001)
    nelec = [rng.poisson(450., size=(25, 25)).astype(np.float64) for _ in range(5)]
    imgs = [engine.Image(nelec=n, **synth.image_kwargs(m)) for n, m in zip(nelec, field["images"])]
    packed = engine.pack_sources([engine.SrcParams(**{k: s[k] for k in
                                  ("u", "a", "b", "t", "theta", "phi", "sigma", "rho", "fluxes")})
                                  for s in field["sources"]])
    fs = engine.FieldSet([imgs], [2])
    fs.set_sources([packed])
    sweep = engine.FieldSweep([fs], n_streams=1, io=True).capture()
    for _ in range(20):
        sweep.run()
    torch.cuda.synchronize()
    n = 200
    t0 = time.perf_counter()
    for _ in range(n):
        sweep.run()
        torch.cuda.current_stream().synchronize()
        fs.h_ll.sum()
    lat = (time.perf_counter() - t0) / n
    fs.loglik_device()
    evals = engine.last_eval_count()
    out = {"workload": "C1: 1 star + 1 galaxy on a 5-band 25x25 stamp, render + Poisson log-likelihood per call "
                       "(sources uploaded, log-likelihoods read back: one CUDA-graph replay + one stream sync)",
           "latency_us": lat * 1e6, "evals_per_call": evals, "value": evals / lat, "unit": "evals/s",
           "note": "latency-bound: %d evals are ~%.2f us of FP32 work; the call costs a graph launch and a sync" %
                   (evals, evals / 1.0e13 * 1e6)}
    if with_cpu:
        from oracle import cpu_path, ref_cpu as ref
        oimgs = [ref.SyntheticImage(nelec=n_, **synth.image_kwargs(m)) for n_, m in zip(nelec, field["images"])]
        osrcs = [ref.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                               sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"]) for s in field["sources"]]
        be = cpu_path.backend()
        cpu_path.set_threads(1)
        best = {"fast": 1e9, "live": 1e9}
        ev = {}
        for _ in range(6):              # 1 warm-up + best of 5
            t0 = time.perf_counter(); _, ev["fast"] = cpu_path.field_loglik(oimgs, osrcs, 0.0, be); t1 = time.perf_counter()
            _, ev["live"] = cpu_path.live_field_loglik(oimgs, osrcs, 0.0); t2 = time.perf_counter()
            best["fast"], best["live"] = min(best["fast"], t1 - t0), min(best["live"], t2 - t1)
        cpu_path.set_threads(cpu_path.host_threads())
        out["cpu_baseline"] = {"kind": be[0], "cores": 1, "unit": "evals/s",
                               "value": ev["fast"] / best["fast"], "latency_us": best["fast"] * 1e6,
                               "live_numpy_route": {"value": ev["live"] / best["live"], "latency_us": best["live"] * 1e6,
                                                    "what": "B-live: util/dists/mog.py:5-21,75-112 restated (oracle/ref_cpu.py), 1 thread"},
                               "sample": "the same stamp, best of 5 after 1 warm-up; B-fast = compiled gmm_like_2d route, 1 thread"}
    return out


# ---------------------------------------------------------------------------
# config 3: 4096 proposals of one galaxy
# ---------------------------------------------------------------------------
def bench_proposals(engine, synth, steps, pk, with_cpu):
    import torch
    P, size = 4096, 64
    field = synth.make_stamp(1003, size=size, n_star=1, n_gal=1)
    rng = np.random.RandomState(1003)
    dev = engine.device()
    nelec = [rng.poisson(450., size=(size, size)).astype(np.float64) for _ in range(5)]
    imgs = [engine.Image(nelec=n, **synth.image_kwargs(m)) for n, m in zip(nelec, field["images"])]
    cur = field["sources"][1]
    props, log_u = synth.make_proposals(1005, cur, P)
    packed = engine.pack_sources([engine.SrcParams(**{k: s[k] for k in
                                  ("u", "a", "b", "t", "theta", "phi", "sigma", "rho", "fluxes")})
                                  for s in [cur] + props])
    boxes = [(0, size, 0, size)] * 5
    bgs = [np.full((size, size), im.epsilon, dtype=np.float32) for im in imgs]
    sc = engine.ProposalScorer(imgs, boxes, bgs, max_proposals=P, guard=1e-2)
    for _ in range(3):
        delta, acc, ng = sc.score(packed, log_u)
    evals = engine.last_eval_count()
    d_src = engine.to_device(packed)
    d_u = torch.from_numpy(log_u).to(dev)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    engine.profile_render(True)
    e0.record()
    for _ in range(steps):
        sc.score_device(d_src, P, None, d_u)
    e1.record()
    torch.cuda.synchronize()
    dev_s = e0.elapsed_time(e1) * 1e-3
    k_ms, k_n = engine.profile_render_read()
    engine.profile_render(False)
    t0 = time.perf_counter()
    for _ in range(steps):
        sc.score(packed, log_u)
    e2e_s = time.perf_counter() - t0
    # a run in which the guard band fires: every proposal within 2.0 of its log-uniform is re-scored in float64
    wide = engine.ProposalScorer(imgs, boxes, bgs, max_proposals=P, guard=2.0)
    d2, a2, ng2 = wide.score(packed, log_u)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        d2, a2, ng2 = wide.score(packed, log_u)
    wide_s = (time.perf_counter() - t0) / 3
    out = {"metric": "MCMC proposals/s per field", "workload": "C3: 4096 proposals x 5-band 64x64 galaxy stamp, "
           "K=42 (dev 8 + exp 6 x 3-comp PSF), accept/reject on device from host log-uniforms",
           "value": P * steps / dev_s, "unit": "proposals/s", "evals_per_s": evals * steps / dev_s,
           "e2e": {"value": P * steps / e2e_s, "unit": "proposals/s",
                   "h2d_bytes_per_step": int(packed.nbytes + log_u.nbytes), "d2h_bytes_per_step": P * 9 + 4},
           "steps": steps, "accepted": int(acc.sum()), "guard_rescored": int(ng),
           "guard_band_run": {"guard": 2.0, "rescored_in_float64": int(ng2), "proposals_per_s": P / wide_s,
                              "decisions_equal_narrow_guard": bool(np.array_equal(a2, acc)),
                              "max_abs_delta_shift": float(np.max(np.abs(d2 - delta)))},
           "roofline": _roofline("render_kernel<float, POISSON>", evals * steps, k_ms, k_n, pk, 30. / 8., 4. / 8., 36. / 8.,
                                 "1 eval = 1 component at 1 pixel; per component and lane (8 px in 2 rows): 12 FFMA + 5 FMUL2 + "
                                 "4 FADD2 (30 FP32 operations) + 4 MUFU.EX2 + 2 LDS + pack/unpack MOVs = 36 issue slots")}
    if with_cpu:
        from oracle import cpu_path, ref_cpu as ref
        oimgs = [ref.SyntheticImage(nelec=n, **synth.image_kwargs(m)) for n, m in zip(nelec, field["images"])]
        mk = lambda s: ref.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                                     sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"])
        be = cpu_path.backend()
        used = cpu_path.set_threads(cpu_path.host_threads()) if be[0] == "reference" else 1
        r = cpu_path.timed_proposals(mk(cur), [mk(p) for p in props[:400]], oimgs, boxes,
                                     [b.astype(np.float64) for b in bgs], budget_s=6.0, be=be)
        out["cpu_baseline"] = {"value": r["proposals"] / r["seconds"], "unit": "proposals/s", "cores": used, "kind": be[0],
                               "evals_per_s": r["evals"] / r["seconds"],
                               "sample": "%d of the 4096 proposals scored one after the other (sources.py:275-306 around the "
                                         "compiled gmm_like_2d route), %.1f s" % (r["proposals"], r["seconds"])}
    return out


# ---------------------------------------------------------------------------
# config 4: crowded field, full Gibbs sweep on the device
# ---------------------------------------------------------------------------
def bench_crowded(engine, synth, steps, pk, with_cpu):
    import torch
    S, P = 5000, 16
    field = synth.make_crowded_field(1004, H=H, W=W, n_src=S)
    B, lam0 = synth.make_eigen_basis(1004)
    imgs, packed, eig, sw = synth.instantiate_crowded(engine, field, B, lam0, simulate_seed=1004, proposals=P, seed=1004)
    for _ in range(2):
        sw.sweep()
    sw.stats()
    rng = np.random.RandomState(4)
    lu = np.log(rng.uniform(size=S * P))
    phases = [("eigen_fluxes", lambda: eig.project()), ("allocate", lambda: sw.allocate(1)),
              ("epsilon", lambda: sw.resample_epsilon(2)), ("fluxes", lambda: sw.resample_fluxes(3)),
              ("propose", lambda: sw.propose(4)), ("score_conditional", lambda: sw.score(0, lu)),
              ("select", lambda: sw.select(0)), ("backgrounds", lambda: sw.backgrounds()),
              ("propose_2", lambda: sw.propose(5)), ("score_mh_background", lambda: sw.score(1, lu)),
              ("select_2", lambda: sw.select(1)), ("field_loglik", lambda: sw.loglik())]
    ph, evals, render_ms, render_n, render_evals = {}, {}, 0., 0, 0
    for name, fn in phases:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        prof = name.startswith("score")
        if prof:
            engine.profile_render(True)
        torch.cuda.synchronize()
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ph[name] = e0.elapsed_time(e1)
        if prof:
            ms, n = engine.profile_render_read()
            engine.profile_render(False)
            render_ms += ms; render_n += n
        if name in ("allocate", "score_conditional", "score_mh_background", "field_loglik"):
            evals[name] = engine.last_eval_count()
            if prof:
                render_evals += evals[name]
    # N2: the location slice samplers of all 5000 sources in one launch (not part of the sweep above)
    sw.allocate(7)
    sw.slice_locations(seed=10)            # warm-up (first launch sets the kernel's shared-memory attributes)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    status, sl_evals = sw.slice_locations(seed=11)
    torch.cuda.synchronize()
    slice_s = time.perf_counter() - t0
    slice_leg = {"what": "Source.resample_location (sources.py:308-319) for all 5000 sources in ONE launch "
                         "(cg_slice_sample_locations), host-supplied uniforms; wall clock incl. the upload of the uniforms",
                 "ms": slice_s * 1e3, "kernel_ms": sw.slice_stats["kernel_ms"],
                 "likelihood_evaluations": int(sl_evals.sum()),
                 "evaluations_per_s": float(sl_evals.sum() / (sw.slice_stats["kernel_ms"] * 1e-3)),
                 "uniforms_per_source": {"block": 256, "max_used": int(sw.slice_stats["used"].max())},
                 "sources_finished_on_device": int((status == 0).sum()),
                 "comparisons_decided_in_float64": int(sw.slice_stats["refined"].sum())}
    # HMC of every galaxy (celeste_mcmc.py:281-323) in one launch, a SHORT trajectory (2 leapfrog steps = 35 float64
    # likelihoods per galaxy instead of the reference's 50 steps = 563); the state is restored afterwards
    keep = sw.d_cur.clone()
    _, hmc_acc = sw.hmc_galaxies(seed=12, step=1e-9, n_steps=2)
    sw.d_cur.copy_(keep)
    hmc_leg = {"what": "hmc_sample_galaxy_params for all galaxies in ONE launch (cg_hmc_galaxy_params), 2 leapfrog steps; "
                       "a likelihood = the galaxy's 5 capped footprints x 42 components in float64, no culling",
               "galaxies": int(hmc_acc.size), "leapfrog_steps": 2, "kernel_ms": sw.hmc_stats["kernel_ms"],
               "float64_likelihoods": int(sw.hmc_stats["likelihoods"]),
               "likelihoods_per_s": float(sw.hmc_stats["likelihoods"] / (sw.hmc_stats["kernel_ms"] * 1e-3)),
               "accepted": int(hmc_acc.sum())}
    # backgrounds: every active neighbour rendered on the target's box (sources.py:481)
    offs, idx = sw.h_nbr
    st = sw.h_static
    Ksrc = np.where(packed["type"] == 0, 3, 42).astype(np.int64)
    box_px = ((st["x1"] - st["x0"]).astype(np.int64) * (st["y1"] - st["y0"])).reshape(-1)
    ksum = np.add.reduceat(np.concatenate([Ksrc[idx], [0]]), np.minimum(offs[:-1], idx.size))
    ksum[offs[1:] == offs[:-1]] = 0
    evals["backgrounds"] = int((box_px * ksum).sum())
    evals_sweep = sum(evals.values())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        sw.sweep()
    e1.record()
    res = sw.stats()
    wall = (time.perf_counter() - t0) / steps
    ms = e0.elapsed_time(e1) / steps
    out = {"workload": "C4: 5000 sources (10% black-body stars, 5% eigen-spectrum point sources, galaxies) with footprints "
                       "forced to ~50x50 on a 2048x1489x5 field; one sweep = eigen fluxes -> photon allocation -> epsilon, "
                       "fluxes ~ Gamma -> 16 proposals/source scored conditionally on the photons -> backgrounds -> 16 "
                       "proposals/source scored by the full MH ratio -> field log-likelihood, all on the device",
           "sweep_ms": ms, "sweep_ms_host_clock": wall * 1e3, "value": evals_sweep / (ms * 1e-3), "unit": "evals/s",
           "proposals_per_s": 2 * S * P / (ms * 1e-3), "evals_per_sweep": evals_sweep, "evals_by_phase": evals,
           "phase_ms": {k: round(v, 3) for k, v in ph.items()}, "workspace_bytes": int(sw.workspace_bytes),
           "slice_sampler_on_device": slice_leg, "hmc_on_device": hmc_leg,
           "h2d_bytes_per_sweep": 2 * S * P * 8, "accepted_last_sweep": res["accepted"].tolist(),
           "guard_rescored_last_sweep": res["guard"].tolist(),
           "roofline": _roofline("render_kernel<float, COND | POISSON> (the two scoring passes: %.0f%% of the sweep)" %
                                 (100. * render_ms / sum(ph.values())), render_evals, render_ms, render_n, pk,
                                 30. / 8., 4. / 8., 36. / 8., "as the proposals leg")}
    if with_cpu:
        from oracle import cpu_path, ref_cpu as ref
        from desi_mcmc_b200 import _ffi
        ref.set_filter_tables_npz(os.path.join(_ffi.PKG, "data", "filter_curves.npz"))
        be = cpu_path.backend()
        used = cpu_path.set_threads(cpu_path.host_threads()) if be[0] == "reference" else 1
        states = sw.states()
        oimgs = [ref.SyntheticImage(nelec=im.nelec.cpu().numpy().astype(np.float64), **synth.image_kwargs(m))
                 for im, m in zip(imgs, field["images"])]
        bg = sw.bg
        picks = [k for k in range(S) if packed["flags"][k] == 0][:40]
        t_tot, n_prop, ev = 0., 0, 0

        def mk(rec):
            return ref.SrcParams(u=rec["u"].copy(), a=int(rec["type"]), theta=float(rec["theta"]), phi=float(rec["phi"]),
                                 sigma=float(rec["sigma"]), rho=float(rec["rho"]), fluxes=rec["flux"].copy())
        for k in picks:
            bxs = [(int(b["x0"]), int(b["x1"]), int(b["y0"]), int(b["y1"])) for b in st[k]]
            bgs = [bg[k * 5 + b, :(bx[1] - bx[0]) * (bx[3] - bx[2])].cpu().numpy().astype(np.float64).reshape(bx[3] - bx[2], bx[1] - bx[0])
                   for b, bx in enumerate(bxs)]
            r = cpu_path.timed_proposals(mk(states[k, 0]), [mk(states[k, p]) for p in range(1, P + 1)], oimgs, bxs, bgs,
                                         budget_s=0.5, be=be)
            t_tot += r["seconds"]; n_prop += r["proposals"]; ev += r["evals"]
            if t_tot > 10.:
                break
        out["cpu_baseline"] = {"value": n_prop / t_tot, "unit": "proposals/s", "cores": used, "kind": be[0],
                               "evals_per_s": ev / t_tot,
                               "sample": "%d MH proposals of the first free-flux sources scored on their cached 50x50 boxes "
                                         "with the backgrounds of this run (sources.py:275-306 around the compiled gmm_like_2d "
                                         "route), %.1f s" % (n_prop, t_tot)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--total-fields", type=int, default=64,
                    help="config 5: fields sharded round-robin over the ranks (strong scaling)")
    ap.add_argument("--fields-per-gpu", type=int, default=None,
                    help="weak scaling instead: this many fields on every rank")
    ap.add_argument("--fields-per-call", type=int, default=8,
                    help="fields batched into one cg_fields_loglik call")
    ap.add_argument("--streams", type=int, default=4,
                    help="side streams the field sets of a step are spread over (the set-up kernels of one set run in the tail of "
                         "the previous set's tile kernel; the roofline leg always times the kernel alone on one stream)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step kernel by kernel")
    ap.add_argument("--no-overlap", action="store_true", help="all-reduce on the compute stream (exposed)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-proposals", action="store_true")
    ap.add_argument("--no-crowded", action="store_true")
    ap.add_argument("--no-stamp", action="store_true")
    ap.add_argument("--main-only", action="store_true", help="skip every secondary leg and CPU baseline")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.main_only:
        args.no_cpu_baseline = args.no_proposals = args.no_crowded = args.no_stamp = True

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    # rank 0 prints exactly ONE line on stdout: library banners (NCCL's version
    # line goes to fd 1) are diverted to stderr for the whole run
    out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from desi_mcmc_b200 import engine, synth, _ffi

    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    engine.init(local)
    if os.environ.get("CG_CULL"):
        engine.set_option(1, int(os.environ["CG_CULL"]))
    if os.environ.get("CG_RECUR"):                       # A/B switch of the row recurrence (CG_OPT_ROW_RECURRENCE)
        engine.set_option(2, int(os.environ["CG_RECUR"]))
    dev = engine.device()
    weak = args.fields_per_gpu is not None
    if weak:
        my_fields = [rank * args.fields_per_gpu + f for f in range(args.fields_per_gpu)]
        total_fields = args.fields_per_gpu * world
    else:
        total_fields = args.total_fields
        my_fields = list(range(rank, total_fields, world))             # dist.shard_fields
    F = len(my_fields)
    if F < 1:
        raise SystemExit("rank %d owns no field: --total-fields must be >= the number of GPUs" % rank)

    # ---- synthetic fields of this rank (config 5 seeds 1005 + f), pixels drawn on device
    fields, images, srcs_host = [], [], []
    gen = torch.Generator(device=dev)
    for fid in my_fields:
        fld = synth.make_field(1005 + fid, H, W, N_SRC, STAR_FRAC)
        imgs = [engine.Image(nelec=torch.zeros((H, W), dtype=torch.float32, device=dev),
                             **synth.image_kwargs(m)) for m in fld["images"]]
        packed = engine.pack_sources([engine.SrcParams(u=s["u"], a=s["a"], theta=s["theta"], phi=s["phi"],
                                                       sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"])
                                      for s in fld["sources"]])
        _, models = engine.field_loglik(imgs, packed, bound_cap=BOUND_CAP, max_patch_px=MAX_PATCH_PX,
                                        return_model=True)
        gen.manual_seed(1005 + fid)
        for im, lam in zip(imgs, models):
            im.set_nelec(torch.poisson(lam, generator=gen))
        del models
        fields.append(fld)
        images.append(imgs)
        srcs_host.append(packed)
    engine._state["ws"].clear()
    torch.cuda.synchronize()

    # field sets: FPC fields per cg_fields_loglik call, all viewing ONE source array and ONE
    # log-likelihood vector of the rank (the payload kernels read them as a whole)
    FPC = max(1, min(args.fields_per_call, F, _ffi.CG_MAX_FIELDS))
    groups = [list(range(g, min(g + FPC, F))) for g in range(0, F, FPC)]
    isz = _ffi.SRC_DTYPE.itemsize
    rank_srcs = torch.zeros(F * N_SRC * isz, dtype=torch.uint8, device=dev)
    rank_hsrcs = torch.zeros(F * N_SRC * isz, dtype=torch.uint8).pin_memory()
    rank_ll = torch.zeros(F * 5, dtype=torch.float64, device=dev)
    sets = []
    for g in groups:
        a, b = g[0] * N_SRC * isz, (g[-1] + 1) * N_SRC * isz
        sets.append(engine.FieldSet([images[f] for f in g], [N_SRC] * len(g), bound_cap=BOUND_CAP,
                                    max_patch_px=MAX_PATCH_PX, d_srcs=rank_srcs[a:b], h_srcs=rank_hsrcs[a:b],
                                    ll=rank_ll[g[0] * 5:(g[-1] + 1) * 5]))
    for fs, g in zip(sets, groups):
        fs.set_sources([srcs_host[f] for f in g])

    # the rank's contribution to the per-sweep all-reduce (dist.SweepStats layout, ~2.7 KB)
    prng = np.random.RandomState(7)

    def colour_prior():
        means = prng.normal(0., .6, size=(PRIOR_K, 5)) + np.array([0., 0., 0., 0., 2.5])
        A = prng.normal(size=(PRIOR_K, 5, 5)) * .3
        return means, np.einsum("kij,klj->kil", A, A) + .2 * np.eye(5), prng.dirichlet(np.ones(PRIOR_K))
    payload = engine.SweepPayload(rank_srcs, F * N_SRC, rank_ll, colour_prior(), colour_prior(), n_buffers=2)
    side = torch.cuda.Stream(device=dev)
    ar_done = [torch.cuda.Event() for _ in range(2)]
    state = {"k": 0}

    sweeps = [engine.FieldSweep(sets, n_streams=args.streams, parallel=payload.prior_stats,
                                after=(lambda k=k: payload.pack(k))) for k in range(2)]
    sweeps_e2e = [engine.FieldSweep(sets, n_streams=args.streams, parallel=payload.prior_stats,
                                    after=(lambda k=k: payload.pack(k)), io=True) for k in range(2)]
    sweep_alone = engine.FieldSweep(sets, n_streams=1, parallel=payload.prior_stats, after=(lambda: payload.pack(0)))

    def exchange(k):
        """NCCL all-reduce of send buffer k on the side stream: it overlaps the next step."""
        if world == 1:
            return
        cur = torch.cuda.current_stream()
        if args.no_overlap:
            dist.all_reduce(payload.vec[k])
            return
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            dist.all_reduce(payload.vec[k])
            ar_done[k].record(side)

    def step(which):
        k = state["k"] & 1
        state["k"] += 1
        if world > 1 and not args.no_overlap:
            torch.cuda.current_stream().wait_event(ar_done[k])     # buffer k is free again (step - 2 reduced)
        which[k].run()
        exchange(k)

    # evaluations per step (constant): count once per set
    evals_step, issued_step = 0, 0
    for fs in sets:
        fs.loglik_device()
        evals_step += engine.last_eval_count()
        issued_step += engine.last_issued_count()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- roofline leg: the same step launched kernel by kernel on one stream, every
    # field_tile_kernel launch bracketed by CUDA events on that stream (events cannot
    # be timed inside a graph replay, and overlapped kernels would share the SMs)
    for _ in range(args.warmup):
        sweep_alone._enqueue_compute()
    barrier()
    engine.launch_count(reset=True)
    engine.profile_render(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        sweep_alone._enqueue_compute()
    e1.record()
    barrier()
    serial_ms = e0.elapsed_time(e1)
    launches = engine.launch_count()        # kernels of args.steps steps (a graph replay launches the same ones)
    render_ms, render_n = engine.profile_render_read()
    engine.profile_render(False)
    ll_serial = rank_ll.cpu().numpy().copy()

    # ---- device-resident leg (the headline `value`): the captured step
    if not args.no_graph:
        try:
            for sw in sweeps:
                sw.capture()
        except Exception as exc:
            sys.stderr.write("CUDA graph capture failed (%s); launching kernel by kernel\n" % exc)
            for sw in sweeps:
                sw.graph = None
            args.no_graph = True
            torch.cuda.synchronize()
    for _ in range(args.warmup):
        step(sweeps)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0.record()
    for _ in range(args.steps):
        step(sweeps)
    e1.record()
    barrier()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    ll_ref = rank_ll.cpu().numpy().copy()
    assert np.array_equal(ll_serial, ll_ref), "graph replay and serial launches differ"
    stats_vec = payload.vec[(state["k"] - 1) & 1].cpu().numpy().copy()

    # ---- end-to-end leg: host source arrays -> pinned -> device, ll -> host, every step
    def step_e2e():
        # the sampler's state lives in the pinned stage (pinned_sources); upload, the field kernels,
        # the payload and the download are one graph replay; the host reads the log-likelihoods
        step(sweeps_e2e)
        torch.cuda.current_stream().synchronize()
        return float(sum(fs.h_ll.sum() for fs in sets))

    for fs, g in zip(sets, groups):       # host state written once into the pinned views
        for k, f in enumerate(g):
            fs.pinned_sources(k, srcs_host[f].shape[0])[:] = srcs_host[f]
    if not args.no_graph:
        for sw in sweeps:
            sw.graph = None
        try:
            for sw in sweeps_e2e:
                sw.capture()
        except Exception as exc:
            sys.stderr.write("CUDA graph capture of the end-to-end step failed (%s)\n" % exc)
            for sw in sweeps_e2e:
                sw.graph = None
            torch.cuda.synchronize()
    for _ in range(args.warmup):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        total_ll = step_e2e()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    clocks = sampler.stop()
    ll_e2e = np.concatenate([fs.h_ll.numpy() for fs in sets])
    assert np.array_equal(ll_e2e, ll_ref), "e2e and device-resident results differ"
    if world > 1:       # the all-reduced vector: its log-likelihood entry is the sum over ALL ranks' images
        torch.cuda.synchronize()
        chk = torch.tensor([float(ll_ref.sum())], dtype=torch.float64, device=dev)
        dist.all_reduce(chk)
        got = float(payload.vec[(state["k"] - 1) & 1][0].item())
        assert abs(got - float(chk.item())) <= 1e-9 * abs(got), "all-reduced log-likelihood != sum over ranks"

    total_evals = evals_step * world          # every rank holds the same shapes, different seeds
    if world > 1:
        t = torch.tensor([float(evals_step)], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        total_evals = float(t.item())
    value = total_evals * args.steps / (dev_ms * 1e-3)
    e2e_value = total_evals * args.steps / e2e_s

    if rank == 0:
        pk = _peaks(engine)
        legs = {}
        with_cpu = not args.no_cpu_baseline and world == 1
        if not args.no_stamp:
            legs["config1"] = bench_stamp(engine, synth, with_cpu)
        if not args.no_proposals:
            legs["proposals"] = bench_proposals(engine, synth, max(3, min(args.steps, 10)), pk, with_cpu)
        if not args.no_crowded:
            legs["config4"] = bench_crowded(engine, synth, max(2, min(args.steps, 5)), pk, with_cpu)
        # Per component and lane (8 pixels) the packed row-recurrence loop (ft_eval2r, SASS of
        # recur_probe_kernel) issues 1 FFMA + 6 FFMA2 + 1 FMUL + 4 FMUL2 + 3 FADD2 (28 FP32
        # operations per lane in 15 instructions) + 3 MUFU.EX2 + 2 LDS.128 + 2 MOV = 22 issue slots.
        sm = engine.sm_count()
        traffic = None
        try:        # dram__bytes_read.sum + dram__bytes_write.sum per 8-field launch, written by tools/ncu_traffic.py
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            traffic = {"bytes_per_launch": tj["dram_bytes_per_launch"] * FPC / float(tj["fields_per_launch"]),
                       "source": tj["source"]}
        except Exception:
            pass
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        roof = _roofline("field_tile_kernel", evals_step * args.steps, render_ms, render_n, pk, 28. / 8., 3. / 8., 22. / 8.,
                         "1 eval = 1 component at 1 pixel = 3.5 FP32 operations (1.875 instructions, packed FFMA2/FMUL2/FADD2) "
                         "+ 0.375 MUFU.EX2 + 0.25 LDS.128 + 0.25 MOV: 2.75 issue slots",
                         {"issued": (issued_step * args.steps / (render_ms * 1e-3)) if render_ms > 0 else None,
                          "note": "achieved = algorithmic evals (footprint pixels x components, SURVEY 8d) / kernel time; "
                                  "issued = lane-evaluations executed after per-warp component culling",
                          "xu_peak": pk["xu"],
                          "mix_ceiling": {"value": pk["mix"], "note": "evals/s of the inner loop (ft_eval2r) in isolation, cg_pipe_probe(6)"},
                          "kernel_share_of_step": render_ms / serial_ms if serial_ms else None,
                          "serial_ms_per_step": serial_ms / args.steps,
                          "traffic": traffic["bytes_per_launch"] if traffic else None,
                          "traffic_source": traffic["source"] if traffic else "no ncu capture summarised in profiles/traffic.json",
                          "traffic_algorithmic": FPC * 5 * H * W * 4,
                          "hbm_peak_gbs": peaks.get("hbm_gbs")})
        roof["issued_frac"] = (roof["issued"] / roof["peak"]) if roof.get("issued") else None
        line = {
            "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak" if weak else "strong", "vs_baseline": None,
            "dtype": "f32 rates, f64 log-likelihood accumulation", "data": "synthetic",
            "config": {"workload": WORKLOAD, "fields_per_gpu": F, "fields_total": total_fields,
                       "fields_per_call": FPC, "streams": sweeps[0].n_streams,
                       "launch": "kernel by kernel" if args.no_graph else "one CUDA graph replay per step (+ the NCCL all-reduce on a side stream)",
                       "evals_per_step_per_gpu": evals_step, "issued_lane_evals_per_step_per_gpu": issued_step,
                       "pixels_per_step_per_gpu": F * 5 * H * W,
                       "l2": "inputs larger than L2 (%d MB of pixel data per GPU per step)" % (F * 5 * H * W * 4 >> 20),
                       "e2e_inputs": "packed source parameters copied pinned host -> device every step; image pixels resident",
                       "allreduce": {"doubles": payload.size, "bytes": payload.size * 8,
                                     "content": "sum ll | sky photons per band | n_star, n_gal | colour-prior sufficient statistics "
                                                "(2 priors x K=%d x (1 + 5 + 15)), packed on the device by cg_sweep_stats" % PRIOR_K,
                                     "overlap": "none (single GPU)" if world == 1 else
                                     ("on the compute stream" if args.no_overlap else "side stream, double-buffered: step k's all-reduce runs under step k+1")},
                       "parallelism": ("fields sharded round-robin over ranks (%s scaling); one float64 all-reduce per step" %
                                       ("weak" if weak else "strong")) if world > 1 else "single GPU"},
            "e2e": {"value": e2e_value, "unit": "evals/s", "ms_per_step": 1e3 * e2e_s / args.steps,
                    "h2d_bytes_per_step": F * N_SRC * _ffi.SRC_DTYPE.itemsize, "d2h_bytes_per_step": F * 5 * 8},
            "gpu_launches": launches,
            "roofline": roof,
            "clocks": clocks,
            "total_loglik": total_ll,
            "sweep_stats": {"sum_ll": float(stats_vec[0]), "n_star": float(stats_vec[6]), "n_gal": float(stats_vec[7]),
                            "prior_mass": float(stats_vec[8:8 + PRIOR_K].sum())},
        }
        line.update(legs)
        if with_cpu:
            line["cpu_baseline"] = {k: v for k, v in cpu_baseline(fields[0], args.cpu_seconds).items()
                                    if k not in ("evals", "seconds")}
        out.write(json.dumps(line) + "\n")
        out.flush()
    if world > 1:
        # Tear-down: graphs first, then the communicator.  A stalled tear-down must be VISIBLE: the
        # watchdog ends the process with a non-zero status.
        wd = threading.Timer(300.0, lambda: (sys.stderr.write("NCCL tear-down stalled\n"), os._exit(3)))
        wd.daemon = True
        wd.start()
        torch.cuda.synchronize()
        for sw in sweeps + sweeps_e2e:
            sw.graph = None
        dist.barrier()
        torch.cuda.synchronize()
        dist.destroy_process_group()
        wd.cancel()


if __name__ == "__main__":
    main()
