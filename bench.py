#!/usr/bin/env python
"""bench.py -- throughput of the B200 render + Poisson log-likelihood path.

Step = one full-field log-likelihood pass (BASELINE.json config 2: synthetic
Stripe-82 frame 2048x1489 x 5 bands, 500 mixed sources, footprints capped at
101x101) over every field this rank owns.  Fields are independent, so ranks
shard them with no pixel traffic (weak scaling, F fields per GPU; 8 GPUs x 8 =
the 64 fields of config 5) and exchange one small float64 all-reduce per step
(global log-likelihood + prior sufficient statistics).

  value  component evaluations/s, source parameters already in HBM
  e2e    the same through the host API: packed source parameters copied
         pinned-host -> device every step, per-image log-likelihoods copied
         back.  Image pixels stay resident (uploaded once when the Image is
         constructed, as FitsImage reads its frame once).

`--impl reference` times the reference's own CPU route (compiled Cython
gmm_like_2d from oracle/_ref, all host threads) on a bounded sample.
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
WORKLOAD = ("C2 full-field log-likelihood: synthetic Stripe-82 frame 2048x1489 x 5 bands, 500 mixed "
            "star/galaxy sources (60% stars), footprints capped at 101x101")


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


def cpu_baseline(field, budget_s):
    """The reference's CPU route on a bounded sample of field 0 (rank 0 only)."""
    from desi_mcmc_b200 import synth
    from oracle import cpu_path, ref_cpu as ref
    be = cpu_path.backend()
    rng = np.random.RandomState(field["seed"])
    imgs = [ref.SyntheticImage(nelec=rng.poisson(430., size=(H, W)).astype(np.float64),
                               **synth.image_kwargs(m)) for m in field["images"]]
    srcs = [ref.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                          sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"]) for s in field["sources"]]
    r = cpu_path.timed_sample(imgs, srcs, BOUND_CAP, budget_s=budget_s, be=be)
    return {"value": r["evals"] / r["seconds"], "unit": "evals/s", "cores": cpu_path.threads_used(be[0]),
            "host_cores": os.cpu_count(), "kind": be[0],
            "sample": "field 0: first %d of %d sources x %d bands rendered with %s gmm_like_2d "
                      "(float64) + full-frame Poisson log-likelihood; %.3g evals in %.1f s"
                      % (r["n_src"], len(srcs), r["n_img"],
                         "the compiled reference (oracle/_ref, OpenMP)" if be[0] == "reference"
                         else "the NumPy port", r["evals"], r["seconds"]),
            "evals": r["evals"], "seconds": r["seconds"]}


def bench_proposals(engine, synth, steps):
    """Second half of BASELINE.json's metric: MCMC proposals/s (config 3: 4096
    proposals of one galaxy on a 5-band 64x64 stamp, K = 42), device-resident and
    end to end (host proposal arrays in, accept decisions out)."""
    import torch
    P, size = 4096, 64
    field = synth.make_stamp(1003, size=size, n_star=1, n_gal=1)
    rng = np.random.RandomState(1003)
    dev = engine.device()
    imgs = [engine.Image(nelec=torch.from_numpy(rng.poisson(450., size=(size, size)).astype(np.float32)),
                         **synth.image_kwargs(m)) for m in field["images"]]
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
    e0.record()
    for _ in range(steps):
        sc.score_device(d_src, P, None, d_u)
    e1.record()
    torch.cuda.synchronize()
    dev_s = e0.elapsed_time(e1) * 1e-3
    t0 = time.perf_counter()
    for _ in range(steps):
        sc.score(packed, log_u)
    e2e_s = time.perf_counter() - t0
    return {"metric": "MCMC proposals/s per field", "workload": "C3: 4096 proposals x 5-band 64x64 galaxy stamp, "
            "K=42 (dev 8 + exp 6 x 3-comp PSF), accept/reject on device from host log-uniforms",
            "value": P * steps / dev_s, "unit": "proposals/s", "evals_per_s": evals * steps / dev_s,
            "e2e": {"value": P * steps / e2e_s, "unit": "proposals/s",
                    "h2d_bytes_per_step": int(packed.nbytes + log_u.nbytes), "d2h_bytes_per_step": P * 9 + 4},
            "steps": steps, "accepted": int(acc.sum()), "guard_rescored": int(ng)}


def run_reference(args, rank):
    """--impl reference: CPU arm, rank 0 only."""
    if rank != 0:
        return
    from desi_mcmc_b200 import synth
    field = synth.make_field(1005, H, W, N_SRC, STAR_FRAC)
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_baseline(field, per_step)
    ev, sec, last = 0, 0.0, None
    for _ in range(args.steps):
        last = cpu_baseline(field, per_step)
        ev += last["evals"]
        sec += last["seconds"]
    val = ev / sec
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "evals/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sec / max(1, args.steps),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": WORKLOAD, "sample_per_step": last["sample"]},
            "cpu_baseline": {"value": val, "unit": "evals/s", "cores": last["cores"], "kind": last["kind"],
                             "sample": last["sample"]},
            "e2e": {"value": val, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--fields-per-gpu", type=int, default=8)
    ap.add_argument("--streams", type=int, default=2,
                    help="side streams the field sets of a step are spread over (gather of one set "
                         "overlaps the render of the next)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step kernel by kernel "
                    "instead of replaying the captured CUDA graph")
    ap.add_argument("--fields-per-call", type=int, default=8,
                    help="fields batched into one cg_fields_loglik call (patches of a call stay in L2)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-proposals", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

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
    F = args.fields_per_gpu

    # ---- synthetic fields of this rank (config 5 seeds 1005+f), pixels drawn on device
    fields, images, srcs_host = [], [], []
    gen = torch.Generator(device=dev)
    for f in range(F):
        fid = rank * F + f
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

    # field sets: FPC fields per cg_fields_loglik call (one prepare/render/gather launch each)
    FPC = max(1, min(args.fields_per_call, F, _ffi.CG_MAX_FIELDS))
    groups = [list(range(g, min(g + FPC, F))) for g in range(0, F, FPC)]
    sets = [engine.FieldSet([images[f] for f in g], [N_SRC] * len(g), bound_cap=BOUND_CAP,
                            max_patch_px=MAX_PATCH_PX) for g in groups]
    for fs, g in zip(sets, groups):
        fs.set_sources([srcs_host[f] for f in g])
    stats = torch.zeros(64, dtype=torch.float64, device=dev)     # [sum ll, n_star, ...]
    n_star = float(sum(1 for fld in fields for s in fld["sources"] if s["a"] == 0))

    # one float64 all-reduce per step (global log-likelihood + counts), enqueued
    # behind the field sets and captured into the step's CUDA graph with them
    template = torch.zeros(64, dtype=torch.float64, device=dev)
    template[1] = n_star

    def exchange():
        stats.copy_(template)
        stats[0:1] = torch.stack([fs.ll.sum() for fs in sets]).sum(0, keepdim=True)
        dist.all_reduce(stats)

    sweep = engine.FieldSweep(sets, n_streams=args.streams, after=exchange if world > 1 else None)

    def step_device():
        sweep.run()

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
    # render_kernel launch bracketed by CUDA events on that stream (events cannot
    # be timed inside a graph replay, and overlapped kernels would share the SMs)
    for _ in range(args.warmup):
        for fs in sets:
            fs.loglik_device()
    barrier()
    engine.launch_count(reset=True)
    engine.profile_render(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        for fs in sets:
            fs.loglik_device()
    e1.record()
    barrier()
    serial_ms = e0.elapsed_time(e1)
    launches = engine.launch_count()        # kernels of args.steps steps (a graph replay launches the same ones)
    render_ms, render_n = engine.profile_render_read()
    engine.profile_render(False)
    ll_serial = np.concatenate([fs.ll.cpu().numpy() for fs in sets])

    # ---- device-resident leg (the headline `value`): the captured step
    if not args.no_graph:
        try:
            sweep.capture()
        except Exception as exc:                      # e.g. a collective that cannot be captured
            sys.stderr.write("CUDA graph capture failed (%s); launching kernel by kernel\n" % exc)
            sweep.graph = None
            args.no_graph = True
            torch.cuda.synchronize()
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    ll_ref = np.concatenate([fs.ll.cpu().numpy() for fs in sets])
    assert np.array_equal(ll_serial, ll_ref), "graph replay and serial launches differ"

    # ---- end-to-end leg: host source arrays -> pinned -> device, ll -> host
    def step_e2e():
        # the sampler's state lives in the pinned stage (pinned_sources); upload, the
        # field kernels, the all-reduce and the download are one graph replay
        sweep_e2e.run()
        torch.cuda.current_stream().synchronize()
        return float(sum(fs.h_ll.sum() for fs in sets))

    for fs, g in zip(sets, groups):       # host state written once into the pinned views
        for k, f in enumerate(g):
            fs.pinned_sources(k, srcs_host[f].shape[0])[:] = srcs_host[f]
    sweep_e2e = engine.FieldSweep(sets, n_streams=args.streams, after=exchange if world > 1 else None, io=True)
    if not args.no_graph:
        sweep.graph = None                 # one live graph per communicator
        try:
            sweep_e2e.capture()
        except Exception as exc:
            sys.stderr.write("CUDA graph capture of the end-to-end step failed (%s)\n" % exc)
            sweep_e2e.graph = None
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

    prop = None
    if rank == 0 and not args.no_proposals:
        prop = bench_proposals(engine, synth, max(3, min(args.steps, 10)))

    total_evals = evals_step * world          # every rank holds the same shapes, different seeds
    if world > 1:
        t = torch.tensor([float(evals_step)], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        total_evals = float(t.item())
    value = total_evals * args.steps / (dev_ms * 1e-3)
    e2e_value = total_evals * args.steps / e2e_s

    if rank == 0:
        xu_peak = engine.pipe_probe(0)
        fma_peak = engine.pipe_probe(1)
        mix_peak = engine.pipe_probe(6)         # the row-recurrence inner loop in isolation, 24 warps/SM
        fma2_peak = engine.pipe_probe(7)        # FP32 FMA operations/s issued as packed FFMA2
        # Per component and lane (8 pixels) the packed row-recurrence loop (ft_eval2r, SASS of
        # recur_probe_kernel) issues 1 FFMA + 6 FFMA2 + 1 FMUL + 4 FMUL2 + 3 FADD2 (28 FP32
        # operations per lane in 15 instructions) + 3 MUFU.EX2 + 2 LDS.128 + 2 MOV = 22 issue slots.
        # Three ceilings, all from live probes on this GPU: the FP32 pipe (operations/s, the better
        # of the FFMA and FFMA2 probes, / 3.5 per evaluation), the XU pipe (MUFU.EX2/s / 0.375) and
        # instruction issue (one warp instruction per SM sub-partition per clock = the FFMA probe's
        # instruction rate, / 2.75 slots per evaluation).  The lowest one is the roofline.
        FP32_PER_EVAL, SLOTS_PER_EVAL, MUFU_PER_EVAL = 28.0 / 8.0, 22.0 / 8.0, 3.0 / 8.0
        bounds = {"fp32 pipe": max(fma_peak, fma2_peak) / FP32_PER_EVAL,
                  "xu pipe": xu_peak / MUFU_PER_EVAL,
                  "instruction issue": fma_peak / SLOTS_PER_EVAL}
        bound_name = min(bounds, key=bounds.get)
        issue_peak = bounds[bound_name]
        achieved = evals_step * args.steps / (render_ms * 1e-3) if render_ms > 0 else None
        sm = engine.sm_count()
        nominal = sm * 16 * (clocks["sm_mhz"] or 0) * 1e6
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 rates, f64 log-likelihood accumulation",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "fields_per_gpu": F, "fields_total": F * world,
                       "fields_per_call": FPC, "streams": sweep.n_streams,
                       "launch": "kernel by kernel" if args.no_graph else "one CUDA graph replay per step",
                       "evals_per_step_per_gpu": evals_step, "issued_lane_evals_per_step_per_gpu": issued_step,
                       "pixels_per_step_per_gpu": F * 5 * H * W,
                       "l2": "inputs larger than L2 (%d MB of pixel data per GPU per step)" % (F * 5 * H * W * 4 >> 20),
                       "e2e_inputs": "packed source parameters copied pinned host -> device every step; image pixels resident",
                       "parallelism": "fields sharded over ranks; one float64 all-reduce per step" if world > 1 else "single GPU"},
            "e2e": {"value": e2e_value, "unit": "evals/s", "ms_per_step": 1e3 * e2e_s / args.steps,
                    "h2d_bytes_per_step": F * N_SRC * _ffi.SRC_DTYPE.itemsize, "d2h_bytes_per_step": F * 5 * 8},
            "gpu_launches": launches,
            "roofline": {"kernel": "field_tile_kernel", "bound": bound_name,
                         "achieved": achieved, "peak": issue_peak,
                         "unit": "evals/s (1 eval = 1 component at 1 pixel = 3.5 FP32 operations (1.875 instructions, "
                                 "packed FFMA2/FMUL2/FADD2) + 0.375 MUFU.EX2 + 0.25 LDS.128 + 0.25 MOV: 2.75 issue slots)",
                         "frac": (achieved / issue_peak) if achieved else None,
                         "bounds": bounds,
                         "peak_source": "measured live on this GPU (cg_pipe_probe 0, 1, 7): FP32 operations/s "
                                        "(FFMA %.4g, as FFMA2 %.4g) / 3.5, MUFU.EX2/s / 0.375, FFMA instructions/s / 2.75; "
                                        "the lowest is the peak.  Nominal FP32: %d SMs x 128/clk x %.0f MHz / 3.5 = %.3g" %
                                        (fma_peak, fma2_peak, sm, clocks["sm_mhz"] or 0,
                                         sm * 128 * (clocks["sm_mhz"] or 0) * 1e6 / FP32_PER_EVAL),
                         "issued": (issued_step * args.steps / (render_ms * 1e-3)) if render_ms > 0 else None,
                         "issued_frac": (issued_step * args.steps / (render_ms * 1e-3) / issue_peak) if render_ms > 0 else None,
                         "note": "achieved = algorithmic evals (footprint pixels x components, SURVEY 8d) / kernel time; "
                                 "issued = lane-evaluations actually executed after per-warp component culling. "
                                 "Up to v11 every eval cost one MUFU.EX2 and the bound was the XU pipe "
                                 "(`xu_frac` = achieved / that peak, kept for comparison); the row recurrence "
                                 "needs 0.375 MUFU.EX2 per eval; v13 packs the two rows of a lane into FP32x2 "
                                 "instructions (same operations, 15 instead of 28 FP32 issue slots per component and lane)",
                         "xu_peak": xu_peak, "xu_frac": (achieved / xu_peak) if achieved else None,
                         "mix_ceiling": {"value": mix_peak, "frac_of_peak": mix_peak / issue_peak,
                                         "note": "evals/s the kernel's inner loop (ft_eval2r: 2 LDS.128 + 1 FFMA + 6 FFMA2 + 1 FMUL + 4 FMUL2 "
                                                 "+ 3 FADD2 + 3 MUFU.EX2 + 2 MOV per component and lane) reaches in isolation "
                                                 "(cg_pipe_probe(6): no global memory, every lane busy); compare with `issued`"},
                         "fp32_fma_peak": fma_peak, "fp32_fma2_peak": fma2_peak, "kernel_ms_total": render_ms, "kernel_launches": render_n,
                         "kernel_share_of_step": render_ms / serial_ms if serial_ms else None,
                         "serial_ms_per_step": serial_ms / args.steps,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one timed 8-field launch from the
                         # ncu --set full capture summarised in profiles/r01_prof_tile_v14_*: 663.0 + 4.0 MB
                         # (algorithmic: 4 B of nelec per pixel = 488 MB); scaled to this launch's fields
                         "traffic": 666.97e6 * FPC / 8.0,
                         "traffic_algorithmic": FPC * 5 * H * W * 4,
                         "hbm_peak_gbs": peaks.get("hbm_gbs")},
            "clocks": clocks,
            "total_loglik": total_ll,
        }
        if prop is not None:
            line["proposals"] = prop
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = {k: v for k, v in cpu_baseline(fields[0], args.cpu_seconds).items()
                                    if k not in ("evals", "seconds")}
        out.write(json.dumps(line) + "\n")
        out.flush()
    if world > 1:
        # Tear-down: the step graph holds the captured NCCL all-reduce and must go
        # before the communicator.  The result is already printed; a watchdog ends
        # the process if NCCL's tear-down stalls (seen once with a live graph).
        wd = threading.Timer(20.0, lambda: os._exit(0))
        wd.daemon = True
        wd.start()
        sweep.graph = None
        sweep_e2e.graph = None
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
        wd.cancel()


if __name__ == "__main__":
    main()
