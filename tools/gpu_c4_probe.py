"""Per-phase CUDA-event timing of the config-4 sweep (5000 sources, 2048x1489x5, 16 proposals/source)."""
import sys, time, json, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from desi_mcmc_b200 import engine, synth

def build(n_src=5000, H=1489, W=2048, P=16, seed=1004, background_pass=True):
    engine.init(0)
    dev = engine.device()
    fld = synth.make_crowded_field(seed, H=H, W=W, n_src=n_src)
    B, lam0 = synth.make_eigen_basis(seed)
    imgs = [engine.Image(nelec=torch.zeros((H, W), dtype=torch.float32, device=dev), **synth.image_kwargs(m)) for m in fld["images"]]
    srcs = [engine.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"], sigma=s["sigma"],
                             rho=s["rho"], fluxes=s["fluxes"], flux_row=s["eigen"]["row"] if s["eigen"] else None) for s in fld["sources"]]
    packed = engine.pack_sources(srcs)
    es = [s["eigen"] for s in fld["sources"] if s["eigen"]]
    eig = engine.EigenSpectra([e["z"] for e in es], [e["w"] for e in es], [e["m"] for e in es], B, lam0)
    eig.project()
    _, models = engine.field_loglik(imgs, packed, bound_cap=25., max_patch_px=53 * 53, return_model=True)
    gen = torch.Generator(device=dev); gen.manual_seed(seed)
    for im, lam in zip(imgs, models):
        im.set_nelec(torch.poisson(lam, generator=gen))
    del models
    t0 = time.time()
    sw = engine.GibbsSweep(imgs, packed, eigen=eig, proposals=P, seed=seed, background_pass=background_pass)
    return sw, time.time() - t0

if __name__ == "__main__":
    sw, setup_s = build()
    print("setup %.1f s, workspace %.2f GB, neighbours %d" % (setup_s, sw.workspace_bytes / 1e9, sw.h_nbr[1].size))
    S, P = sw.n_src, sw.P
    rng = np.random.RandomState(0)
    lu = np.log(rng.uniform(size=S * P))
    phases = [("eigen", lambda: sw.eigen.project()), ("allocate", lambda: sw.allocate(1)), ("epsilon", lambda: sw.resample_epsilon(2)),
              ("fluxes", lambda: sw.resample_fluxes(3)), ("propose", lambda: sw.propose(4)), ("score_cond", lambda: sw.score(0, lu)),
              ("select", lambda: sw.select(0)), ("backgrounds", lambda: sw.backgrounds()), ("propose2", lambda: sw.propose(5)),
              ("score_bg", lambda: sw.score(1, lu)), ("select2", lambda: sw.select(1)), ("loglik", lambda: sw.loglik())]
    for rep in range(3):
        out = {}
        for name, fn in phases:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            out[name] = round(e0.elapsed_time(e1), 3)
            if name in ("allocate", "score_cond", "score_bg", "backgrounds", "loglik"):
                out[name + "_evals"] = engine.last_eval_count(); out[name + "_issued"] = engine.last_issued_count()
        print(json.dumps(out))
    torch.cuda.synchronize(); t0 = time.time()
    for _ in range(5): sw.sweep()
    st = sw.stats(); dt = (time.time() - t0) / 5
    print("sweep %.2f ms" % (dt * 1e3), {k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in st.items()})
