"""Times the on-device location slice sampler on the config-4 field (5000 sources, 5 images)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gpu_c4_probe import build
import torch

if __name__ == "__main__":
    sw, _ = build(background_pass=False)
    sw.sweep()
    for rep in range(3):
        torch.cuda.synchronize()
        sw.slice_locations(seed=100 + rep)
        st = sw.slice_stats
        print(json.dumps({"kernel_ms": round(st["kernel_ms"], 3), "evals": int(st["evals"].sum()),
                          "refined": int(st["refined"].sum()), "max_used": int(st["used"].max())}))
    for rep in range(2):
        keep = sw.d_cur.clone()
        _, acc = sw.hmc_galaxies(seed=12, step=1e-9, n_steps=2)
        sw.d_cur.copy_(keep)
        print(json.dumps({"variant": os.environ.get("CG_LIB_VARIANT", "base"), "hmc_kernel_ms": round(sw.hmc_stats["kernel_ms"], 2),
                          "likelihoods": int(sw.hmc_stats["likelihoods"]), "accepted": int(acc.sum())}))
