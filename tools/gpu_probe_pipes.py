"""Prints the single-pipe and loop probes (cg_pipe_probe) of the loaded library."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from desi_mcmc_b200 import engine
engine.init(0)
names = {0: "MUFU.EX2 /s", 1: "FFMA lane-ops/s", 7: "FFMA2 ops/s (2 per lane)", 8: "FMUL2 ops/s", 9: "FADD2 ops/s",
         6: "recurrence loop evals/s @3 CTAs", 16: "recurrence loop evals/s @8 CTAs",
         20: "(2 FFMA2 + LOP3) groups/s", 21: "(2 FFMA + LOP3) groups/s", 22: "(4 FFMA + LOP3) groups/s"}
print("variant", os.environ.get("CG_LIB_VARIANT", "base"))
for w in (0, 1, 7, 8, 9, 6, 16, 20, 21, 22):
    v = engine.pipe_probe(w)
    print("  probe %2d %-34s %.4g" % (w, names[w], v) + ("   = %.2f cycles per group" % (clk / (v / 32 / sched)) if w >= 20 else ""))
    if w == 1:
        sched = 4 * engine.sm_count()
        clk = v / 32 / sched          # one FFMA per scheduler and cycle
