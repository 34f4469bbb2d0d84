python - <<'PY'
from desi_mcmc_b200 import engine
engine.init(0)
peak = engine.pipe_probe(0)
print("MUFU peak %.4g /s, FFMA %.4g /s" % (peak, engine.pipe_probe(1)))
for w in (2, 3, 4, 5, 12, 13, 14, 15):
    v = engine.pipe_probe(w)
    print("mix probe %2d: %.4g MUFU/s = %.3f of peak" % (w, v, v / peak))
PY
