timeout 300 python tools/_narrow_probe.py 2>&1 | tail -8
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 30 --no-cpu-baseline > gpurun_out/b9.json 2>gpurun_out/b9.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/b9.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("%.4g"%d["value"], "ms/step %.4f"%d["ms_per_step"], "kernel us %.1f"%(1e3*r["kernel_ms_total"]/r["kernel_launches"]), "frac %.3f"%r["frac"], "prop %.4g"%d["proposals"]["value"], d["total_loglik"])
PY
