# variants of the row-recurrence loop (libs prebuilt under build/variants: lib_<name>.so)
cp desi_mcmc_b200/libceleste_b200.so /tmp/lib_default.so
for v in ${VARIANTS:-pq0 pq1}; do
  cp build/variants/lib_$v.so desi_mcmc_b200/libceleste_b200.so
  echo "=== $v"
  timeout 600 python -m pytest tests -m gpu -q 2>&1 | grep -E "^E  +Assertion|passed|failed" | cut -c1-200
  timeout 300 python tools/gpu_err_probe.py 2>&1 | tail -2
  python bench.py --steps 30 --no-cpu-baseline > gpurun_out/var13_$v.json 2>gpurun_out/var13_$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/var13_$v.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("%.4g"%d["value"], "ms/step %.4f"%d["ms_per_step"], "kernel us %.1f"%(1e3*r["kernel_ms_total"]/r["kernel_launches"]), "mix %.4g"%r["mix_ceiling"]["value"], "prop %.4g"%d["proposals"]["value"], repr(d["total_loglik"]))
PY
done
cp /tmp/lib_default.so desi_mcmc_b200/libceleste_b200.so
