# usage: bash tools/gpu_variant_issued.sh <variant> ...   -- 8-field kernel time and issued lane-evals per variant (timing experiments)
for v in "$@"; do
  if [ "$v" = "base" ]; then unset CG_LIB_VARIANT; else export CG_LIB_VARIANT=$v; fi
  python bench.py --main-only --fields-per-gpu 8 --steps 30 > gpurun_out/var_${v}.json 2> gpurun_out/var_${v}.err || tail -5 gpurun_out/var_${v}.err
  python - <<PY | tee -a gpurun_out/variants.log
import json
d = json.load(open("gpurun_out/var_${v}.json"))
r = d["roofline"]
print("VARIANT ${v}: ms/step %.4f kernel %.1f us issued lane-evals per launch %.4g (algorithmic %.4g)" % (
    d["ms_per_step"], 1e3 * r["kernel_ms_total"] / r["kernel_launches"], d["config"]["issued_lane_evals_per_step_per_gpu"], d["config"]["evals_per_step_per_gpu"]))
PY
done
