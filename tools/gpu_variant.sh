# usage: bash tools/gpu_variant.sh <variant> ...   -- 8-field timing line (+ probes) for each libceleste_b200_<variant>.so ("base" = the product library)
for v in "$@"; do
  if [ "$v" = "base" ]; then unset CG_LIB_VARIANT; else export CG_LIB_VARIANT=$v; fi
  python bench.py --main-only --fields-per-gpu 8 --steps 30 > gpurun_out/var_${v}.json 2> gpurun_out/var_${v}.err || tail -5 gpurun_out/var_${v}.err
  python - <<PY | tee -a gpurun_out/variants.log
import json
d = json.load(open("gpurun_out/var_${v}.json"))
r = d["roofline"]
print("VARIANT ${v}: ms/step %.4f kernel %.1f us frac %.3f mix %.4g total_ll %.17g | %s" % (
    d["ms_per_step"], 1e3 * r["kernel_ms_total"] / r["kernel_launches"], r["frac"], r["mix_ceiling"]["value"], d["total_loglik"], r["peak_source"][-90:]))
PY
done
