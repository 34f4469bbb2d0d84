# usage: bash tools/gpu_variant_streams.sh "<variants>" "<streams list>" [extra bench args]   -- the 64-field step per variant and stream count
for v in $1; do
  if [ "$v" = "base" ]; then unset CG_LIB_VARIANT; else export CG_LIB_VARIANT=$v; fi
  for s in $2; do
    python bench.py --main-only --steps 10 --streams $s $3 > gpurun_out/vs.json 2> gpurun_out/vs.err || tail -5 gpurun_out/vs.err
    python - <<PY | tee -a gpurun_out/variant_streams.log
import json
d = json.load(open("gpurun_out/vs.json"))
r = d["roofline"]
print("VARIANT ${v} streams ${s} $3: value %.4g e2e %.4g ms/step %.4f kernel %.1f us frac %.3f total_ll %.17g" % (
    d["value"], d["e2e"]["value"], d["ms_per_step"], 1e3 * r["kernel_ms_total"] / r["kernel_launches"], r["frac"], d["total_loglik"]))
PY
  done
done
