# usage: bash tools/gpu_legs.sh <variant> ...   -- the secondary bench legs (C3 proposals, C4 sweep, slice sampler) per library variant
for v in "$@"; do
  if [ "$v" = "base" ]; then unset CG_LIB_VARIANT; else export CG_LIB_VARIANT=$v; fi
  python bench.py --no-stamp --no-cpu-baseline --steps 5 > gpurun_out/legs_${v}.json 2> gpurun_out/legs_${v}.err || tail -5 gpurun_out/legs_${v}.err
  python - <<PY | tee -a gpurun_out/legs.log
import json
d = json.load(open("gpurun_out/legs_${v}.json"))
p, c = d["proposals"], d["config4"]
print("LEGS ${v}: C3 %.4g proposals/s (e2e %.4g) frac %.3f | C4 sweep %.2f ms frac %.3f score %.2f + %.2f ms | slice %.1f ms" % (
    p["value"], p["e2e"]["value"], p["roofline"]["frac"], c["sweep_ms"], c["roofline"]["frac"],
    c["phase_ms"]["score_conditional"], c["phase_ms"]["score_mh_background"], c["slice_sampler_on_device"]["kernel_ms"]))
PY
done
