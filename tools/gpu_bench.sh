# usage: bash tools/gpu_bench.sh <tag> [ncu]   -- bench line (+ reference arm), optionally launch list + full capture of field_tile_kernel
tag=$1
python bench.py > gpurun_out/bench_${tag}.json 2> gpurun_out/bench_${tag}.err || { echo "bench failed"; tail -20 gpurun_out/bench_${tag}.err; }
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_${tag}.json"))
    r = d["roofline"]
    print("value %.4g e2e %.4g ms/step %.4f frac %.3f issued_frac %.3f kernel_ms/launch %.1f us share %.3f" % (
        d["value"], d["e2e"]["value"], d["ms_per_step"], r["frac"], r["issued_frac"], 1e3 * r["kernel_ms_total"] / r["kernel_launches"], r["kernel_share_of_step"]))
    for k in ("config1", "proposals", "config4"):
        if k in d:
            x = d[k]
            print(k, {kk: x[kk] for kk in ("latency_us", "value", "sweep_ms", "proposals_per_s") if kk in x}, "frac", x.get("roofline", {}).get("frac"), "cpu", x.get("cpu_baseline", {}).get("value"))
    print("cpu", d.get("cpu_baseline", {}).get("value"), d.get("cpu_baseline", {}).get("single_thread", {}).get("value"))
except Exception as e:
    print("no bench line:", e)
PY
if [ "$2" = "ncu" ]; then
  python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_${tag}_reference.json 2>> gpurun_out/bench_${tag}.err
  CMD="python bench.py --steps 2 --warmup 3 --no-graph --main-only --fields-per-gpu 8"
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${tag}.csv $CMD > gpurun_out/ncu_${tag}.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:field_tile_kernel -s 10 -c 1 -o gpurun_out/tile_full_${tag} -f $CMD > gpurun_out/ncu_tile_full_${tag}.log 2>&1
  ncu -i gpurun_out/tile_full_${tag}.ncu-rep --page details --csv > gpurun_out/prof_tile_${tag}_details.csv 2>/dev/null
  ncu -i gpurun_out/tile_full_${tag}.ncu-rep --page source --csv > gpurun_out/tile_source_${tag}.csv 2>/dev/null
  ncu -i gpurun_out/tile_full_${tag}.ncu-rep --page raw --csv > gpurun_out/tile_raw_${tag}.csv 2>/dev/null
fi
if [ "$2" = "ncu" ]; then
  ncu -i gpurun_out/tile_full_${tag}.ncu-rep --page source --csv --print-source cuda > gpurun_out/tile_cuda_${tag}.csv 2>/dev/null
fi
