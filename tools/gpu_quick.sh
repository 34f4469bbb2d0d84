# usage: bash tools/gpu_quick.sh <tag> [pytest -k expression]   -- field-path parity tests + the 8-field timing line
tag=$1
sel=${2:-"field or smoke or model or row_recurrence or linear"}
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x -k "$sel" 2>&1 | tail -5 | tee gpurun_out/quick_${tag}.log
python bench.py --main-only --fields-per-gpu 8 --steps 30 > gpurun_out/quick_${tag}.json 2>> gpurun_out/quick_${tag}.log
python - <<PY | tee -a gpurun_out/quick_${tag}.log
import json
d = json.load(open("gpurun_out/quick_${tag}.json"))
r = d["roofline"]
print("TAG ${tag}: value %.4g e2e %.4g ms/step %.4f frac %.3f kernel %.1f us share %.3f serial %.4f total_ll %.17g" % (
    d["value"], d["e2e"]["value"], d["ms_per_step"], r["frac"], 1e3 * r["kernel_ms_total"] / r["kernel_launches"], r["kernel_share_of_step"], r["serial_ms_per_step"], d["total_loglik"]))
PY
python tools/gpu_err_probe.py 2>&1 | tail -3 | tee -a gpurun_out/quick_${tag}.log
