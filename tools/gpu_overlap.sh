for cfg in "--fields-per-gpu 8 --fields-per-call 8 --streams 1" "--fields-per-gpu 8 --fields-per-call 4 --streams 2" "--fields-per-gpu 8 --fields-per-call 2 --streams 2" "--fields-per-gpu 8 --fields-per-call 2 --streams 4" "--total-fields 64 --streams 1" "--total-fields 64 --streams 2" "--total-fields 64 --streams 3"; do
python bench.py --main-only --steps 20 $cfg > gpurun_out/ov.json 2>/dev/null
python - "$cfg" <<PY
import json,sys
d=json.load(open("gpurun_out/ov.json")); r=d["roofline"]
print("%-60s value %.4g e2e %.4g ms/step %.4f" % (sys.argv[1], d["value"], d["e2e"]["value"], d["ms_per_step"]))
PY
done 2>&1 | tee gpurun_out/overlap1.log
