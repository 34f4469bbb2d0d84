timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -8
python bench.py --steps 30 > gpurun_out/b_recur1.json 2>gpurun_out/b_recur1.err
CG_RECUR=0 python bench.py --steps 30 > gpurun_out/b_recur0.json 2>gpurun_out/b_recur0.err
python - <<'PY'
import json
for f in ["gpurun_out/b_recur1.json","gpurun_out/b_recur0.json"]:
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel_ms_total"]/d["roofline"]["kernel_launches"], d["total_loglik"])
PY
