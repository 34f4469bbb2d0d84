python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for cfg in "2 2" "2 1" "1 2" "4 2" "8 1" "1 4" "2 4"; do set -- $cfg
 echo "== fpc $1 streams $2"; python bench.py --fields-per-call $1 --streams $2 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | tee -a gpurun_out/bench_var.jsonl | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['serial_ms_per_step'], d['clocks'])"
done
echo "== nograph fpc2 s1"; python bench.py --fields-per-call 2 --streams 1 --no-graph --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | tee -a gpurun_out/bench_var.jsonl | cut -c1-200
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_v8.csv python bench.py --steps 2 --warmup 3 --fields-per-gpu 2 --fields-per-call 1 --no-graph --streams 1 --no-cpu-baseline --no-proposals > gpurun_out/ncu_v8.log 2>&1
grep -E "prepare_kernel|render_kernel|field_reduce|field_final" gpurun_out/launches_v8.csv | awk -F'","' '{print $5, $NF}' | tail -8
