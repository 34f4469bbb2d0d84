python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
rm -f gpurun_out/bench_var.jsonl
for cfg in "2 1 --unfused" "2 1" "2 2" "8 1" "1 1"; do set -- $cfg
 echo "== fpc $1 streams $2 $3"; python bench.py --fields-per-call $1 --streams $2 $3 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | tee -a gpurun_out/bench_var.jsonl | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['issued_frac'], d['roofline']['serial_ms_per_step'], d['total_loglik'])"
done
tail -5 gpurun_out/bench_var.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_v9.csv python bench.py --steps 2 --warmup 3 --fields-per-gpu 2 --fields-per-call 1 --no-graph --streams 1 --no-cpu-baseline --no-proposals > gpurun_out/ncu_v9.log 2>&1
grep -E "prepare_kernel|bin_order|field_tile" gpurun_out/launches_v9.csv | awk -F'","' '{print $5, $NF}' | tail -8
