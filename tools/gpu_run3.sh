python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for cfg in "8 1" "4 2" "2 2" "1 1"; do set -- $cfg
 echo "== fpc $1 streams $2 $3"; python bench.py --fields-per-call $1 --streams $2 $3 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | tee -a gpurun_out/bench_var.jsonl | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['issued_frac'], d['roofline']['serial_ms_per_step'], d['total_loglik'])"
done
tail -5 gpurun_out/bench_var.err
