python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
for cfg in "8 1" "1 1"; do set -- $cfg
 echo "== fpc $1 streams $2"; python bench.py --fields-per-call $1 --streams $2 --steps 30 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['value'], d['ms_per_step'], r['frac'], r['issued_frac'], r['kernel_ms_total']/r['kernel_launches']*1e3, 'us/launch', d['total_loglik'])"
done
tail -3 gpurun_out/bench_var.err
