for fpc in 8; do
python bench.py --fields-per-call $fpc --steps 30 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['value'], d['ms_per_step'], d['e2e']['value'], r['frac'], r['issued_frac'], r['kernel_ms_total']/r['kernel_launches']*1e3, '%.6f' % d['total_loglik'])"
done
