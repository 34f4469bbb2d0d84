for c in 0 1; do
 echo "== cull $c"; CG_CULL=$c python bench.py --fields-per-call 8 --streams 1 --steps 20 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['value'], d['ms_per_step'], r['frac'], r['issued_frac'], r['kernel_ms_total']/r['kernel_launches']*1e3, 'us/launch', d['total_loglik'])"
done
