cp desi_mcmc_b200/libceleste_b200.so /tmp/base.so
for v in base T24 T21; do
  if [ $v = base ]; then cp /tmp/base.so desi_mcmc_b200/libceleste_b200.so; else cp gpurun_variants/lib_$v.so desi_mcmc_b200/libceleste_b200.so; fi
  echo "== $v"; python bench.py --steps 30 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['value'], d['ms_per_step'], r['frac'], r['issued_frac'], d['config']['issued_lane_evals_per_step_per_gpu'], '%.6f' % d['total_loglik'])"
done
cp /tmp/base.so desi_mcmc_b200/libceleste_b200.so
