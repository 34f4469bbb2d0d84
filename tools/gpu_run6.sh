# A/B of tile-kernel occupancy variants: swap the library file, run the bench
cp desi_mcmc_b200/libceleste_b200.so /tmp/base.so
for v in base 128_7 128_6 192_4; do
  if [ $v = base ]; then cp /tmp/base.so desi_mcmc_b200/libceleste_b200.so; else cp gpurun_variants/lib_$v.so desi_mcmc_b200/libceleste_b200.so; fi
  for fpc in 8 1; do
  echo "== $v fpc $fpc"; python bench.py --fields-per-call $fpc --streams 1 --steps 30 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['value'], d['ms_per_step'], r['frac'], r['kernel_ms_total']/r['kernel_launches']*1e3, 'us/launch')"
  done
done
cp /tmp/base.so desi_mcmc_b200/libceleste_b200.so
