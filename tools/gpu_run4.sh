for v in 0 1 2 3; do for fpc in 1 8; do
 echo "== variant $v fpc $fpc"; CG_TILE_VARIANT=$v python bench.py --fields-per-call $fpc --streams 1 --no-graph --steps 20 --no-cpu-baseline --no-proposals 2>>gpurun_out/bench_var.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['ms_per_step'], r['frac'], r['kernel_ms_total']/r['kernel_launches']*1e3, 'us/launch')"
done; done
