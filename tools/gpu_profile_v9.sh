# round-1 v9 evidence: bench line, reference arm, launch list, full capture of the dominant kernel
set -x
python bench.py > gpurun_out/r01_bench_v9.json 2> gpurun_out/bench_v9.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_bench_v9_reference.json 2>> gpurun_out/bench_v9.err
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-proposals"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_v9.csv $CMD > gpurun_out/ncu_v9.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:field_tile_kernel -s 4 -c 1 -o gpurun_out/tile_full -f $CMD > gpurun_out/ncu_tile_full.log 2>&1
ncu -i gpurun_out/tile_full.ncu-rep --page details --csv > gpurun_out/r01_prof_tile_v9_details.csv 2>/dev/null
ncu -i gpurun_out/tile_full.ncu-rep --page source --csv > gpurun_out/tile_source.csv 2>/dev/null
ncu -i gpurun_out/tile_full.ncu-rep --page raw --csv > gpurun_out/tile_raw.csv 2>/dev/null
tail -c 600 gpurun_out/r01_bench_v9.json
