timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 | tee gpurun_out/r01_pytest_gpu_v14.log
# round-1 v14 evidence: bench line, reference arm, launch list, full capture of an 8-field timed launch
python bench.py > gpurun_out/r01_bench_v14.json 2> gpurun_out/bench_v14.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_bench_v14_reference.json 2>> gpurun_out/bench_v14.err
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-proposals"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_v14.csv $CMD > gpurun_out/ncu_v14.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:field_tile_kernel -s 10 -c 1 -o gpurun_out/tile_full -f $CMD > gpurun_out/ncu_tile_full.log 2>&1
ncu -i gpurun_out/tile_full.ncu-rep --page details --csv > gpurun_out/r01_prof_tile_v14_details.csv 2>/dev/null
ncu -i gpurun_out/tile_full.ncu-rep --page source --csv > gpurun_out/tile_source.csv 2>/dev/null
ncu -i gpurun_out/tile_full.ncu-rep --page raw --csv > gpurun_out/tile_raw.csv 2>/dev/null
tail -c 300 gpurun_out/r01_bench_v14.json
