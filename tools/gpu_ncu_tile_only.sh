# usage: bash tools/gpu_ncu_tile_only.sh <tag>   -- one full capture of field_tile_kernel on the 8-field step
tag=$1
CMD="python bench.py --steps 2 --warmup 3 --no-graph --main-only --fields-per-gpu 8"
ncu --set full --clock-control none --import-source on -k regex:field_tile_kernel -s 10 -c 1 -o gpurun_out/tile_full_${tag} -f $CMD > gpurun_out/ncu_tile_full_${tag}.log 2>&1
ncu -i gpurun_out/tile_full_${tag}.ncu-rep --page details --csv > gpurun_out/prof_tile_${tag}_details.csv 2>/dev/null
ncu -i gpurun_out/tile_full_${tag}.ncu-rep --page source --csv > gpurun_out/tile_source_${tag}.csv 2>/dev/null
ncu -i gpurun_out/tile_full_${tag}.ncu-rep --page raw --csv > gpurun_out/tile_raw_${tag}.csv 2>/dev/null
rm -f gpurun_out/tile_full_${tag}.ncu-rep
tail -2 gpurun_out/ncu_tile_full_${tag}.log
