# one --set full capture of field_tile_kernel (third launch = after warm-up), details + source pages
ncu --set full --clock-control none --import-source on -k regex:field_tile_kernel -s 3 -c 1 -o gpurun_out/tile_full -f python bench.py --steps 2 --warmup 3 --fields-per-gpu 2 --fields-per-call 1 --no-graph --streams 1 --no-cpu-baseline --no-proposals > gpurun_out/ncu_tile_full.log 2>&1
ncu -i gpurun_out/tile_full.ncu-rep --page details --csv > gpurun_out/tile_details.csv 2>/dev/null
ncu -i gpurun_out/tile_full.ncu-rep --page source --csv > gpurun_out/tile_source.csv 2>/dev/null
ncu -i gpurun_out/tile_full.ncu-rep --page raw --csv > gpurun_out/tile_raw.csv 2>/dev/null
ls -la gpurun_out/ | tail -8
