ncu --set full --clock-control none --import-source on -k regex:alloc_tile_kernel -s 2 -c 1 -o gpurun_out/alloc_full -f python tools/gpu_c4_probe.py > gpurun_out/ncu_alloc.log 2>&1
ncu -i gpurun_out/alloc_full.ncu-rep --page details --csv > gpurun_out/prof_alloc_details.csv 2>/dev/null
ncu -i gpurun_out/alloc_full.ncu-rep --page source --csv > gpurun_out/alloc_source.csv 2>/dev/null
ncu -i gpurun_out/alloc_full.ncu-rep --page raw --csv > gpurun_out/alloc_raw.csv 2>/dev/null
rm -f gpurun_out/alloc_full.ncu-rep
tail -3 gpurun_out/ncu_alloc.log
