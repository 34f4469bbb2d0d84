CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-proposals"
ncu --set full --clock-control none --import-source on -k regex:"prepare_kernel|bin_order_kernel" -s 18 -c 2 -o gpurun_out/prep_full -f $CMD > gpurun_out/ncu_prep_full.log 2>&1
ncu -i gpurun_out/prep_full.ncu-rep --page details --csv > gpurun_out/prep_details.csv 2>/dev/null
ncu -i gpurun_out/prep_full.ncu-rep --page source --csv > gpurun_out/prep_source.csv 2>/dev/null
ncu -i gpurun_out/prep_full.ncu-rep --page raw --csv > gpurun_out/prep_raw.csv 2>/dev/null
tail -3 gpurun_out/ncu_prep_full.log
