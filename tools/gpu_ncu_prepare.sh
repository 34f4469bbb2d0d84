# one full capture of prepare_kernel<float> on the 8-field step
CMD="python bench.py --steps 2 --warmup 3 --no-graph --main-only --fields-per-gpu 8"
ncu --set full --clock-control none --import-source on -k regex:prepare_kernel -s 6 -c 1 -o gpurun_out/prep_full -f $CMD > gpurun_out/ncu_prep.log 2>&1
ncu -i gpurun_out/prep_full.ncu-rep --page details --csv > gpurun_out/prof_prep_details.csv 2>/dev/null
ncu -i gpurun_out/prep_full.ncu-rep --page source --csv > gpurun_out/prep_source.csv 2>/dev/null
tail -2 gpurun_out/ncu_prep.log
