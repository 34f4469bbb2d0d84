python tools/gpu_slice_probe.py > gpurun_out/slice_probe.log 2>&1; tail -4 gpurun_out/slice_probe.log
ncu --set full --clock-control none --import-source on -k regex:slice_location -c 1 -o gpurun_out/slice_full -f python tools/gpu_slice_probe.py > gpurun_out/ncu_slice.log 2>&1
ncu -i gpurun_out/slice_full.ncu-rep --page details --csv > gpurun_out/prof_slice_details.csv 2>/dev/null
ncu -i gpurun_out/slice_full.ncu-rep --page source --csv > gpurun_out/slice_source.csv 2>/dev/null
rm -f gpurun_out/slice_full.ncu-rep
tail -2 gpurun_out/ncu_slice.log
