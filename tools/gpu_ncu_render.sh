ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 6 -c 1 -o gpurun_out/render_full -f python tools/gpu_c4_probe.py > gpurun_out/ncu_render.log 2>&1
ncu -i gpurun_out/render_full.ncu-rep --page details --csv > gpurun_out/prof_render_details.csv 2>/dev/null
ncu -i gpurun_out/render_full.ncu-rep --page source --csv > gpurun_out/render_source.csv 2>/dev/null
ncu -i gpurun_out/render_full.ncu-rep --page raw --csv > gpurun_out/render_raw.csv 2>/dev/null
rm -f gpurun_out/render_full.ncu-rep
tail -2 gpurun_out/ncu_render.log
