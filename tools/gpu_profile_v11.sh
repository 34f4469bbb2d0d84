python bench.py > gpurun_out/r01_bench_v11.json 2> gpurun_out/bench_v11.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_bench_v11_reference.json 2>> gpurun_out/bench_v11.err
tail -c 1200 gpurun_out/r01_bench_v11.json; tail -3 gpurun_out/bench_v11.err
