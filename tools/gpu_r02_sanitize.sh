# round 2: baseline GPU tests + compute-sanitizer memcheck / racecheck on the field, scoring and allocation paths
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3 | tee gpurun_out/r02_pytest_gpu_base.log
SEL='field_loglik or score_proposals or sample_source'
timeout 1500 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_gpu_parity.py tests/test_gpu_compat.py -q -x -k "$SEL" > gpurun_out/r02_sanitizer_memcheck.log 2>&1
echo "memcheck rc=$?" | tee -a gpurun_out/r02_sanitizer_memcheck.log
timeout 1500 compute-sanitizer --tool racecheck --print-limit 20 python -m pytest tests/test_gpu_parity.py tests/test_gpu_compat.py -q -x -k "$SEL" > gpurun_out/r02_sanitizer_racecheck.log 2>&1
echo "racecheck rc=$?" | tee -a gpurun_out/r02_sanitizer_racecheck.log
tail -5 gpurun_out/r02_sanitizer_memcheck.log gpurun_out/r02_sanitizer_racecheck.log
