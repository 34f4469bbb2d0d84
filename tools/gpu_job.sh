# usage: bash tools/gpu_job.sh <tag> [pytest args...]   -- runs pytest -m gpu with the given selection, log under gpurun_out/
tag=$1; shift
timeout 1700 python -m pytest -m gpu -q -x "$@" 2>&1 | tail -40 | tee gpurun_out/pytest_${tag}.log
