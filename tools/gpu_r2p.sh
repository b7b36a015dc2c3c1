set -x
mkdir -p gpurun_out/fin4
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/fin4/pytest_gpu.log 2>&1; echo rc=$?; tail -2 gpurun_out/fin4/pytest_gpu.log
