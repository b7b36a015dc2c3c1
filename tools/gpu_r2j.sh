set -x
mkdir -p gpurun_out/r2j
timeout 1500 python -m pytest tests -m gpu -q --durations=6 > gpurun_out/r2j/pytest_gpu.log 2>&1; echo rc=$?; tail -12 gpurun_out/r2j/pytest_gpu.log | cut -c1-200
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j/smoke.log 2>&1; echo smoke rc=$?; tail -7 gpurun_out/r2j/smoke.log
timeout 600 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_config_scale.py > gpurun_out/r2j/pytest_gpu_2.log 2>&1; echo rc=$?; tail -2 gpurun_out/r2j/pytest_gpu_2.log | cut -c1-200
