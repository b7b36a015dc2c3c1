set -x
mkdir -p gpurun_out/r2w
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -8 | tee gpurun_out/r2w/no_alloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | tee gpurun_out/r2w/pytest_gpu.log
