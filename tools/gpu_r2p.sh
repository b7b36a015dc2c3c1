set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu -k "allocates_no_device_memory" 2>&1 | tail -2
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python -m pytest tests/test_gpu_clt.py -x -q -m gpu -k "four_group" 2>&1 | tail -2
