set -x
timeout 1500 python -m pytest tests/test_gpu_clt.py tests/test_gpu_config_scale.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
