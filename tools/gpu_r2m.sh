set -x
timeout 600 python -m pytest tests/test_gpu_clt.py -m gpu -x -q -k "word or lutopt" 2>&1 | tail -3
timeout 300 python tools/perf_probe.py streams 2>&1 | grep lutopt | cut -c1-300
