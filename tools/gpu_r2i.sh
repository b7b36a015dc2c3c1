set -x
mkdir -p gpurun_out/r2i
timeout 600 python -m pytest tests/test_gpu_clt.py tests/test_gpu_rnghunt.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/perf_probe.py streams 2>&1 | grep lutopt | cut -c1-300
timeout 300 python - <<'PY'
import sys; sys.path.insert(0, '.')
from basebandboard_b200 import lib
c = lib.Context(0)
for cnt in (1 << 21, 1 << 22, 1 << 24):
    c.rnghunt_sweep(192, 1, 0, 1 << 16, k=16)
    for rep in range(2):
        r = c.rnghunt_sweep(192, 1, 0, cnt, k=16)
        print("sweep 2^%d: %.2f ms %.2f M/s launches %d" % (cnt.bit_length() - 1, c.last_device_ms, cnt / c.last_device_ms / 1e3, c.last_launches))
c.close()
PY
