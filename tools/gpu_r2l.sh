set -x
mkdir -p gpurun_out/r2l
timeout 900 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_rnghunt.py tests/test_gpu_cli.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python - <<'PY'
import sys; sys.path.insert(0, '.')
from basebandboard_b200 import lib
c = lib.Context(0)
for cnt in (1 << 21, 1 << 24):
    c.rnghunt_sweep(192, 1, 0, 1 << 16, k=16)
    for rep in range(2):
        r = c.rnghunt_sweep(192, 1, 0, cnt, k=16)
        print("sweep 2^%d: %.2f ms %.2f M/s launches %d best %s" % (cnt.bit_length() - 1, c.last_device_ms, cnt / c.last_device_ms / 1e3, c.last_launches, r["best"][:3]))
c.close()
PY
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2l/bench_default_n2.json 2> gpurun_out/r2l/bench_default_n2.err; python -c "
import json; d=json.load(open('gpurun_out/r2l/bench_default_n2.json')); h=d['secondary']['rnghunt']; print(d['value'], d['result_check']['reduced_hist_total']==d['result_check']['reduced_expected'], h['value'], h['ms_per_step'], h['result_check']['best'][:4])"
for w in prbs_detect prbs_shaped_slicer; do
  timeout 400 $TR bench.py --gpus 2 --steps 3 --warmup 3 --workload $w > gpurun_out/r2l/bench_${w}_n2.json 2> gpurun_out/r2l/bench_${w}_n2.err; cut -c1-120 gpurun_out/r2l/bench_${w}_n2.json; tail -1 gpurun_out/r2l/bench_${w}_n2.err | cut -c1-200
done
