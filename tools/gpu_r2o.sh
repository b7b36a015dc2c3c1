set -x
mkdir -p gpurun_out/r2o
timeout 600 python -m pytest tests/test_gpu_prbs.py tests/test_gpu_clt.py -m gpu -x -q 2>&1 | tail -3
for w in prbs_stream clt_stream lutopt_stream tx_stream; do
  timeout 400 python bench.py --workload $w > gpurun_out/r2o/bench_$w.json 2> gpurun_out/r2o/bench_$w.err; tail -1 gpurun_out/r2o/bench_$w.err; python -c "
import json; d=json.load(open('gpurun_out/r2o/bench_$w.json')); print('$w', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'pinned', round(d['e2e_pinned']['value'],1))"
done
