set -x
mkdir -p gpurun_out/r2x
timeout 900 python -m pytest tests/test_gpu_clt.py tests/test_gpu_prbs.py tests/test_shaper.py -x -q -m gpu 2>&1 | tail -3
for w in prbs_slicer prbs_shaped_slicer; do timeout 300 python bench.py --workload $w > gpurun_out/r2x/bench_$w.json 2>/dev/null; cut -c1-130 gpurun_out/r2x/bench_$w.json; python -c "
import json; d=json.load(open('gpurun_out/r2x/bench_$w.json')); print(d['value'], d.get('roofline',{}).get('frac'), d['e2e']['value'])"; done
