# final evidence refresh on one GPU
set -x
mkdir -p gpurun_out/fin3
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/fin3/pytest_gpu.log 2>&1; echo rc=$?; tail -3 gpurun_out/fin3/pytest_gpu.log | cut -c1-200
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/fin3/smoke.log 2>&1; echo smoke rc=$?
timeout 900 python bench.py > gpurun_out/fin3/bench_default.json 2> gpurun_out/fin3/bench_default.err; echo default rc=$?
for w in lutopt_stream rnghunt clt_stream tx_stream prbs_stream prbs_slicer prbs_shaped_slicer prbs_detect; do
  timeout 400 python bench.py --workload $w > gpurun_out/fin3/bench_$w.json 2> gpurun_out/fin3/bench_$w.err; echo $w rc=$?
done
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/fin3/bench_reference.json 2>/dev/null; echo ref rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/fin3/bench_default_2p34_launches.csv python bench.py --log2-samples 34 --steps 2 --warmup 1 > gpurun_out/fin3/ncu_default.log 2>&1; echo ncu rc=$?
for f in gpurun_out/fin3/bench_*.json; do echo $f; cut -c1-150 $f; done
python -c "
import json; d=json.load(open('gpurun_out/fin3/bench_default.json')); h=d['secondary']['rnghunt']; print('CLT', d['value'], d['e2e']['value'], d['roofline']['frac']); print('hunt', h['value'], h['e2e']['value'], h['roofline']['frac'], h['result_check']['best'][:4])"
