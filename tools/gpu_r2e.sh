set -x
mkdir -p gpurun_out/r2e
timeout 600 python -m pytest tests/test_gpu_clt.py tests/test_shaper.py tests/test_tools.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/perf_probe.py streams 2>&1 | grep lutopt | cut -c1-300
for w in clt_stream lutopt_stream prbs_stream; do
  timeout 300 python bench.py --workload $w > gpurun_out/r2e/bench_$w.json 2> gpurun_out/r2e/bench_$w.err; cut -c1-1000 gpurun_out/r2e/bench_$w.json; tail -2 gpurun_out/r2e/bench_$w.err
done
