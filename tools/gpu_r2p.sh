set -x
mkdir -p gpurun_out/r2v
timeout 900 python -m pytest tests/test_gpu_clt.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py --no-secondary > gpurun_out/r2v/bench_fused.json 2> gpurun_out/r2v/bench_fused.err; cut -c1-400 gpurun_out/r2v/bench_fused.json
timeout 300 python tools/perf_probe.py streams 2>&1 | tee gpurun_out/r2v/streams.jsonl | cut -c1-200
timeout 300 python bench.py --workload prbs_slicer > gpurun_out/r2v/bench_prbs_slicer.json 2>/dev/null; cut -c1-300 gpurun_out/r2v/bench_prbs_slicer.json
