# One GPU validation pass: tests, smoke, the secondary bench workloads.
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -8 gpurun_out/smoke.log
for w in prbs_stream prbs_detect rnghunt prbs_slicer; do
  python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; cat gpurun_out/bench_$w.json; tail -2 gpurun_out/bench_$w.err
done
python bench.py --workload clt_stream --log2-samples 33 --steps 3 --warmup 3 > gpurun_out/bench_clt_stream_i8.json 2> gpurun_out/bench_clt_stream.err; cat gpurun_out/bench_clt_stream_i8.json
python bench.py --workload clt_stream --log2-samples 32 --noise-var-stream 8 --steps 3 --warmup 3 > gpurun_out/bench_clt_stream_i16.json 2>> gpurun_out/bench_clt_stream.err; cat gpurun_out/bench_clt_stream_i16.json
tail -3 gpurun_out/bench_clt_stream.err
python bench.py --workload tx_stream --steps 3 --warmup 3 > gpurun_out/bench_tx_stream.json 2>> gpurun_out/bench_clt_stream.err; cat gpurun_out/bench_tx_stream.json
python tools/perf_probe.py streams rnghunt > gpurun_out/probe_streams_hunt.jsonl 2>/dev/null; cut -c1-200 gpurun_out/probe_streams_hunt.jsonl
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; cut -c1-200 gpurun_out/bench_default.json
