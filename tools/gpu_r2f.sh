# round 2 evidence pass on one GPU: tests, smoke, default bench (both halves of the metric), every secondary workload
set -x
mkdir -p gpurun_out/r2f
timeout 1200 python -m pytest tests -m gpu -q --durations=6 > gpurun_out/r2f/pytest_gpu.log 2>&1; tail -12 gpurun_out/r2f/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f/smoke.log 2>&1; echo smoke rc=$?; tail -6 gpurun_out/r2f/smoke.log
timeout 300 python tools/perf_probe.py rnghunt 2>&1 | grep pow8 | cut -c1-200
timeout 900 python bench.py > gpurun_out/r2f/bench_default.json 2> gpurun_out/r2f/bench_default.err; echo default rc=$?; cat gpurun_out/r2f/bench_default.json
for w in prbs_stream prbs_detect prbs_slicer prbs_shaped_slicer tx_stream rnghunt clt_stream lutopt_stream; do
  timeout 400 python bench.py --workload $w > gpurun_out/r2f/bench_$w.json 2> gpurun_out/r2f/bench_$w.err; echo $w rc=$?
done
for k in 7 15 23; do
  timeout 300 python bench.py --workload prbs_detect --k $k > gpurun_out/r2f/bench_prbs_detect_k$k.json 2>/dev/null; echo k$k rc=$?
done
timeout 300 python bench.py --workload clt_stream --noise-var-stream 8 > gpurun_out/r2f/bench_clt_stream_i16.json 2>/dev/null; echo i16 rc=$?
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2f/bench_reference.json 2>/dev/null; echo ref rc=$?
for f in gpurun_out/r2f/bench_*.json; do echo $f; cut -c1-160 $f; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2f/bench_default_launches.csv python bench.py --steps 1 --warmup 1 --log2-samples 34 --hunt-log2 22 --no-cpu-baseline > /dev/null 2>&1; echo launches rc=$?
timeout 300 python tools/ncu_hunt.py > gpurun_out/r2f/hunt_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"hunt_front_pow_kernel" -c 1 -o gpurun_out/r2f/prof_hunt python tools/ncu_hunt.py > gpurun_out/r2f/hunt_ncu.log 2>&1; cat gpurun_out/r2f/hunt_plain.log
