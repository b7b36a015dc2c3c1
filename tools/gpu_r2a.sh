# round 2, call A: new rnghunt pipeline + chunked host streams -- parity, config-scale tests, throughput, ncu
set -x
mkdir -p gpurun_out/r2a
nvidia-smi -L; nproc
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_config_scale.py 2>&1 | tail -5
timeout 900 python -m pytest tests/test_gpu_config_scale.py -m gpu -x -q --durations=8 2>&1 | tail -16
timeout 300 python tools/perf_probe.py rnghunt > gpurun_out/r2a/probe_hunt.jsonl 2>&1; cat gpurun_out/r2a/probe_hunt.jsonl
for w in prbs_stream tx_stream clt_stream; do
  timeout 300 python bench.py --workload $w > gpurun_out/r2a/bench_$w.json 2> gpurun_out/r2a/bench_$w.err; cut -c1-700 gpurun_out/r2a/bench_$w.json; tail -2 gpurun_out/r2a/bench_$w.err
done
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r2a/bench_default.json 2> gpurun_out/r2a/bench_default.err; cat gpurun_out/r2a/bench_default.json; tail -3 gpurun_out/r2a/bench_default.err
timeout 300 python tools/ncu_hunt.py > gpurun_out/r2a/hunt_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"hunt_front_warp_kernel|hunt_check32_kernel" -c 3 -o gpurun_out/r2a/prof_hunt python tools/ncu_hunt.py > gpurun_out/r2a/hunt_ncu.log 2>&1
tail -2 gpurun_out/r2a/hunt_ncu.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2a/hunt_launches.csv python tools/ncu_hunt.py > /dev/null 2>&1
