# round 2, call B: power kernel for rnghunt, pulser/CLI tests, lane-pair word stores
set -x
mkdir -p gpurun_out/r2b
timeout 900 python -m pytest tests/test_gpu_rnghunt.py tests/test_gpu_cli.py tests/test_shaper.py tests/test_gpu_clt.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python -m pytest tests/test_gpu_config_scale.py -m gpu -x -q --durations=8 -k "rnghunt or prbs" 2>&1 | tail -12
timeout 300 python tools/perf_probe.py rnghunt > gpurun_out/r2b/probe_hunt.jsonl 2>&1; cut -c1-210 gpurun_out/r2b/probe_hunt.jsonl
timeout 300 python tools/perf_probe.py streams > gpurun_out/r2b/probe_streams.jsonl 2>&1; cut -c1-300 gpurun_out/r2b/probe_streams.jsonl
timeout 300 python bench.py --workload lutopt_stream > gpurun_out/r2b/bench_lutopt_stream.json 2> gpurun_out/r2b/bench_lutopt_stream.err; cut -c1-900 gpurun_out/r2b/bench_lutopt_stream.json; tail -2 gpurun_out/r2b/bench_lutopt_stream.err
timeout 300 python bench.py --workload clt_stream > gpurun_out/r2b/bench_clt_stream.json 2> gpurun_out/r2b/bench_clt_stream.err; cut -c1-900 gpurun_out/r2b/bench_clt_stream.json
timeout 300 python tools/ncu_hunt.py > gpurun_out/r2b/hunt_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"hunt_front_pow_kernel" -c 1 -o gpurun_out/r2b/prof_hunt python tools/ncu_hunt.py > gpurun_out/r2b/hunt_ncu.log 2>&1
tail -2 gpurun_out/r2b/hunt_ncu.log; cat gpurun_out/r2b/hunt_plain.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2b/hunt_launches.csv python tools/ncu_hunt.py > /dev/null 2>&1
