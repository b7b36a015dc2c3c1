# round 2, call A: new rnghunt pipeline -- parity, config-scale tests, throughput, ncu
set -x
mkdir -p gpurun_out/r2a
nvidia-smi -L; nproc
timeout 600 python -m pytest tests/test_gpu_rnghunt.py tests/test_gpu_cli.py -m gpu -x -q 2>&1 | tail -5
timeout 900 python -m pytest tests/test_gpu_config_scale.py -m gpu -x -q --durations=8 2>&1 | tail -16
timeout 300 python tools/perf_probe.py rnghunt > gpurun_out/r2a/probe_hunt.jsonl 2>&1; cat gpurun_out/r2a/probe_hunt.jsonl
timeout 300 python bench.py --workload rnghunt --log2-candidates 22 > gpurun_out/r2a/bench_rnghunt.json 2> gpurun_out/r2a/bench_rnghunt.err; cat gpurun_out/r2a/bench_rnghunt.json
timeout 300 python tools/ncu_hunt.py > gpurun_out/r2a/hunt_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"hunt_front_warp_kernel|hunt_check32_kernel" -c 3 -o gpurun_out/r2a/prof_hunt python tools/ncu_hunt.py > gpurun_out/r2a/hunt_ncu.log 2>&1
tail -2 gpurun_out/r2a/hunt_ncu.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2a/hunt_launches.csv python tools/ncu_hunt.py > /dev/null 2>&1
