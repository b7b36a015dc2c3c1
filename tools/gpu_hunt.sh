set -x
python -m pytest tests/test_gpu_rnghunt.py -m gpu -x -q 2>&1 | tail -3
python tools/perf_probe.py rnghunt > gpurun_out/probe_hunt.jsonl 2>&1; cat gpurun_out/probe_hunt.jsonl
python tools/ncu_hunt.py > gpurun_out/hunt_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hunt_front_kernel -c 1 -o gpurun_out/prof_hunt_front python tools/ncu_hunt.py > gpurun_out/hunt_ncu.log 2>&1
tail -2 gpurun_out/hunt_ncu.log
