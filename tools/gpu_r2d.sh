# round 2, call D: shaper kernels without local memory, word-store policies, pow kernel DRAM writes, clt_stream e2e
set -x
mkdir -p gpurun_out/r2d
timeout 600 python -m pytest tests/test_shaper.py tests/test_gpu_prbs.py tests/test_gpu_rnghunt.py -m gpu -x -q 2>&1 | tail -4
for v in "" build/variants/lib_wst1.so build/variants/lib_wst2.so; do
  BBB_B200_LIB=$v timeout 200 python tools/perf_probe.py streams 2>&1 | grep lutopt | cut -c1-260
done
for w in clt_stream tx_stream prbs_shaped_slicer lutopt_stream; do
  timeout 300 python bench.py --workload $w > gpurun_out/r2d/bench_$w.json 2> gpurun_out/r2d/bench_$w.err; cut -c1-1200 gpurun_out/r2d/bench_$w.json; tail -2 gpurun_out/r2d/bench_$w.err
done
timeout 300 python tools/ncu_hunt.py > gpurun_out/r2d/hunt_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"hunt_front_pow_kernel" -c 1 -o gpurun_out/r2d/prof_hunt python tools/ncu_hunt.py > gpurun_out/r2d/hunt_ncu.log 2>&1
tail -1 gpurun_out/r2d/hunt_ncu.log
timeout 300 python tools/ncu_targets_dec.py > gpurun_out/r2d/tx_plain.log 2>&1; tail -2 gpurun_out/r2d/tx_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"shaper_stream_kernel|shaped_slicer_err_kernel|clt256_words2_kernel" -c 4 -o gpurun_out/r2d/prof_tx python tools/ncu_targets_dec.py > gpurun_out/r2d/ncu_tx.log 2>&1; tail -1 gpurun_out/r2d/ncu_tx.log
