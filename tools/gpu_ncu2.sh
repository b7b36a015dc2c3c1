set -x
python tools/ncu_targets.py > gpurun_out/ncu2_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'prbs_stream_kernel|prbs_detect_kernel|hunt_front_kernel|hunt_check_kernel|clt256_s8_kernel|clt256_words_kernel' -c 8 -o gpurun_out/prof_secondary python tools/ncu_targets.py > gpurun_out/ncu2.log 2>&1
tail -5 gpurun_out/ncu2.log; cat gpurun_out/ncu2_plain.log | tail -2
