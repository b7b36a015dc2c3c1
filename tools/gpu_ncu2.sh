set -x
python tools/ncu_targets.py > gpurun_out/ncu2_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'prbs_stream_kernel|prbs_detect_kernel|hunt_front_kernel|hunt_check_kernel|clt256_s8v2_kernel|clt256_words_kernel|shaper_stream_kernel|shaped_slicer_err_kernel' -c 14 -o gpurun_out/prof_secondary python tools/ncu_targets.py > gpurun_out/ncu2.log 2>&1
tail -5 gpurun_out/ncu2.log; cat gpurun_out/ncu2_plain.log | tail -2
