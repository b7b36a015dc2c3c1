set -x
mkdir -p gpurun_out/r2n
timeout 200 python tools/ncu_words.py
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"clt256_words2_kernel" -c 1 -o gpurun_out/r2n/prof_words python tools/ncu_words.py > gpurun_out/r2n/ncu.log 2>&1; tail -1 gpurun_out/r2n/ncu.log
