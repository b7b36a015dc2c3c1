set -x
mkdir -p gpurun_out/r2n
WORDS_VARIANT=4 timeout 200 python tools/ncu_words.py
WORDS_VARIANT=4 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"clt256_words4_kernel" -c 1 -o gpurun_out/r2n/prof_words4 python tools/ncu_words.py > gpurun_out/r2n/ncu4.log 2>&1; tail -1 gpurun_out/r2n/ncu4.log
