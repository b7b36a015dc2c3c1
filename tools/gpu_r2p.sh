set -x
mkdir -p gpurun_out/r2u
BLOCKS=148,140,136,132,128,124,120,112,104,96,128 WORDS_VARIANT=0 timeout 120 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r2u/probe_words7_sweep2.txt
