set -x
mkdir -p gpurun_out/r3a
BLOCKS=128,120,116,112,108,104,100,96,88,128,112,104 WORDS_VARIANT=0 timeout 120 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r3a/probe_words7_sweep3.txt
