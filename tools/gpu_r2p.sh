set -x
mkdir -p gpurun_out/r3a
BLOCKS=128,148,140 WORDS_VARIANT=14 timeout 120 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r3a/probe_words7_halfline.txt
BLOCKS=128,148 WORDS_VARIANT=0 timeout 120 python tools/probe_words_blocks.py 2>&1 | tee -a gpurun_out/r3a/probe_words7_halfline.txt
