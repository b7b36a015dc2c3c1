set -x
mkdir -p gpurun_out/r3a
timeout 300 python -m pytest tests/test_gpu_clt.py -x -q -m gpu -k "word_stream" 2>&1 | tail -3
BLOCKS=37,128,148,0 WORDS_VARIANT=13 timeout 120 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r3a/probe_words9.txt
BLOCKS=37,128,148 WORDS_VARIANT=0 timeout 120 python tools/probe_words_blocks.py 2>&1 | tee -a gpurun_out/r3a/probe_words9.txt
