set -x
mkdir -p gpurun_out/r3c
timeout 200 python -m pytest tests/test_gpu_clt.py -x -q -m gpu -k "word_stream" 2>&1 | tail -2
BLOCKS=37,0,0 WORDS_VARIANT=0 timeout 100 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r3c/probe_words7_unroll2.txt
