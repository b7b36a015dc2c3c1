set -x
mkdir -p gpurun_out/r2q
timeout 600 python -m pytest tests/test_gpu_clt.py -x -q -m gpu -k "word_stream" 2>&1 | tail -5
WORDS_VARIANT=6 timeout 300 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r2q/probe_words6_blocks.txt
WORDS_VARIANT=0 timeout 300 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r2q/probe_words2_blocks.txt
