set -x
mkdir -p gpurun_out/r2y
timeout 300 python -m pytest tests/test_gpu_clt.py -x -q -m gpu -k "word_stream" 2>&1 | tail -3
for v in 12 0 12 0; do BLOCKS=37,0 WORDS_VARIANT=$v timeout 120 python tools/probe_words_blocks.py 2>&1 | tee -a gpurun_out/r2y/probe_words7_straight.txt; done
