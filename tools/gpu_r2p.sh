set -x
timeout 300 python tools/probe_words_blocks.py
