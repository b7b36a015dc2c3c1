set -x
mkdir -p gpurun_out/r3b
timeout 900 python -m pytest tests/test_gpu_clt.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
BLOCKS=0,0 WORDS_VARIANT=0 timeout 120 python tools/probe_words_blocks.py 2>&1 | tail -2
timeout 600 python bench.py --workload lutopt_stream > gpurun_out/r3b/bench_lutopt_stream.json 2> gpurun_out/r3b/err.log; cut -c1-200 gpurun_out/r3b/bench_lutopt_stream.json
timeout 600 python bench.py --workload lutopt_stream --log2-steps 30 > gpurun_out/r3b/bench_lutopt_stream_2p30.json 2>> gpurun_out/r3b/err.log; cut -c1-200 gpurun_out/r3b/bench_lutopt_stream_2p30.json
WORDS_VARIANT=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"clt256_words7_kernel" -c 1 -o gpurun_out/r3b/prof_words7 python tools/ncu_words.py > gpurun_out/r3b/ncu7.log 2>&1; tail -1 gpurun_out/r3b/ncu7.log
