set -x
mkdir -p gpurun_out/r2r
timeout 900 python -m pytest tests/test_gpu_clt.py tests/test_gpu_api.py -x -q -m gpu 2>&1 | tail -3
BLOCKS=0,148,0 WORDS_VARIANT=0 timeout 120 python tools/probe_words_blocks.py 2>&1 | tee gpurun_out/r2r/probe_default.txt
timeout 600 python bench.py --workload lutopt_stream > gpurun_out/r2r/bench_lutopt_stream.json 2> gpurun_out/r2r/bench_lutopt_stream.err; tail -c 1500 gpurun_out/r2r/bench_lutopt_stream.json
timeout 200 python tools/ncu_words.py
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"clt256_words7_kernel" -c 1 -o gpurun_out/r2r/prof_words7 python tools/ncu_words.py > gpurun_out/r2r/ncu7.log 2>&1; tail -2 gpurun_out/r2r/ncu7.log
