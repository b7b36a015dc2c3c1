set -x
mkdir -p gpurun_out/r2k
for v in "" build/variants/lib_share1.so; do
  BBB_B200_LIB=$v timeout 300 python bench.py --log2-samples 38 --steps 3 --warmup 3 --no-cpu-baseline --no-secondary 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['value'], d['roofline']['frac'], d['result_check']['sumsq'])"
done
BBB_B200_LIB=build/variants/lib_share1.so timeout 600 python -m pytest tests/test_gpu_clt.py tests/test_shaper.py tests/test_gpu_prbs.py -m gpu -x -q 2>&1 | tail -3
for v in "" build/variants/lib_share1.so; do
  BBB_B200_LIB=$v timeout 300 python bench.py --workload prbs_slicer --log2-bits 34 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v slicer', d['value'])"
  BBB_B200_LIB=$v timeout 300 python bench.py --workload clt_stream 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v clt_stream', d['value'])"
done
