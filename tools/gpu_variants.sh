#!/bin/bash
# bench every build/variants/lib_*.so on the fused CLT workload (experiments)
for f in build/variants/lib_*.so; do
  n=$(basename $f .so)
  ok=$(BBB_B200_LIB=$PWD/$f python -m pytest tests/test_gpu_clt.py -m gpu -x -q -k "spec256_matches_oracle or int8_stream" 2>&1 | tail -1)
  v=$(BBB_B200_LIB=$PWD/$f python bench.py --log2-samples ${LOG2:-36} --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), d['roofline']['frac'])")
  echo "$n $v | $ok" | tee -a gpurun_out/variants.log
done
