# host-side AddressSanitizer pass: libbbb_b200.so built with -fsanitize=address (build/asan/lib_asan.so) under the GPU tests
set -x
mkdir -p gpurun_out/asan
ASAN=$(gcc -print-file-name=libasan.so); ls -la $ASAN
export ASAN_OPTIONS=protect_shadow_gap=0:detect_leaks=0:abort_on_error=0:halt_on_error=1:log_path=gpurun_out/asan/asan
for t in tests/test_gpu_clt.py tests/test_gpu_prbs.py tests/test_gpu_rnghunt.py tests/test_shaper.py tests/test_zz_no_alloc.py; do
  LD_PRELOAD=$ASAN BBB_B200_LIB=$PWD/build/asan/lib_asan.so timeout 900 python -m pytest $t -m gpu -q -x -p no:cacheprovider > gpurun_out/asan/$(basename $t .py).log 2>&1; echo rc=$?; tail -4 gpurun_out/asan/$(basename $t .py).log | cut -c1-300
done
ls gpurun_out/asan; head -60 gpurun_out/asan/asan.* 2>/dev/null | cut -c1-250
