#!/bin/bash
# Builds build/variants/lib_<name>.so for several generator settings of the one-pass clt256 kernel
# (experiments; run with BBB_B200_LIB=... bench.py).  usage: tools/build_variants.sh "name:ENV=V ENV=V" ...
set -e
mkdir -p build/variants
cp basebandboard_b200/csrc/clt256_v2_gen.inc build/variants/_saved.inc
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}
  env $envs python tools/gen_clt256_v2.py 2>&1 | tail -1
  make basebandboard_b200/libbbb_b200.so 2>&1 | grep -i "error" || true
  grep -A2 "clt256_hist2" build/clt_spec.ptxas.log | tail -2
  cp basebandboard_b200/libbbb_b200.so build/variants/lib_$name.so
done
cp build/variants/_saved.inc basebandboard_b200/csrc/clt256_v2_gen.inc
make basebandboard_b200/libbbb_b200.so 2>&1 | grep -i "error" || true
