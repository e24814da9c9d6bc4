#!/bin/bash
set -u
mkdir -p gpurun_out
for v in 2 3; do
  GATO_B200_LIB=$PWD/build/dev$v/libgato_b200.so timeout 300 python tests/checks/k5_check.py > gpurun_out/k5_check_ctas$v.log 2>&1; echo "ctas=$v rc=$?"; grep -E "^k[25]_chunk" gpurun_out/k5_check_ctas$v.log | cut -c1-110
done
