#!/bin/bash
set -u
mkdir -p gpurun_out
for v in a b c; do
export GATO_B200_LIB=$PWD/build/dev_$v/libgato_b200.so
timeout 300 python tests/checks/k5_check.py > gpurun_out/k5_check_$v.log 2>&1; echo "variant $v rc=$?"; grep -E "^k5_chunk(0|8192)" gpurun_out/k5_check_$v.log | cut -c1-100
D="python tests/checks/k5_profile_driver.py 8192 5 8192 3"
ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/k5_launches_$v.csv $D > gpurun_out/ncu1.log 2>&1
grep -o 'gato_[a-z_0-9]*<[^>]*>\|"ns","[0-9]*"' gpurun_out/k5_launches_$v.csv | paste - - | tail -3 | head -2
done
