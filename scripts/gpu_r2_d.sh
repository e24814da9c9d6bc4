#!/bin/bash
set -u
mkdir -p gpurun_out
export GATO_B200_LIB=$PWD/build/dev/libgato_b200.so
timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "kernel5" > gpurun_out/pytest_k5.log 2>&1; echo "pytest k5 rc=$?"; tail -3 gpurun_out/pytest_k5.log
timeout 300 python tests/checks/k5_check.py > gpurun_out/k5_check_dev.log 2>&1; echo "rc=$?"; grep -E "^k[25]_chunk" gpurun_out/k5_check_dev.log | cut -c1-100 | tail -9; grep -o '"k5_vs_k2.*' gpurun_out/k5_check_dev.log
D="python tests/checks/k5_profile_driver.py 8192 5 8192 3"
ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/k5_launches.csv $D > gpurun_out/ncu1.log 2>&1
grep -o 'gato_[a-z_0-9]*<[^>]*>\|"ns","[0-9]*"' gpurun_out/k5_launches.csv | paste - - | tail -4
