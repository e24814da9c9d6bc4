#!/bin/bash
# Round 2, pass B: kernel 5 first light (tests of kernel 5 first, then its timing check), then the whole GPU suite.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "kernel5" > gpurun_out/pytest_k5.log 2>&1; echo "pytest k5 rc=$?"
tail -30 gpurun_out/pytest_k5.log
timeout 600 python tests/checks/k5_check.py > gpurun_out/k5_check.log 2>&1; echo "k5 check rc=$?"; tail -12 gpurun_out/k5_check.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
