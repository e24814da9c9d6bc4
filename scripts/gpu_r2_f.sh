#!/bin/bash
# Round 2, pass F: tensor-core experiment, chunk sweep of kernel 5, new boundary tests.
set -u
mkdir -p gpurun_out
./build/tf32_experiment > gpurun_out/r02_tf32_experiment.json 2>&1; echo "tf32 rc=$?"; cat gpurun_out/r02_tf32_experiment.json
timeout 300 python tests/checks/k5_check.py > gpurun_out/k5_check.log 2>&1; echo "k5 rc=$?"; grep -E "^k[25]_chunk" gpurun_out/k5_check.log | cut -c1-100; tail -1 gpurun_out/k5_check.log > gpurun_out/r02_k5_chunk_sweep.json
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "cmake or thread or dlpack or multiplan" > gpurun_out/pytest_new.log 2>&1; echo "pytest new rc=$?"; tail -8 gpurun_out/pytest_new.log
