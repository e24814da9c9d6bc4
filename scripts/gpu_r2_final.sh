#!/bin/bash
# Round 2, final check: what the driver runs at round end -- the GPU suite, smoke(), bench.py (both arms) at N = 1.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_default.json 2> gpurun_out/bench_ref_default.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref_default.json
T0=$(date +%s); timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$? wall $(( $(date +%s) - T0 )) s"; cut -c1-400 gpurun_out/bench_default.json
