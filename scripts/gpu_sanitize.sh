#!/bin/bash
# compute-sanitizer passes over every kernel on small inputs (tests/checks/sanitize_run.py); logs under gpurun_out/.
set -u
mkdir -p gpurun_out
S=/usr/local/cuda/bin/compute-sanitizer
timeout 600 $S --tool memcheck --error-exitcode 9 python tests/checks/sanitize_run.py > gpurun_out/sanitize_memcheck.log 2>&1; echo "memcheck rc=$?"
timeout 600 $S --tool synccheck --error-exitcode 9 python tests/checks/sanitize_run.py > gpurun_out/sanitize_synccheck.log 2>&1; echo "synccheck rc=$?"
timeout 900 $S --tool racecheck --error-exitcode 9 python tests/checks/sanitize_run.py k2 > gpurun_out/sanitize_racecheck_k2.log 2>&1; echo "racecheck(k2) rc=$?"
timeout 600 $S --tool initcheck --error-exitcode 9 python tests/checks/sanitize_run.py k2 > gpurun_out/sanitize_initcheck_k2.log 2>&1; echo "initcheck(k2) rc=$?"
tail -n 6 gpurun_out/sanitize_*.log
