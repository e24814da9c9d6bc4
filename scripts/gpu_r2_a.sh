#!/bin/bash
# Round 2, pass A: state check (GPU tests), reference harness dumps + kernel-only timing, ncu of the harness,
# compute-sanitizer of the reference (documents SURVEY 9.6).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python oracle/make_ref_dumps.py > gpurun_out/ref_dumps.log 2>&1; echo "dumps rc=$?"; cat gpurun_out/ref_dumps.log | tail -8
timeout 600 python oracle/ref_kernel_time.py > gpurun_out/r02_reference_kernel_only.json 2> gpurun_out/ref_time.err; echo "time rc=$?"
cat gpurun_out/r02_reference_kernel_only.json
cat > /tmp/refncu.py <<'PY'
import sys; sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import refharness
from gato_python_b200.problems import iiwa_like
p = iiwa_like(K=50, seed=0)
h = refharness.RefHarness(14, 7, 50); h.upload(p)
for _ in range(3): h.run(mode=1)
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02_ncu_reference_launches.csv python /tmp/refncu.py > gpurun_out/ncu_ref.log 2>&1; echo "ncu ref rc=$?"
cat > /tmp/refsan.py <<'PY'
import sys; sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import refharness
from gato_python_b200.problems import iiwa_like
p = iiwa_like(K=50, seed=0)
h = refharness.RefHarness(14, 7, 50); h.set_padding(0); h.upload(p)
h.run(mode=1)
PY
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python /tmp/refsan.py > gpurun_out/r02_sanitizer_reference_memcheck.log 2>&1; echo "sanitizer ref rc=$?"
tail -30 gpurun_out/r02_sanitizer_reference_memcheck.log
ls -la gpurun_out gpurun_out/golden
