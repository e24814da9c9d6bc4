#!/bin/bash
# Round 2, pass H: the shared-matrix path of kernel 5 from the development build (build/dev), then a short bench.
set -u
mkdir -p gpurun_out
export GATO_B200_LIB=$PWD/build/dev/libgato_b200.so
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "kernel5 or shared" > gpurun_out/pytest_h.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_h.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-single > gpurun_out/bench_h.json 2> gpurun_out/bench_h.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_h.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_h.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "e2e_shared", d["e2e_shared_matrix"]["value"])
print("shared_resident", d["shared_matrix_resident"])
PY
