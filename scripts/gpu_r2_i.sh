#!/bin/bash
# Round 2, pass I: development build (build/dev) of the first half of kernel 5 -- phase trace, kernel-5 parity tests, short bench.
set -u
mkdir -p gpurun_out
export GATO_B200_LIB=$PWD/build/dev/libgato_b200.so
python tests/checks/s_trace.py 8192 > gpurun_out/s_trace.json 2> gpurun_out/s_trace.err; tail -3 gpurun_out/s_trace.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/s_trace.json"))
print("cycles/system", d["cycles_per_system"])
for k, v in d["phases_cycles_per_system"].items():
    print(f"  {v:8.0f}  {k}")
PY
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -q -x -k "kernel5 or shared" > gpurun_out/pytest_i.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_i.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-single --parity-sample 8 > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_i.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_i.json").read().strip().splitlines()[-1])
print("value", d["value"], "halves", d["roofline"]["kernel_5_halves_ms_per_step"], "parity", d["parity_sample"]["pass"])
print("shared_resident", d["shared_matrix_resident"]["value"])
PY
