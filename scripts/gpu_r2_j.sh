#!/bin/bash
# Round 2, pass J: whole GPU suite and both bench arms with the production build (bank-sorted scatter, shared-matrix path).
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_ours.err
timeout 600 python bench.py --kernel 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ours_kernel2.json 2> gpurun_out/bench_k2.err; echo "bench k2 rc=$?"
python - <<'PY'
import json
for f in ("bench_ours", "bench_ours_kernel2"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, "value", d["value"], "e2e", d["e2e"]["value"], "parity", d["parity_sample"]["pass"], "single", d.get("single_solve", {}).get("kernel_us_median"))
    print("  shared", d["shared_matrix_resident"]["value"], d["e2e_shared_matrix"]["value"])
PY
