"""Kernel-only time of the UNMODIFIED reference kernels on a B200 (TEST INFRASTRUCTURE / BASELINE ONLY):
BASELINE.md section 2.1 rows R3 (five kernels, pre-allocated buffers, no malloc/free/sync/printf) and R4 (us per PCG
iteration), through oracle/ref_harness.cu.

    gpurun -- 'python oracle/ref_kernel_time.py > gpurun_out/r02_reference_kernel_only.json'
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from gato_python_b200.problems import iiwa_like, pendulum5  # noqa: E402

if __name__ == "__main__":
    import refharness

    out = {}
    for name, prob in (("pendulum5", pendulum5()), ("iiwa50_s0", iiwa_like(K=50, seed=0)),
                       ("iiwa256_s0", iiwa_like(K=256, seed=0))):
        if refharness.harness_path(prob["S"], prob["C"], prob["K"]) is None:
            out[name] = None
            continue
        out[name] = refharness.kernel_only_summary(prob["S"], prob["C"], prob["K"], prob,
                                                   reps=1000 if prob["K"] <= 64 else 300)
    print(json.dumps(out, indent=1))
