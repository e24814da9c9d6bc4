"""Generate tests/golden/ref_*.npz by running the UNMODIFIED reference CUDA build (oracle/_ref,
see build_ref.py) on a B200 (TEST INFRASTRUCTURE ONLY).

    gpurun -- 'python oracle/make_golden.py'      # writes gpurun_out/golden/*.npz; copy to tests/golden/

Each fixture holds the generator arguments (inputs are regenerated from the seed by
gato_python_b200.problems, and their CRC is stored), the reference's lambda and dz, and the
iteration count parsed from its stdout.  The reference's reductions use global atomics, so its
lambda is not bit-reproducible run to run; tests compare within 1e-4 relative.
"""
import json
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from gato_python_b200.problems import iiwa_like, pendulum5  # noqa: E402

CASES = {
    "pendulum5": dict(gen="pendulum5", kw={}),
    "iiwa50_s0": dict(gen="iiwa_like", kw=dict(K=50, seed=0)),
    "iiwa50_s1": dict(gen="iiwa_like", kw=dict(K=50, seed=1)),
    "iiwa50_dense_s1": dict(gen="iiwa_like", kw=dict(K=50, seed=1, dense_G=True, dense_AB=True)),
    "iiwa50_mild_rho1e-1": dict(gen="iiwa_like", kw=dict(K=50, seed=2, preset="mild", rho=1e-1)),
    "iiwa50_rho1e-4": dict(gen="iiwa_like", kw=dict(K=50, seed=3, rho=1e-4)),
    "iiwa256_s0": dict(gen="iiwa_like", kw=dict(K=256, seed=0)),
}


def make_problem(case):
    return pendulum5() if case["gen"] == "pendulum5" else iiwa_like(**case["kw"])


def input_crc(prob):
    h = 0
    for k in ("G_row", "G_col", "G_val", "C_row", "C_col", "C_val", "g", "c"):
        h = zlib.crc32(np.ascontiguousarray(prob[k]).tobytes(), h)
    return h


if __name__ == "__main__":
    import refrun

    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    for name, case in CASES.items():
        prob = make_problem(case)
        d = refrun.ref_module_dir(prob["S"], prob["C"], prob["K"])
        r = refrun.run_module(d, prob, testiters=1)
        # the last S entries of dz come from an out-of-bounds read in the reference
        np.savez(os.path.join(out, f"ref_{name}.npz"), lam=r["lam"], dz=r["dz"], iters=r["iters"],
                 crc=np.uint32(input_crc(prob)), case=json.dumps(case))
        print(name, "iters", r["iters"], "crc", input_crc(prob))
