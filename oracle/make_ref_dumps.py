"""Dump the intermediates of the UNMODIFIED reference kernels (oracle/ref_harness.cu -> libref_harness.so) on a
B200 into tests/golden/refdump_*.npz (TEST INFRASTRUCTURE ONLY).

    gpurun -- 'python oracle/make_ref_dumps.py'      # writes gpurun_out/golden/refdump_*.npz; copy to tests/golden/

Each fixture holds, for one seeded problem of gato_python_b200.problems (inputs are regenerated from the seed; their
CRC is stored): d_G_dense and d_C_dense right after gato_convert_kkt_format (src/gato_schur.cuh:902) -- the
bit-exact comparand of `gato_unpack` --, d_S / d_Pinv / d_gamma after gato_form_schur_jacobi + gato_form_ss
(stored NEGATED by the reference), d_G_dense after the Schur kernel (= G^-1), lambda, dz and the
iteration count.  Long-horizon cases keep only the unpack dumps and gamma (S and Pinv are 1.2 MB there; the GPU
test compares them live against the harness when it travelled to the box).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from make_golden import CASES, input_crc, make_problem  # noqa: E402

DUMP_CASES = ["pendulum5", "iiwa50_s0", "iiwa50_dense_s1", "iiwa50_rho1e-4", "iiwa256_s0"]

if __name__ == "__main__":
    import refharness

    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    for name in DUMP_CASES:
        case = CASES[name]
        prob = make_problem(case)
        h = refharness.RefHarness(prob["S"], prob["C"], prob["K"])
        h.upload(prob)
        d = h.run(mode=1)
        d0 = h.run(mode=0)        # the reference's own host functions on the same buffers: same kernels, same bits
        same = {k: bool(np.array_equal(d[k], d0[k])) for k in ("G_dense_pre", "C_dense", "S", "Pinv", "gamma", "G_inv")}
        h.close()
        keep = dict(G_dense_pre=d["G_dense_pre"], C_dense=d["C_dense"], gamma=d["gamma"], lam=d["lam"], dz=d["dz"],
                    iters=d["iters"], iters_host_api=d0["iters"], crc=np.uint32(input_crc(prob)), case=json.dumps(case))
        if prob["K"] <= 64:
            keep.update(S=d["S"], Pinv=d["Pinv"], G_inv=d["G_inv"])
        np.savez_compressed(os.path.join(out, f"refdump_{name}.npz"), **keep)
        print(name, "iters", d["iters"], "host-api iters", d0["iters"], "kernel-vs-host-api identical:", same)
