"""Cycle sums per phase of the FIRST half of kernel 5 (development build with -DGATO_S_TRACE): thread 0 of CTA 0 adds up
clock64() differences over every system it takes in an 8192-system call.  GPU box only."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan, solver  # noqa: E402
from gato_python_b200.problems import iiwa_like_batch_fast  # noqa: E402

dev = torch.device("cuda", 0)
st = torch.cuda.current_stream()
names = {0: "loop head", 1: "clear shared memory + barrier", 10: "wait for the staged values (TMA)", 2: "unpack scatter + barrier",
         3: "R^-1 / Q^-1 (pass 0) + barrier", 4: "phi, BR, theta, gamma (pass 1)", 5: "region copy to global, phi -> registers, barrier",
         6: "FP64 elimination (14 pivots)", 7: "M, gamma^, stores, barrier", 8: "T = phi M^T (FP64)", 9: "psi = M T (FP64) + store",
         11: "end-of-system barriers", 12: "(2a) scatter of C's values", 13: "(2b) scatter of G's values and g"}
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
b = iiwa_like_batch_fast(n, 14, 7, 50, seed=1000)
pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
pl.set_kernel(5)
pl.set_scaled_variant(2)
d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
lam = torch.zeros((n, pl.n_l), device=dev); dz = torch.zeros((n, pl.n_z), device=dev); its = torch.zeros(n, dtype=torch.int32, device=dev)
for _ in range(3):
    pl.solve_batched_device(n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                            d["g"].data_ptr(), pl.n_z, d["c"].data_ptr(), pl.n_l, lam.data_ptr(), dz.data_ptr(),
                            its.data_ptr(), exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=st.cuda_stream)
tr = np.zeros(248 + 24, np.int64)
L = solver.lib()
L.gato_debug_read_trace.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_longlong), C.c_int]
assert L.gato_debug_read_trace(pl._h, n, tr.ctypes.data_as(C.POINTER(C.c_longlong)), len(tr)) == 0
acc = tr[248:][1:]
nsys_cta0 = (n + 147) // 148
tot = float(acc.sum())
print(json.dumps({"systems_taken_by_cta0": nsys_cta0, "cycles_per_system": tot / nsys_cta0,
                  "phases_cycles_per_system": {f"{i}: {names[i]}": float(acc[i]) / nsys_cta0 for i in (0, 1, 10, 12, 13, 2, 3, 4, 5, 6, 7, 8, 9, 11)},
                  "share": {f"{i}": float(acc[i]) / tot for i in (0, 1, 10, 12, 13, 2, 3, 4, 5, 6, 7, 8, 9, 11)}}, indent=1))
pl.close()
