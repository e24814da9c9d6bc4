"""Cycle stamps of the second half of kernel 5 (development build with -DGATO_V5_TRACE): where one PCG iteration's time
goes for a system that runs alone and for one that shares the SM with two others.  GPU box only."""
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
names = ["apply(p)", "dot+tree", "wait barrier A", "sum, alpha, updates, publish r", "wait barrier B", "apply(r)", "dot+tree",
         "wait barrier C", "sum, exit test, beta, p, publish", "wait barrier D"]
out = {}
for n in (1, 8192):
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
    tr = np.zeros(96, np.int64)
    L = solver.lib()
    L.gato_debug_read_trace.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_longlong), C.c_int]
    assert L.gato_debug_read_trace(pl._h, n, tr.ctypes.data_as(C.POINTER(C.c_longlong)), 96) == 0
    tr = tr.reshape(8, 12)[2:7, :11]                 # iterations 2..6
    dt = np.diff(tr, axis=1).mean(axis=0)
    out[f"nsys={n}"] = {"cycles_per_iteration": float((tr[1:, 0] - tr[:-1, 0]).mean()),
                        "phases": {f"{i}: {names[i]}": float(dt[i]) for i in range(10)}}
    pl.close()
print(json.dumps(out, indent=1))
