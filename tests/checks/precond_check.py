"""Block-Jacobi vs stair preconditioner (gato_plan_set_preconditioner) on kernel 2: iterations and device time for the
IIWA-50 single solve and for the 8192-system benchmark batch.  GPU box only; prints one JSON line."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like, iiwa_like_batch_fast  # noqa: E402

dev = torch.device("cuda", 0)
st = torch.cuda.current_stream()
out = {}
prob = iiwa_like(14, 7, 50, seed=0)
pl = Plan(14, 7, 50, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)  # noqa: E731
G, Cv, g1, c1 = t(prob["G_val"]), t(prob["C_val"]), t(prob["g"]), t(prob["c"])
l1 = torch.empty(pl.n_l, device=dev); z1 = torch.empty(pl.n_z, device=dev); i1 = torch.empty(1, dtype=torch.int32, device=dev)
n = 8192
b = iiwa_like_batch_fast(n, 14, 7, 50, seed=1000)
pb = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
pb.set_kernel(2)
d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
lam = torch.zeros((n, pb.n_l), device=dev); dz = torch.zeros((n, pb.n_z), device=dev); its = torch.zeros(n, dtype=torch.int32, device=dev)
for name, which in (("stair", 0), ("block_jacobi", 1)):
    pl.set_preconditioner(which); pb.set_preconditioner(which)
    def run1():
        pl.solve_batched_device(1, G.data_ptr(), 0, Cv.data_ptr(), 0, g1.data_ptr(), 0, c1.data_ptr(), 0, l1.data_ptr(),
                                z1.data_ptr(), i1.data_ptr(), exit_tol=prob["exit_tol"], max_iters=prob["max_iters"],
                                rho=prob["rho"], stream=st.cuda_stream)
    def runb():
        pb.solve_batched_device(n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                                d["g"].data_ptr(), pb.n_z, d["c"].data_ptr(), pb.n_l, lam.data_ptr(), dz.data_ptr(),
                                its.data_ptr(), exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=st.cuda_stream)
    res = {}
    for key, fn, reps in (("single", run1, 200), ("batch8192", runb, 10)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); e1.synchronize()
        res[key + "_us"] = e0.elapsed_time(e1) * 1e3 / reps
    res["single_iters"] = int(i1.item())
    res["batch_mean_iters"] = float(its.float().mean())
    res["batch_solves_per_s"] = n / res["batch8192_us"] * 1e6
    out[name] = res
print(json.dumps(out))
