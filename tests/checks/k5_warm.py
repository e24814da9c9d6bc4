"""True warm start on a device-filling batch: kernel 5 (correction form) against the one-CTA kernel, 8192 IIWA-50 systems,
lambda0 = the solution scaled by 0.9 .. 1.1 per system.  GPU box only."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan, WARM_TRUE  # noqa: E402
from gato_python_b200.problems import iiwa_like_batch_fast  # noqa: E402

n = 8192
b = iiwa_like_batch_fast(n, 14, 7, 50, seed=1000)
dev = torch.device("cuda", 0)
st = torch.cuda.current_stream()
d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
out = {}
lam0 = None
for kernel in (5, 2):
    pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
    pl.set_kernel(kernel)
    lam = torch.zeros((n, pl.n_l), device=dev); dz = torch.zeros((n, pl.n_z), device=dev); its = torch.zeros(n, dtype=torch.int32, device=dev)

    def run(warm, lin):
        pl.solve_batched_device(n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                                d["g"].data_ptr(), pl.n_z, d["c"].data_ptr(), pl.n_l, lam.data_ptr(), dz.data_ptr(), its.data_ptr(),
                                dlam_in=lin.data_ptr() if lin is not None else 0, ldl=pl.n_l if lin is not None else 0,
                                exit_tol=b["exit_tol"], max_iters=b["max_iters"], warm_start=warm, rho=b["rho"], stream=st.cuda_stream)

    def timed(warm, lin):
        for _ in range(3):
            run(warm, lin)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(st)
        for _ in range(10):
            run(warm, lin)
        e1.record(st); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 10, float(its.float().mean())
    cold_ms, cold_it = timed(0, None)
    if lam0 is None:
        rng = np.random.default_rng(1)
        lam0 = (lam * torch.from_numpy((0.9 + 0.2 * rng.random((n, 1))).astype(np.float32)).to(dev)).contiguous()
    warm_ms, warm_it = timed(WARM_TRUE, lam0)
    out[f"kernel {kernel}"] = {"cold": {"ms": cold_ms, "mean_iters": cold_it, "solves_per_s": n / cold_ms * 1e3},
                               "true_warm_start": {"ms": warm_ms, "mean_iters": warm_it, "solves_per_s": n / warm_ms * 1e3}}
    pl.close()
print(json.dumps(out, indent=1))
