"""gato_solve_batched_device_noalloc under CUDA graph capture (after one warm-up call outside the capture): capture one
call on a side stream, replay it on new inputs, compare with the allocating entry point.  GPU box only."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like_batch_fast  # noqa: E402

dev = torch.device("cuda", 0)
n, S, C, K = 32, 14, 7, 50
b = iiwa_like_batch_fast(n, S, C, K, seed=11)
pl = Plan(S, C, K, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
ref = Plan(S, C, K, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
ref.set_kernel(5)
d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
lam = torch.zeros((n, pl.n_l), device=dev); dz = torch.zeros((n, pl.n_z), device=dev); its = torch.zeros(n, dtype=torch.int32, device=dev)


def call(stream):
    pl.solve_batched_device_noalloc(n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                                    d["g"].data_ptr(), pl.n_z, d["c"].data_ptr(), pl.n_l, lam.data_ptr(), dz.data_ptr(), its.data_ptr(),
                                    exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=stream)


side = torch.cuda.Stream()
with torch.cuda.stream(side):
    call(side.cuda_stream)            # warm-up: symbol address, function attributes, occupancy
side.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g, stream=side):
    call(torch.cuda.current_stream().cuda_stream)
for rep, seed in enumerate((12, 13)):
    b2 = iiwa_like_batch_fast(n, S, C, K, seed=seed)
    for k in d:
        d[k].copy_(torch.from_numpy(b2[k]))
    lam.fill_(float("nan")); dz.fill_(float("nan")); its.fill_(-1)
    g.replay()
    torch.cuda.synchronize()
    got = (lam.cpu().numpy().copy(), dz.cpu().numpy().copy(), its.cpu().numpy().copy())
    l2 = torch.zeros_like(lam); z2 = torch.zeros_like(dz); i2 = torch.zeros_like(its)
    ref.solve_batched_device(n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                             d["g"].data_ptr(), pl.n_z, d["c"].data_ptr(), pl.n_l, l2.data_ptr(), z2.data_ptr(), i2.data_ptr(),
                             exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    same = np.array_equal(got[0], l2.cpu().numpy()) and np.array_equal(got[1], z2.cpu().numpy()) and np.array_equal(got[2], i2.cpu().numpy())
    print("replay", rep, "same bits as gato_solve_batched_device:", same, "iters", got[2][:4])
    assert same
print("ok")
