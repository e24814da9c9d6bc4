"""Driver for ncu: a few launches of kernel 5 (or another kernel) on the benchmark batch.  argv: nsys kernel chunk reps"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like_batch_fast  # noqa: E402

n, kernel, chunk, reps = (int(x) for x in sys.argv[1:5])
b = iiwa_like_batch_fast(n, 14, 7, 50, seed=1000)
dev = torch.device("cuda", 0)
pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
pl.set_kernel(kernel)
if kernel == 5:
    pl.set_scaled_chunk(chunk)
d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
lam = torch.zeros((n, pl.n_l), device=dev); dz = torch.zeros((n, pl.n_z), device=dev)
its = torch.zeros(n, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream()
for _ in range(reps):
    pl.solve_batched_device(n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                            d["g"].data_ptr(), pl.n_z, d["c"].data_ptr(), pl.n_l, lam.data_ptr(), dz.data_ptr(),
                            its.data_ptr(), exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"],
                            stream=st.cuda_stream)
torch.cuda.synchronize()
print("mean iters", float(its.float().mean()))
