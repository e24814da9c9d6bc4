"""Where does the host-buffer batched step spend its time?  Same call with max_iters = 0 (transfers + Schur only)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import iiwa_like_batch_fast
nsys = 8192
b = iiwa_like_batch_fast(nsys)
pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=nsys)
host = {k: torch.from_numpy(b[k]).pin_memory() for k in ("G_val", "C_val", "g", "c")}
hv = {k: v.numpy() for k, v in host.items()}
outs = (torch.empty((nsys, pl.n_l)).pin_memory().numpy(), torch.empty((nsys, pl.n_z)).pin_memory().numpy(),
        torch.empty(nsys, dtype=torch.int32).pin_memory().numpy())
def run(mi, chunk, reps=8):
    pl.set_pipeline(chunk)
    for _ in range(2):
        pl.solve_batched(hv["G_val"], hv["C_val"], hv["g"], hv["c"], None, b["exit_tol"], mi, 0, b["rho"], out=outs)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps):
        pl.solve_batched(hv["G_val"], hv["C_val"], hv["g"], hv["c"], None, b["exit_tol"], mi, 0, b["rho"], out=outs)
    return (time.perf_counter() - t0) / reps * 1e3
for chunk in (0, 444, 888, 1184, 1776, 2368):
    print(f"chunk {chunk}: full {run(500, chunk):.3f} ms   max_iters=0 {run(0, chunk):.3f} ms", flush=True)
pl.close()
