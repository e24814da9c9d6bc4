"""TMA-staged persistent v2 vs direct loads: bitwise equality and timing (single solve + batch)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "checks"))
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import iiwa_like, iiwa_like_batch_fast
from v2_check import kernel_only, NAMES

prob = iiwa_like(K=50, seed=0)
pl = Plan(14, 7, 50, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
res = {}
for on in (1, 0):
    pl.set_staging(on)
    lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
    med, mn, its, _ = kernel_only(pl, prob)
    d = pl.solve_debug(prob["G_val"], prob["C_val"], prob["g"], prob["c"], prob["exit_tol"], prob["max_iters"], prob["rho"], dumps=False)
    dt = np.diff(d["clk_ns"][:10])
    res[on] = (lam, dz)
    print(f"[single staging={on}] iters {inf['iters']} kernel-only us median {med:.1f} min {mn:.1f}")
    print("     " + "  ".join(f"{n}={v/1e3:.1f}" for n, v in zip(NAMES, dt)))
print("single bitwise equal:", np.array_equal(res[0][0], res[1][0]), np.array_equal(res[0][1], res[1][1]))
pl.close()
for nsys in (8192, 1000, 149):
    b = iiwa_like_batch_fast(nsys)
    pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=nsys)
    outs = {}
    for on in (1, 0):
        pl.set_staging(on)
        med, mn, its, lam = kernel_only(pl, dict(b), nsys=nsys, reps=8, batch=b)
        outs[on] = (its, lam)
        print(f"[batch {nsys} staging={on}] {med/1e3:.3f} ms -> {nsys/(med*1e-6):.0f} solves/s, mean iters {its.mean():.2f}")
    print("batch bitwise equal:", np.array_equal(outs[0][0], outs[1][0]), np.array_equal(outs[0][1], outs[1][1]))
    pl.close()
