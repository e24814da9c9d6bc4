"""Kernel 3: single-solve latency vs cluster size (CTAs per system)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "checks"))
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import iiwa_like
from v2_check import kernel_only

for K, sizes in ((256, (0, 5, 6, 7, 8, 9, 10, 11, 13, 16)), (108, (0, 2, 3, 4, 6, 8)), (57, (0, 2, 3, 4)), (50, (2, 3, 4)), (432, (0, 8, 12, 16))):
    prob = iiwa_like(K=K, seed=0)
    pl = Plan(14, 7, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    pl.set_kernel(3)
    ref = None
    for n in sizes:
        try:
            pl.set_cluster(n)
        except Exception as e:
            print(f"K={K} ncta={n}: {e}"); continue
        lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
        if ref is None:
            ref = lam
        med, mn, its, _ = kernel_only(pl, prob, reps=40)
        err = float(np.linalg.norm(lam - ref) / np.linalg.norm(ref))
        print(f"K={K} ncta={n:2d} -> cluster {pl.cluster()} iters {inf['iters']} kernel-only us median {med:.1f} min {mn:.1f} "
              f"({med / (inf['iters'] + 1):.3f} us/iter incl. fixed) rel diff vs first {err:.1e}", flush=True)
    pl.close()
