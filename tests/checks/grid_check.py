"""Kernel 4 (the register-resident kernel over a cooperative grid) vs the oracle; timing."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "checks"))
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import iiwa_like
from oracle import gato_oracle as orc
from v2_check import kernel_only, rel

def check(prob, name, kernel):
    S, C, K = prob["S"], prob["C"], prob["K"]
    lam_o, dz_o, info = orc.linsys_solve_info(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"],
        prob["C_col"], prob["C_val"], prob["g"], prob["c"], np.zeros(S*K), 1, prob["exit_tol"], prob["max_iters"], False, prob["rho"])
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    if kernel:
        pl.set_kernel(kernel)
    lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
    med, mn, its, _ = kernel_only(pl, prob, reps=20)
    print(f"[{name}] kernel v{pl.kernel()} cl={pl.cluster()} iters={inf['iters']} (oracle {info['iters']}) lam {rel(lam, lam_o):.2e} dz {rel(dz, dz_o):.2e} "
          f"| kernel-only us median {med:.1f} min {mn:.1f} | {med/(inf['iters']+1):.2f} us/iter incl. fixed", flush=True)
    pl.close()

if __name__ == "__main__":
    check(iiwa_like(S=4, C=2, K=40, seed=0), "S4K40", 4)
    check(iiwa_like(K=57, seed=0), "iiwa57", 4)
    check(iiwa_like(K=256, seed=0), "iiwa256", 4)
    check(iiwa_like(K=1000, seed=0), "iiwa1000", 0)
    check(iiwa_like(K=3000, seed=0), "iiwa3000", 0)
    check(iiwa_like(S=4, C=2, K=2000, seed=1, preset="mild"), "S4K2000", 0)
