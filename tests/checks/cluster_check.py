"""Kernel 3 (the one-CTA kernel over a thread-block cluster) vs the oracle and vs kernel 1; timing."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "checks"))
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import iiwa_like, pendulum5
from oracle import gato_oracle as orc
from v2_check import kernel_only, NAMES, rel

def check(prob, name, kernels=(3, 1)):
    S, C, K = prob["S"], prob["C"], prob["K"]
    lam_o, dz_o, info = orc.linsys_solve_info(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"],
        prob["C_col"], prob["C_val"], prob["g"], prob["c"], np.zeros(S*K), 1, prob["exit_tol"], prob["max_iters"], False, prob["rho"])
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    for kern in kernels:
        try:
            pl.set_kernel(kern)
        except Exception as e:
            print(f"[{name}] kernel {kern}: {e}"); continue
        lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
        med, mn, its, _ = kernel_only(pl, prob, reps=50)
        d = pl.solve_debug(prob["G_val"], prob["C_val"], prob["g"], prob["c"], prob["exit_tol"], prob["max_iters"], prob["rho"], dumps=False)
        dt = np.diff(d["clk_ns"][:10])
        print(f"[{name}] kernel v{pl.kernel()} cl={pl.cluster()} iters={inf['iters']} (oracle {info['iters']}) lam {rel(lam, lam_o):.2e} dz {rel(dz, dz_o):.2e} "
              f"| kernel-only us median {med:.1f} min {mn:.1f} | per iter {dt[7]/1e3/(inf['iters']+1):.3f} us", flush=True)
        print("     " + "  ".join(f"{n}={v/1e3:.1f}" for n, v in zip(NAMES, dt)), flush=True)
    pl.close()

if __name__ == "__main__":
    check(iiwa_like(S=4, C=2, K=2, seed=0), "K2")
    check(pendulum5(), "pendulum5")
    check(iiwa_like(K=12, seed=3), "iiwa12")
    check(iiwa_like(K=50, seed=0), "iiwa50", kernels=(3, 2))
    check(iiwa_like(K=50, seed=4, dense_G=True, dense_AB=True), "iiwa50-dense", kernels=(3, 2))
    check(iiwa_like(K=57, seed=0), "iiwa57")
    check(iiwa_like(K=108, seed=0), "iiwa108")
    check(iiwa_like(K=256, seed=0), "iiwa256")
    check(iiwa_like(K=432, seed=0), "iiwa432")
    check(iiwa_like(S=7, C=3, K=33, seed=2, preset="mild"), "S7K33")
    check(iiwa_like(S=4, C=2, K=512, seed=1, preset="mild"), "S4K512")
