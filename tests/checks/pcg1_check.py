"""Single-reduction PCG (gato_plan_set_pcg(1)) vs the reference update order (0) and the oracle: iteration counts,
errors against an FP64 solve of the same Schur system, kernel-only time; then the whole 8192-system batch."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "checks"))
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import pendulum5, iiwa_like, iiwa_like_batch_fast
from oracle import gato_oracle as orc
from v2_check import kernel_only, rel


def single(prob, name):
    S, C, K = prob["S"], prob["C"], prob["K"]
    lam_o, dz_o, io = orc.linsys_solve_info(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"],
        prob["C_col"], prob["C_val"], prob["g"], prob["c"], np.zeros(S * K), 1, prob["exit_tol"], prob["max_iters"], False, prob["rho"])
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    pl.set_kernel(2)
    for v in (0, 1):
        pl.set_pcg(v)
        lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
        med, mn, _, _ = kernel_only(pl, prob, reps=50)
        print(f"[{name}] pcg {v}: iters {inf['iters']} (oracle {io['iters']}) lam {rel(lam, lam_o):.2e} dz {rel(dz, dz_o):.2e} "
              f"| kernel-only us median {med:.1f} min {mn:.1f}", flush=True)
    pl.close()


def batch(nsys=8192):
    b = iiwa_like_batch_fast(nsys, 14, 7, 50, seed=1000)
    pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=nsys)
    res = {}
    for v in (0, 1):
        pl.set_pcg(v)
        med, mn, its, lam = kernel_only(pl, b, nsys=nsys, reps=10, batch=b)
        res[v] = (its.copy(), lam.reshape(nsys, -1).copy())
        print(f"[batch {nsys}] pcg {v}: {med / 1e3:.3f} ms median ({nsys / med * 1e6 / 1e6:.3f} M solves/s), "
              f"iters mean {its.mean():.2f} max {its.max()}", flush=True)
    d = res[1][0].astype(int) - res[0][0].astype(int)
    vals, cnt = np.unique(d, return_counts=True)
    print("iteration-count difference (pcg1 - pcg0) histogram:", dict(zip(vals.tolist(), cnt.tolist())))
    num = np.linalg.norm(res[1][1] - res[0][1], axis=1); den = np.linalg.norm(res[0][1], axis=1)
    r = num / den
    print(f"lambda rel diff pcg1 vs pcg0 over the batch: median {np.median(r):.2e} max {r.max():.2e}")
    # a sample against the oracle
    worst = np.argsort(-np.abs(d))[:4].tolist() + [0, 1]
    for i in worst:
        lam_o, dz_o, io = orc.linsys_solve_info(14, 7, 50, b["G_row"], b["G_col"], b["G_val"][i], b["C_row"], b["C_col"],
            b["C_val"][i], b["g"][i], b["c"][i], np.zeros(700), 1, b["exit_tol"], b["max_iters"], False, b["rho"])
        print(f"  system {i}: oracle iters {io['iters']}, pcg0 {res[0][0][i]} (lam {rel(res[0][1][i], lam_o):.2e}), "
              f"pcg1 {res[1][0][i]} (lam {rel(res[1][1][i], lam_o):.2e})")
    pl.close()


if __name__ == "__main__":
    single(pendulum5(), "pendulum5")
    for seed in (0, 1):
        single(iiwa_like(K=50, seed=seed), f"iiwa50 seed{seed}")
    single(iiwa_like(K=50, seed=0, rho=1e-1), "iiwa50 rho1e-1")
    single(iiwa_like(K=50, seed=0, rho=1e-4), "iiwa50 rho1e-4")
    single(iiwa_like(K=50, seed=1, dense_G=True), "iiwa50 denseG")
    single(iiwa_like(S=4, C=2, K=100, seed=1, preset="mild"), "S4K100")
    batch()
