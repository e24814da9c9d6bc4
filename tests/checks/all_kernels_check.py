"""One pass over every kernel (1, 2 direct, 2 persistent with TMA staging, 3, 4) on small inputs, each result
checked against the oracle; one line per case.  Written to run under compute-sanitizer (memcheck / racecheck /
synccheck), which this GPU pool keeps closed -- so it runs plain, as the all-kernels correctness pass.

    python tests/checks/all_kernels_check.py [k2]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like, iiwa_like_batch_fast, pendulum5  # noqa: E402
from oracle import gato_oracle as orc  # noqa: E402


def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def single(prob, name, kernel=0, cluster=0, warm=0):
    S, C, K = prob["S"], prob["C"], prob["K"]
    lam_o, dz_o, io = orc.linsys_solve_info(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"],
                                            prob["C_col"], prob["C_val"], prob["g"], prob["c"], np.zeros(S * K), 1,
                                            prob["exit_tol"], prob["max_iters"], False, prob["rho"])
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    if kernel:
        pl.set_kernel(kernel)
    if cluster:
        pl.set_cluster(cluster)
    lam0 = lam_o * 0.5 if warm else None
    lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], lam0, prob["exit_tol"],
                            prob["max_iters"], warm, prob["rho"])
    ok = rel(lam, lam_o) < 1e-4 and rel(dz, dz_o) < 1e-4 and (warm or abs(inf["iters"] - io["iters"]) <= 1)
    print(f"[{name}] kernel {pl.kernel()} cluster {pl.cluster()} warm {warm}: iters {inf['iters']} "
          f"(oracle cold {io['iters']}) lam {rel(lam, lam_o):.2e} dz {rel(dz, dz_o):.2e} {'OK' if ok else 'MISMATCH'}",
          flush=True)
    pl.close()
    return ok


def batched(nsys, name, staging):
    b = iiwa_like_batch_fast(nsys, 14, 7, 50, seed=7)
    pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=nsys)
    pl.set_staging(staging)
    lam, dz, its, _ = pl.solve_batched(b["G_val"], b["C_val"], b["g"], b["c"], None, b["exit_tol"], b["max_iters"], 0,
                                       b["rho"])
    i = nsys - 1
    lam_o, dz_o, io = orc.linsys_solve_info(14, 7, 50, b["G_row"], b["G_col"], b["G_val"][i], b["C_row"], b["C_col"],
                                            b["C_val"][i], b["g"][i], b["c"][i], np.zeros(700), 1, b["exit_tol"],
                                            b["max_iters"], False, b["rho"])
    ok = rel(lam[i], lam_o) < 1e-4 and rel(dz[i], dz_o) < 1e-4 and abs(int(its[i]) - io["iters"]) <= 1
    print(f"[{name}] {nsys} systems staging {staging}: last system iters {int(its[i])} (oracle {io['iters']}) "
          f"lam {rel(lam[i], lam_o):.2e} dz {rel(dz[i], dz_o):.2e} {'OK' if ok else 'MISMATCH'}", flush=True)
    pl.close()
    return ok


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    ok = True
    ok &= single(pendulum5(), "pendulum5")
    ok &= single(iiwa_like(K=50, seed=0), "iiwa50 k2")
    ok &= single(iiwa_like(K=50, seed=0), "iiwa50 k2 true-warm", warm=2)
    ok &= single(iiwa_like(K=50, seed=1, dense_G=True), "iiwa50 dense k2")
    ok &= batched(300, "iiwa50 batch persistent", 1)     # > 148 CTAs: persistent loop + TMA prefetch of the next system
    ok &= batched(20, "iiwa50 batch unstaged", 0)
    if which != "k2":
        ok &= single(iiwa_like(K=50, seed=0), "iiwa50 k1", kernel=1)
        ok &= single(iiwa_like(K=50, seed=0), "iiwa50 k3 x2", kernel=3, cluster=2)
        ok &= single(iiwa_like(K=120, seed=0), "iiwa120 k3", kernel=3)
        ok &= single(iiwa_like(K=256, seed=0), "iiwa256 k3 (9 CTAs)")
        ok &= single(iiwa_like(K=120, seed=0), "iiwa120 k4", kernel=4)
        ok &= single(iiwa_like(S=4, C=2, K=300, seed=1, preset="mild"), "S4K300 k1", kernel=1)
    print("ALL OK" if ok else "SOME MISMATCH", flush=True)
    sys.exit(0 if ok else 1)
