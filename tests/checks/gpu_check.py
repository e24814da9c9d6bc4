"""First-light check on a B200: CUDA path vs the NumPy oracle (and the reference build if present)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gato_python_b200 import Plan
from gato_python_b200.problems import pendulum5, iiwa_like
from oracle import gato_oracle as orc
import refrun

def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)), float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))

def check(prob, name, clusters):
    S, C, K = prob["S"], prob["C"], prob["K"]
    lam_o, dz_o, info = orc.linsys_solve_info(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"],
        prob["C_col"], prob["C_val"], prob["g"], prob["c"], np.zeros(S*K), 1, prob["exit_tol"], prob["max_iters"], False, prob["rho"])
    lam64, dz64 = orc.dense_kkt_solve(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"], prob["C_col"],
                                      prob["C_val"], prob["g"], prob["c"], prob["rho"])
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    Gd, Cd = pl.unpack(prob["G_val"], prob["C_val"], prob["rho"])
    print(f"[{name}] unpack bit-exact: G {np.array_equal(Gd.view(np.uint32), info['G_dense'].view(np.uint32))} "
          f"C {np.array_equal(Cd.view(np.uint32), info['C_dense'].view(np.uint32))}")
    for cl in clusters:
        try:
            pl.set_cluster(cl)
        except Exception as e:
            print(f"[{name}] cl={cl}: {e}"); continue
        lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
        ts = []
        for _ in range(20):
            _, _, i2 = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
            ts.append(i2["device_ms"])
        print(f"[{name}] cl={pl.cluster()} iters={inf['iters']} (oracle {info['iters']}) lam vs oracle {rel(lam, lam_o)} dz {rel(dz, dz_o)} | "
              f"vs fp64 KKT lam {rel(lam, lam64)} dz {rel(dz, dz64)} | bracket ms median {np.median(ts):.4f} min {np.min(ts):.4f}")
    rd = refrun.ref_module_dir(S, C, K)
    if rd:
        try:
            r = refrun.run_module(rd, prob, testiters=20)
            print(f"[{name}] REFERENCE iters={r['iters']} lam vs ours {rel(lam, r['lam'])} dz {rel(dz, r['dz'])} "
                  f"ref-vs-oracle lam {rel(r['lam'], lam_o)} | first_ms {r['first_ms']} avg_ms {r['avg_ms']}")
        except Exception as e:
            print(f"[{name}] reference run failed: {e}")
    dd = refrun.dropin_module_dir(S, C, K)
    if dd:
        r = refrun.run_module(dd, prob, testiters=20)
        print(f"[{name}] DROP-IN module iters={r['iters']} lam vs ours {rel(lam, r['lam'])} first_ms {r['first_ms']} avg_ms {r['avg_ms']}")
    pl.close()

if __name__ == "__main__":
    check(pendulum5(), "pendulum5", [0])
    check(iiwa_like(K=50, seed=0), "iiwa50", [2, 4, 8, 16])
    check(iiwa_like(K=50, seed=1, dense_G=True, dense_AB=True), "iiwa50-dense", [2])
    check(iiwa_like(K=256, seed=0), "iiwa256", [16])
