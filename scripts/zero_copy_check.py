"""gato_solve bracket (H2D + kernel + D2H) with and without zero-copy inputs; bitwise equality."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan
from gato_python_b200.problems import iiwa_like, pendulum5
for name, prob in (("iiwa50", iiwa_like(K=50, seed=0)), ("pendulum5", pendulum5()), ("iiwa12", iiwa_like(K=12, seed=3))):
    pl = Plan(prob["S"], prob["C"], prob["K"], prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    res = {}
    for zc in (1, 0, 1, 0):
        pl.set_zero_copy(zc)
        ts = []
        for _ in range(300):
            lam, dz, info = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"],
                                     prob["max_iters"], 0, prob["rho"])
            ts.append(info["device_ms"] * 1e3)
        res[zc] = (lam, dz)
        print(f"[{name}] zero_copy={zc}: bracket us median {np.median(ts[50:]):.1f} p99 {np.percentile(ts[50:], 99):.1f} iters {info['iters']}")
    print("  bitwise equal:", np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]))
    pl.close()
