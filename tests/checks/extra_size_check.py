"""Even state sizes other than 14 through the two-column-group (H = 2) register mapping of the v2 kernels: the
library is built a second time with one extra (STATE_SIZE, CONTROL_SIZE) pair into build/ (the in-tree library is
left alone), loaded in place of it, and every kernel is checked against the oracle.

    python tests/checks/extra_size_check.py build S C     # here (nvcc): build/libgato_b200_S{S}_C{C}.so
    python tests/checks/extra_size_check.py run S C       # on the GPU box
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import build as b  # noqa: E402
from gato_python_b200 import solver  # noqa: E402


def side_lib(S, C):
    return os.path.join(ROOT, "build", f"libgato_b200_S{S}_C{C}.so")


def build(S, C):
    print(b.build_library(extra_sc=(S, C), out=side_lib(S, C)))


def rel(a, c):
    return float(np.linalg.norm(a - c) / max(np.linalg.norm(c), 1e-30))


def run(S, C):
    solver.LIB_PATH = side_lib(S, C)
    from gato_python_b200 import Plan
    from gato_python_b200.problems import iiwa_like
    from oracle import gato_oracle as orc

    bad = 0
    for K, dense_G, dense_AB, kernel in [(10, False, False, 2), (40, False, False, 2), (40, True, True, 2),
                                         (40, False, False, 1), (40, False, False, 3), (33, True, False, 3),
                                         (150, False, False, 0), (150, False, False, 4)]:
        prob = iiwa_like(S=S, C=C, K=K, seed=3, dense_G=dense_G, dense_AB=dense_AB)
        lam_o, dz_o, io = orc.linsys_solve_info(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"],
                                                prob["C_col"], prob["C_val"], prob["g"], prob["c"], np.zeros(S * K),
                                                1, prob["exit_tol"], prob["max_iters"], False, prob["rho"])
        pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
        try:
            if kernel:
                pl.set_kernel(kernel)
            lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"],
                                    prob["max_iters"], 0, prob["rho"])
            ok = rel(lam, lam_o) < 1e-4 and rel(dz, dz_o) < 1e-4 and abs(inf["iters"] - io["iters"]) <= 1
            print(f"S={S} C={C} K={K} dense_G={dense_G} dense_AB={dense_AB} kernel {pl.kernel()} cluster "
                  f"{pl.cluster()}: iters {inf['iters']} (oracle {io['iters']}) lam {rel(lam, lam_o):.2e} "
                  f"dz {rel(dz, dz_o):.2e} {'OK' if ok else 'MISMATCH'}", flush=True)
            bad += not ok
        except solver.GatoError as e:
            print(f"S={S} C={C} K={K} kernel {kernel}: {e}", flush=True)
        pl.close()
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    mode, S, C = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    build(S, C) if mode == "build" else run(S, C)
