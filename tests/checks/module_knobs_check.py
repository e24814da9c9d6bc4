"""The drop-in module built with build knobs outside the built-in size list (CMakeLists.txt:18 of the reference:
-DSTATE_SIZE -DCONTROL_SIZE -DKNOT_POINTS): `python -m gato_python_b200.build --module S C K` here, then on the GPU box

    python tests/checks/module_knobs_check.py S C K

calls gpu_library.linsys_solve (lists in, lists out, the reference's signature) and compares with the oracle.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200.build import module_dir  # noqa: E402
from gato_python_b200.problems import iiwa_like  # noqa: E402
from oracle import gato_oracle as orc  # noqa: E402

S, C, K = (int(a) for a in sys.argv[1:4])
sys.path.insert(0, module_dir(S, C, K))
import gpu_library  # noqa: E402

bad = 0
for seed, dense in [(0, False), (1, True)]:
    p = iiwa_like(S=S, C=C, K=K, seed=seed, dense_G=dense, dense_AB=dense)
    lam_o, dz_o, io = orc.linsys_solve_info(S, C, K, p["G_row"], p["G_col"], p["G_val"], p["C_row"], p["C_col"],
                                            p["C_val"], p["g"], p["c"], np.zeros(S * K), 1, p["exit_tol"],
                                            p["max_iters"], False, p["rho"])
    lam, dz = gpu_library.linsys_solve(p["G_row"].tolist(), p["G_col"].tolist(), p["G_val"].tolist(),
                                       p["C_row"].tolist(), p["C_col"].tolist(), p["C_val"].tolist(), p["g"].tolist(),
                                       p["c"].tolist(), [0.0] * (S * K), 1, p["exit_tol"], p["max_iters"], False,
                                       p["rho"])
    lam, dz = np.asarray(lam), np.asarray(dz)
    el = np.linalg.norm(lam - lam_o) / np.linalg.norm(lam_o)
    ed = np.linalg.norm(dz - dz_o) / np.linalg.norm(dz_o)
    ok = el < 1e-4 and ed < 1e-4
    bad += not ok
    print(f"gpu_library S={S} C={C} K={K} seed {seed} dense {dense}: lam {el:.2e} dz {ed:.2e} "
          f"(oracle iters {io['iters']}) {'OK' if ok else 'MISMATCH'}", flush=True)
sys.exit(1 if bad else 0)
