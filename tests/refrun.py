"""Run a `gpu_library` module (the reference build under oracle/_ref or the drop-in under build/)
in a subprocess -- the module name and its sizes are fixed at compile time, so each configuration
needs its own interpreter -- and capture the C-level stdout the reference prints its iteration
count to (gpu_library.cu:189-198)."""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_CHILD = r"""
import sys, json, numpy as np
sys.path.insert(0, sys.argv[1])
import gpu_library
d = np.load(sys.argv[2])
out = {}
args = [d[k].tolist() for k in ("G_row","G_col","G_val","C_row","C_col","C_val","g","c","lambda0")]
l, dz = gpu_library.linsys_solve(*args, int(d["testiters"]), float(d["exit_tol"]), int(d["max_iters"]),
                                 bool(d["warm_start"]), float(d["rho"]))
wall = 0.0
nrep = int(sys.argv[4]) if len(sys.argv) > 4 else 0
if nrep:                      # Python wall clock of the list-in / list-out call itself (SURVEY 8d "T_e2e")
    import time, os
    fd = os.dup(1); devnull = os.open(os.devnull, os.O_WRONLY); os.dup2(devnull, 1)   # mute the C-level prints
    a2 = args[:9] + [1] + [float(d["exit_tol"]), int(d["max_iters"]), bool(d["warm_start"]), float(d["rho"])]
    for _ in range(5):
        gpu_library.linsys_solve(*a2)
    t0 = time.perf_counter()
    for _ in range(nrep):
        gpu_library.linsys_solve(*a2)
    wall = (time.perf_counter() - t0) / nrep
    os.dup2(fd, 1)
np.savez(sys.argv[3], lam=np.array(l, np.float32), dz=np.array(dz, np.float32), wall=wall)
"""


def ref_module_dir(S, C, K):
    d = os.path.join(ROOT, "oracle", "_ref", f"S{S}_C{C}_K{K}")
    return d if os.path.isdir(d) and any(f.startswith("gpu_library") for f in os.listdir(d)) else None


def dropin_module_dir(S, C, K):
    d = os.path.join(ROOT, "build", f"S{S}_C{C}_K{K}")
    return d if os.path.isdir(d) and any(f.startswith("gpu_library") for f in os.listdir(d)) else None


def run_module(mod_dir, prob, lambda0=None, testiters=1, warm_start=False, timeout=300, wall_reps=0):
    """-> dict(lam, dz, iters, first_ms, avg_ms, stdout, python_wall_s)"""
    n_l = prob["S"] * prob["K"]
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "in.npz"), os.path.join(td, "out.npz")
        np.savez(fin, G_row=prob["G_row"], G_col=prob["G_col"], G_val=prob["G_val"], C_row=prob["C_row"],
                 C_col=prob["C_col"], C_val=prob["C_val"], g=prob["g"], c=prob["c"],
                 lambda0=np.zeros(n_l, np.float32) if lambda0 is None else np.asarray(lambda0, np.float32),
                 testiters=testiters, exit_tol=prob["exit_tol"], max_iters=prob["max_iters"],
                 warm_start=warm_start, rho=prob["rho"])
        r = subprocess.run([sys.executable, "-c", _CHILD, mod_dir, fin, fout, str(int(wall_reps))], capture_output=True, text=True,
                           timeout=timeout)
        if r.returncode != 0:
            raise RuntimeError(f"gpu_library run failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
        o = np.load(fout)
        m = re.search(r"first run PCG terminated in (\d+) iterations, time:\s+([\d.eE+-]+)", r.stdout)
        a = re.search(r"avg time: ([\d.eE+-]+)", r.stdout)
        return dict(lam=o["lam"], dz=o["dz"], iters=int(m.group(1)) if m else None,
                    first_ms=float(m.group(2)) if m else None, avg_ms=float(a.group(1)) if a else None,
                    stdout=r.stdout, python_wall_s=float(o["wall"]))
