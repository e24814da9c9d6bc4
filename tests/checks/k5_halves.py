"""Kernel 5 on an 8192-system batch of horizon K (argv[1]): ms of each half from the plan's CUDA events, iterations.
Used to compare builds of the second half (threads per CTA, CTAs per SM).  GPU box only."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like_batch_fast  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 50
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
b = iiwa_like_batch_fast(n, 14, 7, K, seed=1000)
dev = torch.device("cuda", 0)
pl = Plan(14, 7, K, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
pl.set_kernel(5)
pl.set_profiling(True)
d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
lam = torch.zeros((n, pl.n_l), device=dev); dz = torch.zeros((n, pl.n_z), device=dev)
its = torch.zeros(n, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream()
h = []
for r in range(12):
    pl.solve_batched_device(n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                            d["g"].data_ptr(), pl.n_z, d["c"].data_ptr(), pl.n_l, lam.data_ptr(), dz.data_ptr(),
                            its.data_ptr(), exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"],
                            stream=st.cuda_stream)
    torch.cuda.synchronize()
    if r >= 4:
        h.append(pl.last_profile()[:2])
h = np.array(h)
print(json.dumps({"K": K, "nsys": n, "first_half_ms": float(h[:, 0].mean()), "second_half_ms": float(h[:, 1].mean()),
                  "mean_iters": float(its.float().mean()), "lib": os.environ.get("GATO_B200_LIB", "production")}))
