"""Kernel 5 (scaled system, two kernels) against kernel 2 on the benchmark batch: parity of the two and device time
per 8192 resident systems, per half of kernel 5 and for a few chunk sizes.  GPU box only."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like_batch_fast  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
b = iiwa_like_batch_fast(n, 14, 7, 50, seed=1000)
dev = torch.device("cuda", 0)
pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=n)
d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
nnzG, nnzC = b["G_val"].shape[1], b["C_val"].shape[1]
st = torch.cuda.current_stream()
out = {}
res = {}


def run(lam, dz, its):
    pl.solve_batched_device(n, d["G_val"].data_ptr(), nnzG, d["C_val"].data_ptr(), nnzC, d["g"].data_ptr(), pl.n_z,
                            d["c"].data_ptr(), pl.n_l, lam.data_ptr(), dz.data_ptr(), its.data_ptr(),
                            exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=st.cuda_stream)


for kernel, chunk in ((2, 0), (5, 0), (5, 1184), (5, 8192), (51, 0), (51, 1184)):
    form = 1 if kernel == 51 else 0
    kernel = 5 if kernel == 51 else kernel
    pl.set_kernel(kernel)
    pl.set_pcg_form(form)
    if kernel == 5:
        pl.set_scaled_chunk(chunk)
    lam = torch.zeros((n, pl.n_l), device=dev); dz = torch.zeros((n, pl.n_z), device=dev)
    its = torch.zeros(n, dtype=torch.int32, device=dev)
    for _ in range(3):
        run(lam, dz, its)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record()
    for _ in range(reps):
        run(lam, dz, its)
    e1.record(); e1.synchronize()
    ms = e0.elapsed_time(e1) / reps
    key = f"k{kernel}{'_onered' if form else ''}_chunk{chunk}"
    out[key] = {"ms_per_batch": ms, "solves_per_s": n / ms * 1e3, "mean_iters": float(its.float().mean())}
    res[key] = (lam.cpu().numpy(), dz.cpu().numpy(), its.cpu().numpy())
    print(key, out[key], flush=True)

l2, z2, i2 = res["k2_chunk0"]
l5, z5, i5 = res["k5_chunk0"]
rl = np.linalg.norm(l5 - l2, axis=1) / np.linalg.norm(l2, axis=1)
rz = np.linalg.norm(z5 - z2, axis=1) / np.linalg.norm(z2, axis=1)
out["k5_vs_k2"] = {"max_rel_lambda": float(rl.max()), "mean_rel_lambda": float(rl.mean()), "max_rel_dz": float(rz.max()),
                   "iter_diff_hist": {int(k): int(v) for k, v in zip(*np.unique(i5.astype(int) - i2.astype(int), return_counts=True))},
                   "nan": bool(np.isnan(l5).any())}
for key in ("k5_chunk1184", "k5_chunk8192"):
    out[key]["bitwise_equal_to_default_chunk"] = bool(np.array_equal(res[key][0], l5) and np.array_equal(res[key][2], i5))
l6, z6, i6 = res["k5_onered_chunk0"]
rl = np.linalg.norm(l6 - l2, axis=1) / np.linalg.norm(l2, axis=1)
rz = np.linalg.norm(z6 - z2, axis=1) / np.linalg.norm(z2, axis=1)
out["k5_onered_vs_k2"] = {"max_rel_lambda": float(rl.max()), "mean_rel_lambda": float(rl.mean()), "max_rel_dz": float(rz.max()),
                          "iter_diff_hist": {int(k): int(v) for k, v in zip(*np.unique(i6.astype(int) - i2.astype(int), return_counts=True))},
                          "nan": bool(np.isnan(l6).any())}
pl.set_pcg_form(0)
print(json.dumps(out))

# single-solve latency, kernel only (device buffers), kernel 2 vs kernel 5
from gato_python_b200.problems import iiwa_like  # noqa: E402
prob = iiwa_like(14, 7, 50, seed=0)
p1 = Plan(14, 7, 50, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)  # noqa: E731
G, Cv, g1, c1 = t(prob["G_val"]), t(prob["C_val"]), t(prob["g"]), t(prob["c"])
l1 = torch.empty(p1.n_l, device=dev); z1 = torch.empty(p1.n_z, device=dev); i1 = torch.empty(1, dtype=torch.int32, device=dev)
single = {}
for kernel in (2, 5):
    p1.set_kernel(kernel)
    def run1():
        p1.solve_batched_device(1, G.data_ptr(), 0, Cv.data_ptr(), 0, g1.data_ptr(), 0, c1.data_ptr(), 0, l1.data_ptr(),
                                z1.data_ptr(), i1.data_ptr(), exit_tol=prob["exit_tol"], max_iters=prob["max_iters"],
                                rho=prob["rho"], stream=st.cuda_stream)
    for _ in range(20):
        run1()
    torch.cuda.synchronize()
    ts = []
    for _ in range(200):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run1(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    single[f"k{kernel}"] = {"kernel_us_median": float(np.median(ts)), "iters": int(i1.item())}
    if kernel == 5:
        for variant in (1, 2):
            p1.set_scaled_variant(variant)
            p1.set_profiling(True)
            hs = []
            for _ in range(50):
                run1()
                hs.append(p1.last_profile()[:2])
            p1.set_profiling(False)
            single[f"k5_variant{variant}_halves_us"] = [float(np.median([h[0] for h in hs]) * 1e3), float(np.median([h[1] for h in hs]) * 1e3)]
        p1.set_scaled_variant(0)
print(json.dumps({"single_solve": single}))
