"""`gato_solve_batched_device_noalloc` (no counterpart in the reference):
kernel 5 with its scratch in a file-scope device array, for callers that cannot call cudaMalloc.

Through the C ABI (ctypes), against the existing entry point `gato_solve_batched_device` with kernel 5 selected, on two
calls in a row (the array keeps the first call's contents: the second call must not depend on them), with different
inputs and a different system count.  Collected last (`zz`).
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like_batch_fast  # noqa: E402

pytestmark = pytest.mark.gpu
TOL = 1e-4          # the bar of the parity tests (lambda, dz relative; iterations +-1); the two paths run the same kernels


def _rel(a, ref):
    a = a.astype(np.float64); ref = ref.astype(np.float64)
    return max(np.linalg.norm(a - ref) / max(np.linalg.norm(ref), 1e-300), np.abs(a - ref).max() / max(np.abs(ref).max(), 1e-300))


def test_noalloc_entry_point_matches_the_allocating_one_twice_in_a_row():
    import torch

    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream().cuda_stream
    S, C, K = 14, 7, 50
    b0 = iiwa_like_batch_fast(4, S, C, K, seed=77)
    pl = Plan(S, C, K, b0["G_row"], b0["G_col"], b0["C_row"], b0["C_col"], max_batch=64)
    cap = pl.noalloc_max_systems()
    assert cap == (8 * 2**20 // 4 - 1024) // (K * 1088) == 38
    ref = Plan(S, C, K, b0["G_row"], b0["G_col"], b0["C_row"], b0["C_col"], max_batch=64)
    ref.set_kernel(5)
    for call, (n, seed) in enumerate([(cap, 501), (7, 502)]):
        b = iiwa_like_batch_fast(n, S, C, K, seed=seed)
        d = {k: torch.from_numpy(b[k]).to(dev) for k in ("G_val", "C_val", "g", "c")}
        out = {}
        for name in ("noalloc", "ref"):
            lam = torch.full((n, pl.n_l), float("nan"), device=dev)
            dz = torch.full((n, pl.n_z), float("nan"), device=dev)
            its = torch.full((n,), -7, dtype=torch.int32, device=dev)
            args = (n, d["G_val"].data_ptr(), b["G_val"].shape[1], d["C_val"].data_ptr(), b["C_val"].shape[1],
                    d["g"].data_ptr(), pl.n_z, d["c"].data_ptr(), pl.n_l, lam.data_ptr(), dz.data_ptr(), its.data_ptr())
            if name == "noalloc":
                pl.solve_batched_device_noalloc(*args, exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=st)
            else:
                ref.solve_batched_device(*args, exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=st)
            torch.cuda.synchronize()
            out[name] = (lam.cpu().numpy(), dz.cpu().numpy(), its.cpu().numpy())
        (l1, z1, i1), (l2, z2, i2) = out["noalloc"], out["ref"]
        assert np.isfinite(l1).all() and np.isfinite(z1).all() and (i1 >= 0).all(), call
        assert np.abs(i1.astype(int) - i2.astype(int)).max() <= 1, call
        for i in range(n):
            assert _rel(l1[i], l2[i]) <= TOL and _rel(z1[i], z2[i]) <= TOL, (call, i, _rel(l1[i], l2[i]), _rel(z1[i], z2[i]))
    # beyond the fixed array: a clear error, nothing launched
    with pytest.raises(ValueError, match="at most 38 systems"):   # GATO_EINVAL
        pl.solve_batched_device_noalloc(cap + 1, 1, 0, 1, 0, 1, 0, 1, 0, 1, 1, 1)
    ref.close()
    pl.close()
