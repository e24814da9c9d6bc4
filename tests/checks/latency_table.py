"""Single-solve latency table (BASELINE.json configs 0-2): ours (kernel-only and the reference's
H2D + solve + D2H bracket) beside the unmodified reference CUDA build's own `avg time` on the same
inputs, when its prebuilt module travelled to the box (oracle/_ref).  Prints one JSON object."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import refrun  # noqa: E402
from gato_python_b200 import Plan  # noqa: E402
from gato_python_b200.problems import iiwa_like, pendulum5  # noqa: E402


def ours(prob, kernel=0, reps=200):
    S, C, K = prob["S"], prob["C"], prob["K"]
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    if kernel:
        pl.set_kernel(kernel)
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)   # noqa: E731
    G, Cv, g, c = t(prob["G_val"]), t(prob["C_val"]), t(prob["g"]), t(prob["c"])
    lam = torch.empty(pl.n_l, device=dev); dz = torch.empty(pl.n_z, device=dev)
    it = torch.empty(1, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def run():
        pl.solve_batched_device(1, G.data_ptr(), 0, Cv.data_ptr(), 0, g.data_ptr(), 0, c.data_ptr(), 0,
                                lam.data_ptr(), dz.data_ptr(), it.data_ptr(), exit_tol=prob["exit_tol"],
                                max_iters=prob["max_iters"], rho=prob["rho"], stream=st)
    for _ in range(20):
        run()
    torch.cuda.synchronize()
    ks = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); e1.synchronize()
        ks.append(e0.elapsed_time(e1) * 1e3)
    bs = []
    for _ in range(reps):
        _, _, info = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"],
                              prob["max_iters"], 0, prob["rho"])
        bs.append(info["device_ms"] * 1e3)
    out = {"kernel": pl.kernel(), "cluster": list(pl.cluster()), "iters": int(it.item()),
           "kernel_us_median": round(float(np.median(ks)), 2), "kernel_us_min": round(float(np.min(ks)), 2),
           "bracket_us_median": round(float(np.median(bs)), 2)}
    out["us_per_iter"] = round(out["kernel_us_median"] / (out["iters"] + 1), 3)
    pl.close()
    return out


def main():
    cases = {"pendulum5": pendulum5(), "iiwa50_s0": iiwa_like(K=50, seed=0), "iiwa256_s0": iiwa_like(K=256, seed=0)}
    table = {}
    for name, prob in cases.items():
        row = {"ours": ours(prob)}
        if prob["K"] == 50:
            row["ours_cluster_kernel"] = ours(prob, kernel=1)
        md = refrun.ref_module_dir(prob["S"], prob["C"], prob["K"])
        if md:
            r = refrun.run_module(md, prob, testiters=50, wall_reps=100)
            row["reference_cuda_build"] = {"iters": r["iters"], "bracket_us_avg50": round(r["avg_ms"] * 1e3, 1),
                                           "first_us": round(r["first_ms"] * 1e3, 1),
                                           "python_list_call_wall_us": round(r["python_wall_s"] * 1e6, 1)}
        dd = refrun.dropin_module_dir(prob["S"], prob["C"], prob["K"])
        if dd:   # the same list-in / list-out Python call through OUR gpu_library module
            r = refrun.run_module(dd, prob, testiters=50, wall_reps=100)
            row["ours_dropin_module"] = {"iters": r["iters"], "bracket_us_avg50": round(r["avg_ms"] * 1e3, 1),
                                         "python_list_call_wall_us": round(r["python_wall_s"] * 1e6, 1)}
        table[name] = row
        print(name, json.dumps(row), file=sys.stderr, flush=True)
    print(json.dumps(table))


if __name__ == "__main__":
    main()
