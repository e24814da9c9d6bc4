"""Per-phase %globaltimer breakdown of one fused solve + kernel-only CUDA-event time."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import pendulum5, iiwa_like

NAMES = ["clear smem", "unpack", "invert Q,R", "schur rows(+B1)", "stair(+B2)", "rows->regs(+B3)", "pcg init", "pcg loop", "dz"]

def kernel_only(pl, prob, nsys=1, reps=200):
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)
    G, Cv, g, c = t(prob["G_val"]), t(prob["C_val"]), t(prob["g"]), t(prob["c"])
    lam = torch.empty(nsys * pl.n_l, device=dev); dz = torch.empty(nsys * pl.n_z, device=dev)
    it = torch.empty(nsys, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    def run():
        pl.solve_batched_device(nsys, G.data_ptr(), 0, Cv.data_ptr(), 0, g.data_ptr(), 0, c.data_ptr(), 0,
                                lam.data_ptr(), dz.data_ptr(), it.data_ptr(), exit_tol=prob["exit_tol"],
                                max_iters=prob["max_iters"], rho=prob["rho"], stream=st)
    for _ in range(10): run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts)), float(np.min(ts)), int(it[0])

def prof(prob, name, cls):
    pl = Plan(prob["S"], prob["C"], prob["K"], prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    for cl in cls:
        pl.set_cluster(cl)
        best = None
        for _ in range(5):
            d = pl.solve_debug(prob["G_val"], prob["C_val"], prob["g"], prob["c"], prob["exit_tol"], prob["max_iters"], prob["rho"])
            clk = d["clk_ns"][:10]; dt = np.diff(clk)
            if best is None or dt.sum() < best.sum(): best = dt
        med, mn, iters = kernel_only(pl, prob)
        print(f"[{name}] cluster={pl.cluster()} iters={d['iters']} kernel-only us: median {med:.1f} min {mn:.1f}; "
              f"in-kernel total {best.sum()/1e3:.1f} us; per PCG iter {best[7]/1e3/(d['iters']+1):.3f} us")
        print("    " + "  ".join(f"{n}={v/1e3:.1f}" for n, v in zip(NAMES, best)))
    pl.close()

if __name__ == "__main__":
    prof(pendulum5(), "pendulum5", [0])
    prof(iiwa_like(K=50, seed=0), "iiwa50", [2, 4, 8, 16])
    prof(iiwa_like(K=256, seed=0), "iiwa256", [16])
