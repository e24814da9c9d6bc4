"""v2 (one CTA per system) vs v1 (cluster) vs the NumPy oracle; kernel-only timing and phase stamps."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from gato_python_b200 import Plan
from gato_python_b200.problems import pendulum5, iiwa_like, iiwa_like_batch_fast
from oracle import gato_oracle as orc

def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))

def kernel_only(pl, prob, nsys=1, reps=100, batch=None):
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev)
    if batch is None:
        G, Cv, g, c = t(prob["G_val"]), t(prob["C_val"]), t(prob["g"]), t(prob["c"]); ld = (0, 0, 0, 0)
    else:
        G, Cv, g, c = t(batch["G_val"]), t(batch["C_val"]), t(batch["g"]), t(batch["c"])
        ld = (G.shape[1], Cv.shape[1], g.shape[1], c.shape[1])
    lam = torch.empty(nsys * pl.n_l, device=dev); dz = torch.empty(nsys * pl.n_z, device=dev)
    it = torch.empty(nsys, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    def run():
        pl.solve_batched_device(nsys, G.data_ptr(), ld[0], Cv.data_ptr(), ld[1], g.data_ptr(), ld[2], c.data_ptr(), ld[3],
                                lam.data_ptr(), dz.data_ptr(), it.data_ptr(), exit_tol=prob["exit_tol"],
                                max_iters=prob["max_iters"], rho=prob["rho"], stream=st)
    for _ in range(5): run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts)), float(np.min(ts)), it.cpu().numpy(), lam.cpu().numpy()

NAMES = ["clear", "unpack", "invert Q,R", "schur rows", "mats->regs", "(-)", "pcg init", "pcg loop", "dz"]
def check(prob, name):
    S, C, K = prob["S"], prob["C"], prob["K"]
    lam_o, dz_o, info = orc.linsys_solve_info(S, C, K, prob["G_row"], prob["G_col"], prob["G_val"], prob["C_row"],
        prob["C_col"], prob["C_val"], prob["g"], prob["c"], np.zeros(S*K), 1, prob["exit_tol"], prob["max_iters"], False, prob["rho"])
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"])
    for kern in (2, 1):
        try:
            pl.set_kernel(kern)
        except Exception as e:
            print(f"[{name}] kernel {kern}: {e}"); continue
        lam, dz, inf = pl.solve(prob["G_val"], prob["C_val"], prob["g"], prob["c"], None, prob["exit_tol"], prob["max_iters"], 0, prob["rho"])
        med, mn, its, _ = kernel_only(pl, prob)
        d = pl.solve_debug(prob["G_val"], prob["C_val"], prob["g"], prob["c"], prob["exit_tol"], prob["max_iters"], prob["rho"], dumps=False)
        dt = np.diff(d["clk_ns"][:10])
        print(f"[{name}] kernel v{pl.kernel()} cl={pl.cluster()} iters={inf['iters']} (oracle {info['iters']}) lam {rel(lam, lam_o):.2e} dz {rel(dz, dz_o):.2e} "
              f"| kernel-only us median {med:.1f} min {mn:.1f} | per iter {dt[7]/1e3/(inf['iters']+1):.3f} us")
        print("     " + "  ".join(f"{n}={v/1e3:.1f}" for n, v in zip(NAMES, dt)))
    pl.close()

if __name__ == "__main__":
    check(pendulum5(), "pendulum5")
    check(iiwa_like(S=4, C=2, K=1, seed=0), "K1")
    check(iiwa_like(S=4, C=2, K=2, seed=0), "K2")
    check(iiwa_like(S=7, C=3, K=33, seed=2, preset="mild"), "S7K33")
    check(iiwa_like(K=50, seed=0), "iiwa50")
    check(iiwa_like(K=50, seed=4, dense_G=True, dense_AB=True), "iiwa50-dense")
    check(iiwa_like(K=50, seed=2, preset="mild", rho=1e-1), "iiwa50-mild")
    check(iiwa_like(K=12, seed=3), "iiwa12")
    # batched throughput
    nsys = 4096
    b = iiwa_like_batch_fast(nsys)
    prob = dict(b)
    pl = Plan(14, 7, 50, b["G_row"], b["G_col"], b["C_row"], b["C_col"], max_batch=nsys)
    outs = {}
    for kern in (2, 1):
        pl.set_kernel(kern)
        med, mn, its, lam = kernel_only(pl, prob, nsys=nsys, reps=5, batch=b)
        outs[kern] = (its, lam)
        print(f"[batch {nsys}] kernel v{kern}: {med/1e3:.3f} ms -> {nsys/(med*1e-6):.0f} solves/s, mean iters {its.mean():.2f}")
    print("v2 vs v1 batch: iters max diff", np.abs(outs[2][0] - outs[1][0]).max(), "lam rel", rel(outs[2][1], outs[1][1]))
    pl.close()
