#!/usr/bin/env python
"""bench.py -- batched IIWA-50 KKT solves/sec (BASELINE.json config 4) and single-solve latency.

A "step" is one pass of the hot path (CSR unpack -> Schur -> PCG -> dz, one fused kernel launch)
over one batch of `--nsys` independent synthetic IIWA-50 KKT systems per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N > 1 is launched by torchrun (one process per GPU); systems are sharded by rank with no
collective on the solve path; the only torch.distributed calls are the barrier and the
max-over-ranks of the device time.  Rank 0 prints ONE JSON line.

* value        solves/s, inputs resident in HBM (CUDA events around K launches).
* e2e          solves/s through the host-buffer C-ABI call `gato_solve_batched` (pinned host
               buffers, H2D and D2H inside the timed region).
* roofline     algorithmic HBM bytes per launch / measured kernel time vs MEASURED_PEAKS.json,
               plus the FP32 figure (the kernel is FP32-pipe/latency bound, not HBM bound).
* cpu_baseline the NumPy oracle port timed on the host cores (bounded sample).
* single_solve IIWA-50 single-solve latency in us (the other half of BASELINE.json's metric).

`--impl reference` times the UNMODIFIED reference CUDA build (oracle/_ref, its own
`gpu_library.linsys_solve` called in a loop -- the reference has no batched entry point and no
CPU path) on a bounded sample of the same workload; if that build is absent it times the NumPy
oracle port on the host cores instead.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

S, C, K = 14, 7, 50
METRIC = "iiwa50_batched_kkt_solves_per_sec"
UNIT = "solves/s"
SEED = 1000
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.4, SURVEY.md section 6 (theoretical FFMA peak)


def flops_per_system(iters):
    """SURVEY.md 8(d) algorithmic flop count for one system that ran `iters` PCG iterations."""
    schur = K * (12 * S ** 3 + 2 * C ** 3 + 2 * S * C ** 2 + 2 * S ** 2 * C + 6 * S ** 2 + 2 * S * C)
    pcg = (np.asarray(iters, np.float64) + 1) * (12 * S * S * K + 10 * S * K)
    dz = 2 * K * (2 * S * S + C * C + S * C)
    return schur + pcg + dz


def bytes_per_system(nnzG, nnzC):
    """values + g + c in, lambda + dz out (indices are shared by the batch)."""
    n_z, n_l = (S + C) * K - C, S * K
    return 4 * (nnzG + nnzC + n_z + n_l) + 4 * (n_l + n_z)


def ncu_traffic(kernel, nsys):
    """dram bytes per launch of `kernel` from the committed ncu capture (profiles/ncu_traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
            e = json.load(f).get(kernel)
        return int(e["bytes"]) if e and int(e["systems_per_launch"]) == nsys else None
    except Exception:
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (NVML; falls back to nvidia-smi)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index, period=0.02):
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
        except Exception:
            self._nvml = None

    def _poll_nvml(self):
        n = self._nvml
        while not self._stop.is_set():
            try:
                self.samples.append(float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
                bits = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for b, name in self.REASONS.items():
                    if bits & b and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def _poll_smi(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for nm, v in zip(names, out[2:]):
                    if v.strip().lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            self._stop.wait(max(self.period, 0.1))

    def start(self):
        self._thr = threading.Thread(target=self._poll_nvml if self._nvml else self._poll_smi, daemon=True)
        self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=10)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# CPU baseline: the NumPy oracle port on the host cores (bounded sample of the same workload)
# ------------------------------------------------------------------------------------------------
def _cpu_worker(job):
    first, n = job
    os.environ["OMP_NUM_THREADS"] = "1"
    from gato_python_b200.problems import iiwa_like_batch_fast
    from oracle import gato_oracle as orc

    b = iiwa_like_batch_fast(n, S, C, K, seed=SEED, first=first)
    lam0 = np.zeros(S * K, np.float32)
    t0 = time.perf_counter()
    its = []
    for i in range(n):
        _, _, info = orc.linsys_solve_info(S, C, K, b["G_row"], b["G_col"], b["G_val"][i], b["C_row"], b["C_col"],
                                           b["C_val"][i], b["g"][i], b["c"][i], lam0, 1, b["exit_tol"],
                                           b["max_iters"], False, b["rho"])
        its.append(info["iters"])
    return time.perf_counter() - t0, its


def cpu_baseline(per_core=96, cores=None, scipy_per_core=0):
    """Oracle port (NumPy Schur + PCG, FP32) over `cores` processes, `per_core` systems each."""
    import multiprocessing as mp

    cores = cores or os.cpu_count() or 1
    ctx = mp.get_context("spawn")
    jobs = [(i * per_core, per_core) for i in range(cores)]
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0          # includes process start-up: report the slowest worker
    busy = max(r[0] for r in res)
    n = per_core * cores
    out = {"value": n / busy, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"{n} IIWA-50 systems of the same generator ({per_core} per process, {cores} processes, "
                     f"NumPy oracle Schur+PCG FP32; slowest worker {busy:.1f} s, wall {wall:.1f} s)",
           "mean_iters": float(np.mean([i for r in res for i in r[1]]))}
    if scipy_per_core > 0:   # C1 (SciPy CG on the sparse Schur complement, stair LinearOperator) and C2 (spsolve)
        try:
            from oracle import scipy_baseline as sb
            out["scipy_schur_cg"] = sb.batched("schur_cg", scipy_per_core, cores, seed=SEED)
            out["scipy_kkt_spsolve_fp64"] = sb.batched("kkt_spsolve", 4 * scipy_per_core, cores, seed=SEED)
        except Exception as e:   # SciPy missing on the box: the NumPy port above is the baseline
            out["scipy_error"] = repr(e)[:200]
    return out


# ------------------------------------------------------------------------------------------------
# reference arm
# ------------------------------------------------------------------------------------------------
_REF_CHILD = r"""
import sys, time, json, os
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import gpu_library
from gato_python_b200.problems import iiwa_like_batch_fast
n, steps, warmup, seed, first = (int(x) for x in sys.argv[3:8])
b = iiwa_like_batch_fast(n, 14, 7, 50, seed=seed, first=first)
Gr, Gc, Cr, Cc = b["G_row"].tolist(), b["G_col"].tolist(), b["C_row"].tolist(), b["C_col"].tolist()
lam0 = [0.0] * 700
tol, mit, rho = float(b["exit_tol"]), int(b["max_iters"]), float(b["rho"])
# The reference's API takes Python sequences.  The NumPy -> list conversion of every system's arrays is done ONCE here,
# outside the timed loop (our arm pays no conversion either); its cost is reported on its own.
t0 = time.perf_counter()
vals = [(b["G_val"][i].tolist(), b["C_val"][i].tolist(), b["g"][i].tolist(), b["c"][i].tolist()) for i in range(n)]
conv_us = (time.perf_counter() - t0) / n * 1e6
def step():
    for Gv, Cv, gv, cv in vals:
        gpu_library.linsys_solve(Gr, Gc, Gv, Cr, Cc, Cv, gv, cv, lam0, 1, tol, mit, False, rho)
for _ in range(warmup): step()
t0 = time.perf_counter()
for _ in range(steps): step()
wall = time.perf_counter() - t0
# single-solve latency as the reference itself reports it (its own cudaEvent bracket, avg of 50)
gpu_library.linsys_solve(Gr, Gc, vals[0][0], Cr, Cc, vals[0][1], vals[0][2], vals[0][3], lam0, 50, tol, mit, False, rho)
sys.stdout.flush()
sys.stderr.write("GATO_REF_RESULT " + json.dumps({"wall": wall, "n": n, "steps": steps, "list_conversion_us": conv_us}) + "\n")
"""


def bench_config(args, world):
    """The workload both arms are quoted on (identical dict in both lines)."""
    return {"workload": f"iiwa50_batch{args.nsys}_per_gpu", "S": S, "C": C, "K": K, "systems_per_gpu": args.nsys,
            "global_batch": args.nsys * world, "per_system_matrices": True, "exit_tol": 1e-6, "rho": 1e-3,
            "max_iters": 500, "parallelism": f"batch-shard x{world} (no collective)"}


def reference_kernel_only():
    """the reference's five kernels on pre-allocated buffers through oracle/ref_harness.cu
    (the reference's src/*.cuh included unchanged), IIWA-50 seed 0.  None when the harness did not travel."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import refharness
        from gato_python_b200.problems import iiwa_like
        if refharness.harness_path(S, C, K) is None:
            return None
        return refharness.kernel_only_summary(S, C, K, iiwa_like(S, C, K, seed=0), reps=300, warmup=20)
    except Exception as e:
        return {"error": repr(e)[:200]}


def reference_arm(args, rank):
    """The UNMODIFIED reference CUDA build through its own gpu_library.linsys_solve, looped over a bounded sample of the
    workload: ONE PROCESS PER GPU ("multi-GPU = N independent processes"), rates summed."""
    if rank != 0:
        return
    import re

    ref_dir = os.path.join(ROOT, "oracle", "_ref", f"S{S}_C{C}_K{K}")
    n = args.ref_sample
    ngpu = max(args.gpus, 1)
    base = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": bench_config(args, ngpu)}
    have = os.path.isdir(ref_dir) and any(f.startswith("gpu_library") for f in os.listdir(ref_dir))
    if have:
        procs = []
        for gi in range(ngpu):
            env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(gi))
            for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
                env.pop(k, None)
            procs.append(subprocess.Popen([sys.executable, "-c", _REF_CHILD, ref_dir, ROOT, str(n), str(args.steps),
                                           str(args.warmup), str(SEED), str(gi * args.nsys)], stdout=subprocess.PIPE,
                                          stderr=subprocess.PIPE, text=True, env=env))
        outs = [p.communicate(timeout=1500) + (p.returncode,) for p in procs]
        res = []
        for so, se, rc in outs:
            m = re.search(r"GATO_REF_RESULT (\{.*\})", se)
            if rc == 0 and m:
                res.append((json.loads(m.group(1)), so))
        if len(res) == ngpu:
            rates = [r["n"] * r["steps"] / r["wall"] for r, _ in res]
            val = float(sum(rates))
            avg = re.findall(r"avg time: ([\d.eE+-]+)", res[0][1])
            its = re.findall(r"first run PCG terminated in (\d+) iterations", res[0][1])
            ko = reference_kernel_only()
            per_call_us = 1e6 / rates[0]
            single = {"ref_bracket_us_avg50": float(avg[-1]) * 1e3 if avg else None,
                      "iters": int(its[-1]) if its else None, "python_call_us": per_call_us,
                      "list_conversion_us": res[0][0]["list_conversion_us"]}
            if ko and "kernel_us_median" in ko:
                single.update(kernel_us_median=ko["kernel_us_median"], us_per_iter=ko["us_per_iter"],
                              kernel_phase_us=ko["phase_us_median"], host_api_us_median=ko["host_api_us_median"],
                              host_overhead_fraction_of_call=1.0 - ko["kernel_us_median"] / per_call_us)
            base.update(value=val, ms_per_step=max(r["wall"] for r, _ in res) / args.steps * 1e3,
                        e2e={"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                        cpu_baseline={"value": val, "unit": UNIT, "cores": ngpu, "kind": "reference",
                                      "sample": f"{n} IIWA-50 systems per step and GPU through the unmodified reference "
                                                f"CUDA build's gpu_library.linsys_solve, {ngpu} process(es), one per "
                                                "GPU, one host thread each (the reference ships no CPU path and no "
                                                "batched entry point); list conversion outside the timed loop"},
                        per_gpu_values=rates, single_solve=single,
                        note="reference = its own CUDA build (oracle/_ref), one process per GPU, called in a loop; "
                             "single_solve.kernel_us_median = its five kernels on pre-allocated buffers "
                             "(oracle/ref_harness.cu): the rest of a call is host overhead (per-solve cudaMalloc / "
                             "cudaFree / cudaDeviceSynchronize / printf, gpu_library.cu:36-81)")
            print(json.dumps(base), flush=True)
            return
        base["reference_cuda_build_error"] = (outs[0][1] or outs[0][0])[-300:]
    # fall back: the NumPy oracle port on the host cores
    cb = cpu_baseline(per_core=max(n // max(os.cpu_count() or 1, 1), 8))
    base.update(value=cb["value"], ms_per_step=None, cpu_baseline=cb,
                e2e={"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                note="reference CUDA build unavailable: NumPy oracle port on host cores")
    print(json.dumps(base), flush=True)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def single_solve_latency(torch, Plan, dev_index, reps=300, knots=K):
    """IIWA-50 single solve: kernel-only (CUDA events around the launch, inputs resident) and the
    reference's bracket (H2D + kernel + D2H, gpu_library.cu:172-187) through gato_solve."""
    from gato_python_b200.problems import iiwa_like

    prob = iiwa_like(S, C, knots, seed=0)
    pl = Plan(S, C, knots, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"], device=dev_index)
    dev = torch.device("cuda", dev_index)
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
    out = {"workload": f"iiwa{knots}_single_solve_seed0", "kernel": pl.kernel(), "iters": int(it.item()),
           "kernel_us_median": float(np.median(ks)), "kernel_us_min": float(np.min(ks)),
           "bracket_us_median": float(np.median(bs)), "bracket_us_p99": float(np.percentile(bs, 99)),
           "cluster": list(pl.cluster()), "reps": reps}
    pl.close()
    return out


def mpc_sequence(Plan, dev_index, nsteps=1000):
    """BASELINE.json config 5: `nsteps` CONSECUTIVE IIWA-50 solves (g, c follow a 1 % random walk, rho swept
    1e-4 .. 1e-1), host buffers in and out.  `cold` = what the reference computes with warm_start=True
    (the flag has no numerical effect there, src/gato_pcg.cuh:300-304); `warm` = the opt-in true warm start
    with lambda chained on the device (gato_solve_sequence); `closed_loop` = one gato_solve per step with
    GATO_WARM_RESIDENT, the form a controller that computes step t+1's g, c from step t's dz would use."""
    from gato_python_b200.problems import iiwa_like
    from gato_python_b200.solver import WARM_REFERENCE, WARM_RESIDENT, WARM_TRUE

    prob = iiwa_like(S, C, K, seed=9)
    rng = np.random.default_rng(0)
    g = np.empty((nsteps, prob["g"].size), np.float32); c = np.empty((nsteps, prob["c"].size), np.float32)
    gi, ci = prob["g"].copy(), prob["c"].copy()
    for t in range(nsteps):
        gi = (gi * (1 + 0.01 * rng.standard_normal(gi.size))).astype(np.float32)
        ci = (ci * (1 + 0.01 * rng.standard_normal(ci.size))).astype(np.float32)
        g[t], c[t] = gi, ci
    rho = np.logspace(-4, -1, nsteps).astype(np.float32)
    pl = Plan(S, C, K, prob["G_row"], prob["G_col"], prob["C_row"], prob["C_col"], device=dev_index)
    out = {"workload": f"iiwa50_mpc_{nsteps}_consecutive_solves_rho_1e-4_to_1e-1", "steps": nsteps}
    for name, mode in (("cold", WARM_REFERENCE), ("warm", WARM_TRUE)):
        pl.solve_sequence(prob["G_val"], prob["C_val"], g[:50], c[:50], rho[:50], None, prob["exit_tol"],
                          prob["max_iters"], mode)
        t0 = time.perf_counter()
        _, _, its, ms = pl.solve_sequence(prob["G_val"], prob["C_val"], g, c, rho, None, prob["exit_tol"],
                                          prob["max_iters"], mode)
        wall = time.perf_counter() - t0
        out[name] = {"solves_per_s": nsteps / wall, "us_per_solve_wall": wall / nsteps * 1e6,
                     "us_per_solve_device": ms * 1e3 / nsteps, "mean_iters": float(its.mean()),
                     "max_iters": int(its.max())}
    t0 = time.perf_counter()
    its = []
    for t in range(nsteps):
        _, _, info = pl.solve(prob["G_val"], prob["C_val"], g[t], c[t], None, prob["exit_tol"], prob["max_iters"],
                              WARM_RESIDENT, float(rho[t]))
        its.append(info["iters"])
    wall = time.perf_counter() - t0
    out["closed_loop_warm_resident"] = {"solves_per_s": nsteps / wall, "us_per_solve_wall": wall / nsteps * 1e6,
                                        "mean_iters": float(np.mean(its))}
    pl.close()
    return out


def parity_sample(b, lam, dz, its, lam2, dz2, its2, n, rank):
    """Oracle check of the timed outputs: `n` systems spread evenly over the batch (first, last, and a seeded choice in
    between).  -> max relative errors (max of 2-norm and max-norm relative) and max iteration-count difference, for the
    device-resident run (lam, dz, its) and the host-buffer run (lam2, dz2, its2)."""
    from oracle import gato_oracle as orc

    nsys = lam.shape[0]
    rng = np.random.default_rng(12345 + rank)
    n = max(min(n, nsys), 1)
    idx = np.unique(np.concatenate([[0, nsys - 1], (np.arange(n) * nsys // n + rng.integers(0, max(nsys // n, 1), n)) % nsys]))[:max(n, 2)]
    lam0 = np.zeros(S * K, np.float32)

    def rel(a, ref):
        ref = ref.astype(np.float64); d = a.astype(np.float64) - ref
        return max(np.linalg.norm(d) / max(np.linalg.norm(ref), 1e-300), np.abs(d).max() / max(np.abs(ref).max(), 1e-300))
    out = dict(n=len(idx), systems=[int(i) for i in idx[:8]] + ["..."], max_rel_lambda=0.0, max_rel_dz=0.0, max_iter_diff=0,
               max_rel_lambda_e2e=0.0, max_rel_dz_e2e=0.0, max_iter_diff_e2e=0)
    for i in idx:
        lo, do, info = orc.linsys_solve_info(S, C, K, b["G_row"], b["G_col"], b["G_val"][i], b["C_row"], b["C_col"],
                                             b["C_val"][i], b["g"][i], b["c"][i], lam0, 1, b["exit_tol"], b["max_iters"],
                                             False, b["rho"])
        out["max_rel_lambda"] = max(out["max_rel_lambda"], rel(lam[i], lo))
        out["max_rel_dz"] = max(out["max_rel_dz"], rel(dz[i], do))
        out["max_iter_diff"] = max(out["max_iter_diff"], abs(int(its[i]) - info["iters"]))
        out["max_rel_lambda_e2e"] = max(out["max_rel_lambda_e2e"], rel(lam2[i], lo))
        out["max_rel_dz_e2e"] = max(out["max_rel_dz_e2e"], rel(dz2[i], do))
        out["max_iter_diff_e2e"] = max(out["max_iter_diff_e2e"], abs(int(its2[i]) - info["iters"]))
    return out


def ours_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    from gato_python_b200 import Plan
    from gato_python_b200.problems import iiwa_like_batch_fast

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # CPU baseline first (rank 0, N == 1 only), before the GPU is busy
    cpu = refk = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(per_core=args.cpu_per_core, scipy_per_core=args.scipy_per_core)
        refk = reference_kernel_only()

    # host placement: run and pin memory on the GPU's own NUMA node (after the CPU baseline used every core)
    numa = None
    if os.environ.get("GATO_B200_NUMA_BIND", "1") != "0":
        from gato_python_b200.sharding import bind_to_gpu_numa
        numa = bind_to_gpu_numa(local_rank)

    nsys = args.nsys
    b = iiwa_like_batch_fast(nsys, S, C, K, seed=SEED, first=rank * nsys)
    nnzG, nnzC = b["G_val"].shape[1], b["C_val"].shape[1]
    pl = Plan(S, C, K, b["G_row"], b["G_col"], b["C_row"], b["C_col"], device=local_rank, max_batch=nsys)
    if args.kernel:
        pl.set_kernel(args.kernel)
    kernel_id = pl.kernel_for(nsys)
    if kernel_id == 5:
        pl.set_profiling(True)      # CUDA events around each half of kernel 5, on the launching stream
    n_l, n_z = pl.n_l, pl.n_z

    # pinned host copies (the e2e path reads these) and device-resident copies (the `value` path)
    host = {k: torch.from_numpy(b[k]).pin_memory() for k in ("G_val", "C_val", "g", "c")}
    d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
    lam = torch.empty((nsys, n_l), device=dev); dz = torch.empty((nsys, n_z), device=dev)
    its = torch.empty(nsys, dtype=torch.int32, device=dev)
    h_lam = torch.empty((nsys, n_l), dtype=torch.float32).pin_memory()
    h_dz = torch.empty((nsys, n_z), dtype=torch.float32).pin_memory()
    h_it = torch.empty(nsys, dtype=torch.int32).pin_memory()
    stream = torch.cuda.current_stream()

    def step():
        pl.solve_batched_device(nsys, d["G_val"].data_ptr(), nnzG, d["C_val"].data_ptr(), nnzC,
                                d["g"].data_ptr(), n_z, d["c"].data_ptr(), n_l, lam.data_ptr(), dz.data_ptr(),
                                its.data_ptr(), exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"],
                                stream=stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record(stream)
    halves, launches = [], 0
    for i in range(args.steps):
        step()
        evs[i + 1].record(stream)
        if kernel_id == 5 and i >= args.steps - 3:      # last steps only: reading the events waits for the step
            a, b2, nl = pl.last_profile()
            halves.append((a, b2))
            launches = nl
    barrier()
    total_ms = evs[0].elapsed_time(evs[-1])
    kern_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    iters_host = its.cpu().numpy()
    lam_host, dz_host = lam.cpu().numpy(), dz.cpu().numpy()      # the TIMED outputs (last step), checked below
    pl.set_profiling(False)

    # ---- end to end: host buffers through gato_solve_batched (H2D + kernel + D2H every step) ----
    hv = {k: v.numpy() for k, v in host.items()}
    outs = (h_lam.numpy(), h_dz.numpy(), h_it.numpy())

    def e2e_step():
        return pl.solve_batched(hv["G_val"], hv["C_val"], hv["g"], hv["c"], None, b["exit_tol"], b["max_iters"],
                                0, b["rho"], out=outs)[3]
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_dev_ms = 0.0
    for _ in range(args.steps):
        e2e_dev_ms += e2e_step()
    barrier()
    e2e_wall = time.perf_counter() - t0
    clocks = sampler.stop()
    e2e_iters = h_it.numpy().copy()   # before the shared-matrix variant below reuses the output buffers
    e2e_lam, e2e_dz = h_lam.numpy().copy(), h_dz.numpy().copy()

    # ---- what the host <-> device links of this box carry at the same time on all ranks: a plain pinned-memory copy of
    # the same bytes both ways (the ceiling of any host-buffer pipeline; shows whether e2e is bound by the host)
    probe_bytes_up = sum(v.numel() * 4 for v in host.values())
    probe_bytes_dn = (h_lam.numel() + h_dz.numel() + h_it.numel()) * 4
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    def copy_probe():
        with torch.cuda.stream(s_up):
            for k2, v in host.items():
                d[k2].copy_(v, non_blocking=True)
        with torch.cuda.stream(s_dn):
            h_lam.copy_(lam, non_blocking=True); h_dz.copy_(dz, non_blocking=True); h_it.copy_(its, non_blocking=True)
    copy_probe()
    barrier()
    t0 = time.perf_counter()
    for _ in range(max(args.steps // 2, 3)):
        copy_probe()
    barrier()
    probe_wall = (time.perf_counter() - t0) / max(args.steps // 2, 3)

    # ---- the other reading of config 4 ("randomised g/c"): ONE matrix pair shared by all systems, only g and
    # c differ -- every system is still unpacked, factored and solved on its own; only the PCIe bytes shrink
    def shared_step():
        return pl.solve_batched(hv["G_val"][0], hv["C_val"][0], hv["g"], hv["c"], None, b["exit_tol"],
                                b["max_iters"], 0, b["rho"], out=outs)[3]
    for _ in range(2):
        shared_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        shared_step()
    barrier()
    shared_wall = time.perf_counter() - t0
    shared_iters = float(h_it.numpy().mean())

    # the same reading with everything resident in HBM: ld = 0 for the matrices.  Kernel 5 then runs its first half ONCE
    # per call (gato_plan_set_shared_matrix_path) -- timed with and without that path, checked against the oracle
    def shared_resident(steps):
        def st():
            pl.solve_batched_device(nsys, d["G_val"].data_ptr(), 0, d["C_val"].data_ptr(), 0, d["g"].data_ptr(), n_z,
                                    d["c"].data_ptr(), n_l, lam.data_ptr(), dz.data_ptr(), its.data_ptr(),
                                    exit_tol=b["exit_tol"], max_iters=b["max_iters"], rho=b["rho"], stream=stream.cuda_stream)
        for _ in range(3):
            st()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            st()
        e1.record(stream)
        barrier()
        return e0.elapsed_time(e1)
    sh_ms = shared_resident(args.steps)
    sh_lam, sh_dz, sh_it = lam.cpu().numpy(), dz.cpu().numpy(), its.cpu().numpy()
    sh_general_ms = None
    if kernel_id == 5:
        pl.set_shared_matrix_path(False)
        sh_general_ms = shared_resident(args.steps)
        pl.set_shared_matrix_path(True)
    bs = dict(b, G_val=np.broadcast_to(b["G_val"][0], b["G_val"].shape), C_val=np.broadcast_to(b["C_val"][0], b["C_val"].shape))
    sh_par = parity_sample(bs, sh_lam, sh_dz, sh_it, sh_lam, sh_dz, sh_it, max(args.parity_sample // 2, 2), rank)

    t = torch.tensor([total_ms, e2e_wall * 1e3, e2e_dev_ms, shared_wall * 1e3, probe_wall * 1e3, sh_ms, sh_general_ms or 0.0,
                      sh_par["max_rel_lambda"], sh_par["max_rel_dz"], float(sh_par["max_iter_diff"])], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_wall_ms, e2e_dev_ms, shared_wall_ms, probe_wall_ms, sh_ms, sh_general_ms, sh_rl, sh_rd, sh_di = (float(x) for x in t.tolist())

    # parity of the TIMED outputs against the oracle: a seeded sample of systems spread over this rank's batch, from
    # the buffers the timed device-resident run and the timed host-buffer run left behind
    parity = parity_sample(b, lam_host, dz_host, iters_host, e2e_lam, e2e_dz, e2e_iters, args.parity_sample, rank)
    pt = torch.tensor([parity["max_rel_lambda"], parity["max_rel_dz"], float(parity["max_iter_diff"]),
                       parity["max_rel_lambda_e2e"], parity["max_rel_dz_e2e"], float(parity["max_iter_diff_e2e"])],
                      dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(pt, op=dist.ReduceOp.MAX)
    pmax = [float(x) for x in pt.tolist()]

    if rank == 0:
        ms_per_step = total_ms / args.steps
        value = nsys * world * args.steps / (total_ms * 1e-3)
        bps = bytes_per_system(nnzG, nnzC)
        peak, peak_src = measured_peaks()
        kavg_ms = float(np.mean(kern_ms))
        ach = bps * nsys / (kavg_ms * 1e-3) / 1e9
        fl = float(np.sum(flops_per_system(iters_host)))
        tfl = fl / (kavg_ms * 1e-3) / 1e12
        h2d = 4 * nsys * (nnzG + nnzC + n_z + n_l)
        d2h = 4 * nsys * (n_l + n_z + 1)
        cfg = bench_config(args, world)
        stats = {"nnzG": nnzG, "nnzC": nnzC, "mean_pcg_iters": float(iters_host.mean()),
                 "max_pcg_iters": int(iters_host.max()), "kernel": kernel_id,
                 "cache": f"inputs {bps * nsys / 1e6:.0f} MB per launch > 126 MB L2 (no flush needed)"}
        if kernel_id == 5 and halves:
            # the dominant kernel of kernel 5 is its second half (PCG + dz); its duration comes from CUDA events recorded
            # around it on the launching stream (gato_plan_set_profiling), summed over the chunks of a step
            h1 = float(np.mean([h[0] for h in halves])); h2 = float(np.mean([h[1] for h in halves]))
            kname, kdur_ms = "gato_pcg_scaled", h2
            launches_per_step = launches
            halves_ms = {"gato_fused_solve_v2<SCALED> (unpack + Schur + FP64 congruence)": h1,
                         "gato_pcg_scaled (PCG on the scaled system + dz)": h2, "share_of_step": h2 / (h1 + h2)}
            scratch = 4 * nsys * K * 1088          # floats the first half writes and the second reads (ScaledLayout)
        else:
            kname = "gato_fused_solve_v2" if kernel_id == 2 else "gato_fused_solve"
            kdur_ms, launches_per_step, halves_ms, scratch = kavg_ms, 1, None, 0
        ach_k = bps * nsys / (kdur_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "stats": stats,
            "e2e": {"value": nsys * world * args.steps / (e2e_wall_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "timing": "wall clock, max over ranks",
                    "device_event_value": nsys * world * args.steps / (e2e_dev_ms * 1e-3),
                    "host_numa_binding": numa,
                    "bytes_per_system": (h2d + d2h) / nsys,
                    "aggregate_gbs": (h2d + d2h) * world * args.steps / (e2e_wall_ms * 1e-3) / 1e9,
                    "pinned_copy_probe": {"aggregate_gbs": (probe_bytes_up + probe_bytes_dn) * world / (probe_wall_ms * 1e-3) / 1e9,
                                          "solves_per_s_ceiling": nsys * world / (probe_wall_ms * 1e-3),
                                          "what": "the same H2D and D2H bytes as one e2e step, plain cudaMemcpyAsync from / "
                                                  "to pinned memory on two streams, all ranks at once, max over ranks: "
                                                  "no host-buffer pipeline on this box can beat it"}},
            "e2e_shared_matrix": {"value": nsys * world * args.steps / (shared_wall_ms * 1e-3), "unit": UNIT,
                                  "h2d_bytes_per_step": 4 * (nnzG + nnzC + nsys * (n_z + n_l)),
                                  "d2h_bytes_per_step": d2h, "mean_pcg_iters": shared_iters,
                                  "note": "same call, G_val/C_val of system 0 shared by all systems (ld = 0), g and c "
                                          "per system; kernel 5 forms and factors the Schur complement once per pipeline "
                                          "chunk, every system runs its own PCG and dz"},
            "shared_matrix_resident": {"value": nsys * world * args.steps / (sh_ms * 1e-3), "unit": UNIT,
                                       "ms_per_step": sh_ms / args.steps,
                                       "per_system_path_value": (nsys * world * args.steps / (sh_general_ms * 1e-3)) if sh_general_ms else None,
                                       "mean_pcg_iters": float(sh_it.mean()),
                                       "parity": {"n": sh_par["n"] * world, "max_rel_lambda": sh_rl, "max_rel_dz": sh_rd,
                                                  "max_iter_diff": int(sh_di), "pass": bool(max(sh_rl, sh_rd) <= 1e-4 and sh_di <= 1)},
                                       "note": "inputs resident, ONE G_val / C_val for the call (ld = 0), g and c per system: "
                                               "kernel 5 runs unpack + Schur + FP64 congruence once, gato_gamma_scaled forms "
                                               "every system's gamma^, gato_pcg_scaled does PCG + dz per system; "
                                               "per_system_path_value = the same call with that path switched off"},
            "parity_sample": {"n": parity["n"] * world, "systems": parity["systems"], "checker": "NumPy oracle (oracle/gato_oracle.py)",
                              "max_rel_lambda": pmax[0], "max_rel_dz": pmax[1], "max_iter_diff": int(pmax[2]),
                              "e2e": {"max_rel_lambda": pmax[3], "max_rel_dz": pmax[4], "max_iter_diff": int(pmax[5])},
                              "bar": "2-norm and max-norm relative <= 1e-4, iterations +-1", 
                              "pass": bool(max(pmax[0], pmax[1], pmax[3], pmax[4]) <= 1e-4 and max(pmax[2], pmax[5]) <= 1)},
            "gpu_launches": args.steps * launches_per_step,
            "roofline": {"bound": "hbm", "achieved": ach_k, "peak": peak, "unit": "GB/s", "frac": ach_k / peak,
                         "traffic": ncu_traffic(kname, nsys), "peak_source": peak_src,
                         "kernel": kname, "algorithmic_bytes_per_launch": bps * nsys,
                         "kernel_ms_avg": kdur_ms, "algorithmic_bytes_per_system": bps,
                         "launches_per_step": launches_per_step, "kernel_5_halves_ms_per_step": halves_ms,
                         "scratch_bytes_per_step_written_then_read": scratch,
                         "whole_step": {"achieved": ach, "frac": ach / peak, "ms": kavg_ms},
                         "fp32": {"achieved_tflops": tfl, "peak_tflops": FP32_PEAK_TFLOPS,
                                  "frac": tfl / FP32_PEAK_TFLOPS, "flops_per_launch": fl,
                                  "peak_source": "theoretical 148 SM x 128 lanes x 2 x 1.965 GHz",
                                  "note": "whole step (all launches), SURVEY 8(d) algorithmic flop count of the "
                                          "reference's formulation; arithmetic intensity ~100 flop/B >> 11.5 flop/B "
                                          "machine balance: FP32-pipe/latency bound, the HBM fraction is a sanity ceiling"}},
            "cpu_baseline": cpu,
            "clocks": clocks,
        }
        if refk is not None:
            line["reference_kernel_only"] = refk
        if world == 1 and not args.no_single:
            line["single_solve"] = single_solve_latency(torch, Plan, local_rank)
            if refk and "kernel_us_median" in refk:
                line["single_solve"]["reference_kernel_only_us"] = refk["kernel_us_median"]
                line["single_solve"]["speedup_vs_reference_kernels"] = refk["kernel_us_median"] / line["single_solve"]["kernel_us_median"]
                line["single_solve"]["reference_us_per_iter"] = refk["us_per_iter"]
            line["single_solve_k256"] = single_solve_latency(torch, Plan, local_rank, reps=100, knots=256)
            line["mpc_sequence"] = mpc_sequence(Plan, local_rank)
        print(json.dumps(line), flush=True)
    pl.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def single_process_multi_gpu(args):
    """`--single-process-gpus N`: ONE process drives N GPUs through gato_solve_batched_multi (one host thread per device,
    contiguous slices of ONE pinned host buffer each way).  The timed region is H2D + solve + the D2H
    gather; prints one JSON line.  Not the driver's contract (that is torchrun, one process per GPU): an extra view."""
    import torch

    from gato_python_b200 import MultiPlan
    from gato_python_b200.problems import iiwa_like_batch_fast

    n_gpu = args.single_process_gpus
    nsys = args.nsys * n_gpu
    b = iiwa_like_batch_fast(nsys, S, C, K, seed=SEED, first=0)
    mp = MultiPlan(S, C, K, b["G_row"], b["G_col"], b["C_row"], b["C_col"], devices=list(range(n_gpu)), max_batch=args.nsys)
    host = {k: torch.from_numpy(b[k]).pin_memory().numpy() for k in ("G_val", "C_val", "g", "c")}
    out = (torch.empty((nsys, mp.n_l)).pin_memory().numpy(), torch.empty((nsys, mp.n_z)).pin_memory().numpy(),
           torch.empty(nsys, dtype=torch.int32).pin_memory().numpy())

    def step():
        return mp.solve_batched(host["G_val"], host["C_val"], host["g"], host["c"], None, b["exit_tol"], b["max_iters"],
                                0, b["rho"], out=out)[3]
    for _ in range(max(args.warmup, 3)):
        step()
    t0 = time.perf_counter()
    dev_ms = [step() for _ in range(args.steps)]
    wall = time.perf_counter() - t0
    par = parity_sample(b, out[0], out[1], out[2], out[0], out[1], out[2], args.parity_sample, 0)
    nnzG, nnzC = b["G_val"].shape[1], b["C_val"].shape[1]
    h2d = 4 * nsys * (nnzG + nnzC + mp.n_z + mp.n_l)
    d2h = 4 * nsys * (mp.n_l + mp.n_z + 1)
    print(json.dumps({"metric": METRIC, "mode": "single process, one host thread per GPU (gato_solve_batched_multi)",
                      "n_gpus": n_gpu, "value": nsys * args.steps / wall, "unit": UNIT, "steps": args.steps,
                      "ms_per_step": wall / args.steps * 1e3, "longest_device_ms_per_step": float(np.mean(dev_ms)),
                      "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "aggregate_gbs": (h2d + d2h) * args.steps / wall / 1e9, "config": bench_config(args, n_gpu),
                      "parity_sample": {k: par[k] for k in ("n", "max_rel_lambda", "max_rel_dz", "max_iter_diff")}}), flush=True)
    mp.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nsys", type=int, default=8192, help="systems per GPU per step")
    ap.add_argument("--ref-sample", type=int, default=64, help="systems per step for --impl reference")
    ap.add_argument("--cpu-per-core", type=int, default=1024, help="systems per process for cpu_baseline")
    ap.add_argument("--scipy-per-core", type=int, default=8, help="systems per process for the SciPy baselines (0 = skip)")
    ap.add_argument("--parity-sample", type=int, default=64, help="systems per rank checked against the oracle")
    ap.add_argument("--kernel", type=int, default=0, help="force a kernel (0 = automatic selection)")
    ap.add_argument("--single-process-gpus", type=int, default=0,
                    help="extra mode: one process drives N GPUs through gato_solve_batched_multi (host buffers)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-single", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    if args.single_process_gpus > 0:
        if rank == 0:
            single_process_multi_gpu(args)
        return
    ours_arm(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
