"""ctypes loader of oracle/_ref/S{S}_C{C}_K{K}/libref_harness.so (oracle/ref_harness.cu: the reference's own
src/*.cuh #included UNCHANGED and driven on pre-allocated buffers).  TEST INFRASTRUCTURE ONLY: used by the
parity tests (dumps of the reference's d_G_dense / d_C_dense / d_S / d_Pinv / d_gamma) and by the reference
legs of bench.py / oracle/ref_kernel_time.py (kernel-only time of the reference's five kernels).

The sizes are compile-time constants of the reference (CMakeLists.txt:18), so there is one library per
configuration; each is an ordinary shared object with its own globals and several can be loaded side by side.
The reference's host functions print to the C stdout (src/gato_schur.cuh:899,981): `quiet()` mutes fd 1."""
import contextlib
import ctypes as ct
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_f = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def harness_path(S, C, K):
    p = os.path.join(ROOT, "oracle", "_ref", f"S{S}_C{C}_K{K}", "libref_harness.so")
    return p if os.path.exists(p) else None


@contextlib.contextmanager
def quiet():
    import sys
    sys.stdout.flush()
    keep = os.dup(1)
    null = os.open(os.devnull, os.O_WRONLY)
    os.dup2(null, 1)
    try:
        yield
    finally:
        os.dup2(keep, 1)
        os.close(keep)
        os.close(null)


class RefHarness:
    def __init__(self, S, C, K):
        p = harness_path(S, C, K)
        if p is None:
            raise FileNotFoundError(f"no reference harness for S={S} C={C} K={K} (python oracle/build_ref.py)")
        self.lib = ct.CDLL(p)
        L = self.lib
        L.refh_last_error.restype = ct.c_char_p
        L.refh_upload.argtypes = [_i, _i, _f, ct.c_int, _i, _i, _f, ct.c_int, _f, _f]
        opt = ct.c_void_p
        L.refh_run.argtypes = [ct.c_int, ct.c_float, ct.c_float, ct.c_int] + [opt] * 8 + [ct.POINTER(ct.c_int)]
        L.refh_time.argtypes = [ct.c_int, ct.c_float, ct.c_float, ct.c_int, ct.c_int, ct.c_int, opt, opt,
                                ct.POINTER(ct.c_int)]
        L.refh_set_padding.argtypes = [ct.c_int]
        s, c, k = ct.c_int(), ct.c_int(), ct.c_int()
        L.refh_sizes(ct.byref(s), ct.byref(c), ct.byref(k))
        assert (s.value, c.value, k.value) == (S, C, K)
        self.S, self.C, self.K = S, C, K
        self.n_l, self.n_z = S * K, (S + C) * K - C
        self.g_dense = (S * S + C * C) * K - C * C
        self.c_dense = (S * S + S * C) * (K - 1)

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed ({rc}): {self.lib.refh_last_error().decode()}")

    def set_padding(self, floats):
        self.lib.refh_set_padding(int(floats))

    def upload(self, prob):
        a = lambda k, t: np.ascontiguousarray(prob[k], t)   # noqa: E731
        self._keep = [a("G_row", np.int32), a("G_col", np.int32), a("G_val", np.float32), a("C_row", np.int32),
                      a("C_col", np.int32), a("C_val", np.float32), a("g", np.float32), a("c", np.float32)]
        k = self._keep
        self._check(self.lib.refh_upload(k[0], k[1], k[2], k[2].size, k[3], k[4], k[5], k[5].size, k[6], k[7]),
                    "refh_upload")
        self.prob = prob

    def run(self, mode=1):
        """One solve of the uploaded problem by the reference's own kernels; every intermediate is returned."""
        S3 = 3 * self.S * self.S * self.K
        out = dict(G_dense_pre=np.empty(self.g_dense, np.float32), C_dense=np.empty(max(self.c_dense, 1), np.float32),
                   S=np.empty(S3, np.float32), Pinv=np.empty(S3, np.float32), gamma=np.empty(self.n_l, np.float32),
                   G_inv=np.empty(self.g_dense, np.float32), lam=np.empty(self.n_l, np.float32),
                   dz=np.empty(self.n_z, np.float32))
        it = ct.c_int(-1)
        p = self.prob
        ptr = lambda a: a.ctypes.data_as(ct.c_void_p)   # noqa: E731
        with quiet():
            rc = self.lib.refh_run(mode, float(p["rho"]), float(p["exit_tol"]), int(p["max_iters"]),
                                   ptr(out["G_dense_pre"]), ptr(out["C_dense"]), ptr(out["S"]), ptr(out["Pinv"]),
                                   ptr(out["gamma"]), ptr(out["G_inv"]), ptr(out["lam"]), ptr(out["dz"]), ct.byref(it))
        self._check(rc, "refh_run")
        out["C_dense"] = out["C_dense"][:self.c_dense]
        out["iters"] = it.value
        return out

    def time(self, mode=1, reps=1000, warmup=50):
        """-> dict(total_us[reps], phase_us[5][reps] (mode 1), iters)"""
        tot = np.zeros(reps, np.float32)
        ph = np.zeros((5, reps), np.float32)
        it = ct.c_int(-1)
        p = self.prob
        with quiet():
            rc = self.lib.refh_time(mode, float(p["rho"]), float(p["exit_tol"]), int(p["max_iters"]), reps, warmup,
                                    tot.ctypes.data_as(ct.c_void_p), ph.ctypes.data_as(ct.c_void_p) if mode == 1 else None,
                                    ct.byref(it))
        self._check(rc, "refh_time")
        return dict(total_us=tot, phase_us=ph if mode == 1 else None, iters=it.value)

    def close(self):
        self.lib.refh_release()


PHASES = ("gato_convert_kkt_format", "gato_form_schur_jacobi", "gato_form_ss", "parallelPCG", "gato_compute_dz")


def kernel_only_summary(S, C, K, prob, reps=1000, warmup=50):
    """BASELINE.md 2.1 rows R3 / R4 for one problem: the reference's five kernels on pre-allocated buffers."""
    h = RefHarness(S, C, K)
    h.upload(prob)
    k = h.time(mode=1, reps=reps, warmup=warmup)
    a = h.time(mode=0, reps=max(reps // 10, 20), warmup=5)
    h.close()
    ph = {n: float(np.median(k["phase_us"][i])) for i, n in enumerate(PHASES)}
    it = k["iters"]
    return {"kernel_us_median": float(np.median(k["total_us"])), "kernel_us_min": float(np.min(k["total_us"])),
            "kernel_us_p99": float(np.percentile(k["total_us"], 99)), "iters": it, "phase_us_median": ph,
            # the loop runs iterations 0..iters inclusive (0-based exit index, src/gato_pcg.cuh:404-411)
            "us_per_iter": ph["parallelPCG"] / (it + 1) if it >= 0 else None,
            "host_api_us_median": float(np.median(a["total_us"])), "reps": reps,
            "what": "five reference kernels launched back to back on pre-allocated buffers, one cudaEvent pair "
                    "(no cudaMalloc/cudaFree, no cudaDeviceSynchronize, no printf); host_api = form_schur + "
                    "solve_pcg<float> + compute_dz as they are, on the same buffers"}
