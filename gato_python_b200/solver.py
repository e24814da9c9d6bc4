"""ctypes binding of the C ABI (include/gato_b200.h) and a host-side mirror of the reference's
operator interface `gpu_library.linsys_solve` (/root/reference/gpu_library.cu:85-87, 236-239).

There is NO CPU fallback here: if `libgato_b200.so` is missing the import of the library fails,
and if no CUDA device is present every solve raises `GatoError`.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# GATO_B200_LIB: another build of the same library (development builds, build.py --dev)
LIB_PATH = os.environ.get("GATO_B200_LIB") or os.path.join(_HERE, "libgato_b200.so")

WARM_OFF, WARM_REFERENCE, WARM_TRUE, WARM_RESIDENT = 0, 1, 2, 3
PRECOND_STAIR, PRECOND_BLOCK_JACOBI = 0, 1

#: every symbol include/gato_b200.h declares
SYMBOLS = [
    "gato_supported_sizes", "gato_last_error", "gato_device_count", "gato_plan_create",
    "gato_plan_create_blocks", "gato_solve_sequence",
    "gato_plan_destroy", "gato_plan_set_cluster", "gato_plan_get_cluster", "gato_plan_set_kernel", "gato_plan_set_pipeline", "gato_plan_set_staging", "gato_plan_set_zero_copy",
    "gato_plan_set_preconditioner", "gato_plan_get_preconditioner",
    "gato_scaled_layout", "gato_scaled_debug", "gato_plan_set_scaled_chunk", "gato_plan_get_kernel_for",
    "gato_plan_set_profiling", "gato_plan_last_profile", "gato_solve_batched_multi",
    "gato_plan_set_scaled_variant", "gato_plan_set_pcg_form", "gato_plan_set_shared_matrix_path",
    "gato_noalloc_max_systems", "gato_solve_batched_device_noalloc",
    "gato_plan_get_kernel", "gato_solve",
    "gato_solve_batched", "gato_solve_batched_device", "gato_unpack", "gato_solve_debug",
]


class GatoError(RuntimeError):
    pass


_lib = None
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)
_ll = C.c_longlong


def lib():
    """Load the native library (fails loudly when it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GatoError(f"{LIB_PATH} not built: run `python -m gato_python_b200.build` "
                        "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.gato_last_error.restype = C.c_char_p
    L.gato_supported_sizes.argtypes = [_ip, C.c_int]
    L.gato_plan_create.argtypes = [C.c_int] * 3 + [_ip, _ip, C.c_int, _ip, _ip, C.c_int, C.c_int, C.c_int,
                                                  C.POINTER(C.c_void_p)]
    L.gato_plan_create_blocks.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_void_p)]
    L.gato_solve_sequence.argtypes = [C.c_void_p, C.c_int, _fp, _ll, _fp, _ll, _fp, _ll, _fp, _ll, _fp, C.c_float,
                                      _fp, _fp, _fp, _ip, C.c_float, C.c_int, C.c_int, _fp]
    L.gato_plan_destroy.argtypes = [C.c_void_p]
    L.gato_plan_destroy.restype = None
    L.gato_plan_set_cluster.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_get_cluster.argtypes = [C.c_void_p, _ip, _ip]
    L.gato_plan_set_kernel.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_set_pipeline.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_set_staging.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_set_zero_copy.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_set_preconditioner.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_get_preconditioner.argtypes = [C.c_void_p]
    L.gato_scaled_layout.argtypes = [C.c_void_p, _ip]
    L.gato_scaled_debug.argtypes = [C.c_void_p, _fp, _fp, _fp, _fp, C.c_float, _fp, _ll]
    L.gato_plan_set_scaled_chunk.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_get_kernel_for.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.gato_solve_batched_multi.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, _fp, _ll, _fp, _ll, _fp, _ll, _fp, _ll,
                                           _fp, _ll, _fp, _fp, _ip, C.c_float, C.c_int, C.c_int, C.c_float, _fp]
    L.gato_plan_set_scaled_variant.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_set_pcg_form.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_set_shared_matrix_path.argtypes = [C.c_void_p, C.c_int]
    L.gato_noalloc_max_systems.argtypes = [C.c_void_p]
    L.gato_solve_batched_device_noalloc.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p, _ll] * 4 + \
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_float, C.c_void_p]
    L.gato_plan_set_profiling.argtypes = [C.c_void_p, C.c_int]
    L.gato_plan_last_profile.argtypes = [C.c_void_p, _fp, _fp, _ip]
    L.gato_plan_get_kernel.argtypes = [C.c_void_p]
    L.gato_solve.argtypes = [C.c_void_p, _fp, _fp, _fp, _fp, _fp, _fp, _fp, C.c_float, C.c_int, C.c_int,
                             C.c_float, _ip, _fp]
    L.gato_solve_batched.argtypes = [C.c_void_p, C.c_int, _fp, _ll, _fp, _ll, _fp, _ll, _fp, _ll, _fp, _ll,
                                     _fp, _fp, _ip, C.c_float, C.c_int, C.c_int, C.c_float, _fp]
    L.gato_solve_batched_device.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p, _ll] * 5 + \
        [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_float, C.c_void_p]
    L.gato_unpack.argtypes = [C.c_void_p, _fp, _fp, C.c_float, _fp, _fp]
    L.gato_solve_debug.argtypes = [C.c_void_p, _fp, _fp, _fp, _fp, _fp, _fp, C.c_float, C.c_int, C.c_float,
                                   _ip, _fp, _fp, _fp, _fp, C.POINTER(C.c_longlong)]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        msg = lib().gato_last_error().decode()
        if rc in (-1, -2):
            raise ValueError(msg)
        raise GatoError(f"{msg} (code {rc})")


def _f(a):
    return a.ctypes.data_as(_fp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None and a.shape != shape:
        raise ValueError(f"expected shape {shape}, got {a.shape}")
    return a


class _DLDevice(C.Structure):
    _fields_ = [("device_type", C.c_int), ("device_id", C.c_int)]


class _DLDataType(C.Structure):
    _fields_ = [("code", C.c_uint8), ("bits", C.c_uint8), ("lanes", C.c_uint16)]


class _DLTensor(C.Structure):
    _fields_ = [("data", C.c_void_p), ("device", _DLDevice), ("ndim", C.c_int), ("dtype", _DLDataType),
                ("shape", C.POINTER(C.c_int64)), ("strides", C.POINTER(C.c_int64)), ("byte_offset", C.c_uint64)]


def _dlpack_as_cai(a, name, device):
    """Read a DLPack tensor (an object with __dlpack__, or a "dltensor" PyCapsule) into the dict shape of
    __cuda_array_interface__.  The capsule is not consumed: it stays owned by its producer and is released with it
    (the caller keeps the tensor alive for the duration of the asynchronous call, as with any device pointer)."""
    cap = a if type(a).__name__ == "PyCapsule" else a.__dlpack__()
    api = C.pythonapi
    api.PyCapsule_IsValid.argtypes = [C.py_object, C.c_char_p]
    api.PyCapsule_IsValid.restype = C.c_int
    api.PyCapsule_GetPointer.argtypes = [C.py_object, C.c_char_p]
    api.PyCapsule_GetPointer.restype = C.c_void_p
    if not api.PyCapsule_IsValid(cap, b"dltensor"):
        raise TypeError(f"{name}: not an unconsumed DLPack capsule")
    t = C.cast(api.PyCapsule_GetPointer(cap, b"dltensor"), C.POINTER(_DLTensor)).contents   # DLManagedTensor starts with it
    if t.device.device_type not in (2, 13):          # kDLCUDA, kDLCUDAManaged
        raise ValueError(f"{name}: DLPack tensor is not on a CUDA device (device_type {t.device.device_type})")
    if t.device.device_type == 2 and t.device.device_id != device:
        raise ValueError(f"{name}: DLPack tensor lives on cuda:{t.device.device_id}, the plan on cuda:{device}")
    if t.dtype.lanes != 1 or (t.dtype.code, t.dtype.bits) not in ((2, 32), (0, 32)):
        raise ValueError(f"{name}: DLPack dtype (code {t.dtype.code}, {t.dtype.bits} bits) is not float32 / int32")
    shape = tuple(int(t.shape[i]) for i in range(t.ndim))
    item = 4
    strides = None if not t.strides else tuple(int(t.strides[i]) * item for i in range(t.ndim))
    cai = {"data": (int(t.data or 0) + int(t.byte_offset), False), "shape": shape, "strides": strides,
           "typestr": "<f4" if t.dtype.code == 2 else "<i4", "_keepalive": cap}
    return cai


def supported_sizes():
    buf = (C.c_int * 32)()
    n = lib().gato_supported_sizes(buf, 16)
    return [(buf[2 * i], buf[2 * i + 1]) for i in range(min(n, 16))]


def device_count():
    return lib().gato_device_count()


class MultiPlan:
    """One process driving several GPUs: a Plan per device over the same sparsity pattern;
    `solve_batched` cuts the batch into contiguous slices, one host thread per device, results gathered by each
    device's own D2H into disjoint slices of ONE host buffer (gato_solve_batched_multi).  No collective."""

    def __init__(self, S, C_, K, G_row, G_col, C_row, C_col, devices=None, max_batch=1):
        devices = list(range(device_count())) if devices is None else list(devices)
        if not devices:
            raise GatoError("no CUDA device")
        self.plans = [Plan(S, C_, K, G_row, G_col, C_row, C_col, device=d, max_batch=max_batch) for d in devices]
        self.devices = devices
        p0 = self.plans[0]
        self.S, self.C, self.K, self.n_z, self.n_l, self.nnzG, self.nnzC = p0.S, p0.C, p0.K, p0.n_z, p0.n_l, p0.nnzG, p0.nnzC

    def solve_batched(self, G_val, C_val, g, c, lambda_in=None, exit_tol=1e-6, max_iters=500, warm_start=WARM_OFF,
                      rho=0.0, out=None):
        """Same arguments and results as Plan.solve_batched; device_ms is the longest device time."""
        g = np.ascontiguousarray(g, np.float32)
        if g.ndim != 2 or g.shape[1] != self.n_z:
            raise ValueError(f"g must be [nsys, {self.n_z}]")
        nsys = g.shape[0]

        def prep(a, n):
            a = np.ascontiguousarray(a, np.float32)
            if a.ndim == 1:
                a = a[None]
            if a.shape[1] != n or a.shape[0] not in (1, nsys):
                raise ValueError(f"expected [{nsys} or 1, {n}], got {a.shape}")
            return a, (0 if (a.shape[0] == 1 and nsys > 1) else n)

        G_val, ldG = prep(G_val, self.nnzG)
        C_val, ldC = prep(C_val, self.nnzC)
        c, ldc = prep(c, self.n_l)
        lam_in, ldl = (prep(lambda_in, self.n_l) if lambda_in is not None else (None, 0))
        if out is None:
            out = (np.empty((nsys, self.n_l), np.float32), np.empty((nsys, self.n_z), np.float32),
                   np.empty(nsys, np.int32))
        lam, dz, iters = out
        for a, shp, dt in ((lam, (nsys, self.n_l), np.float32), (dz, (nsys, self.n_z), np.float32),
                           (iters, (nsys,), np.int32)):
            if a.shape != shp or a.dtype != dt or not a.flags.c_contiguous:
                raise ValueError(f"out array must be C-contiguous {shp} {np.dtype(dt).name}")
        handles = (C.c_void_p * len(self.plans))(*[p._h.value for p in self.plans])
        ms = C.c_float()
        _check(lib().gato_solve_batched_multi(handles, len(self.plans), nsys, _f(G_val), ldG, _f(C_val), ldC, _f(g),
                                              self.n_z, _f(c), ldc, _f(lam_in) if lam_in is not None else None, ldl,
                                              _f(lam), _f(dz), _i(iters), exit_tol, int(max_iters), int(warm_start),
                                              rho, C.byref(ms)))
        return lam, dz, iters, ms.value

    def close(self):
        for p in self.plans:
            p.close()
        self.plans = []


class Plan:
    """Sparsity pattern + persistent device workspace for one (S, C, K) on one GPU."""

    def _sizes(self, S, C_, K, device):
        self.S, self.C, self.K = int(S), int(C_), int(K)
        self.n_z = (self.S + self.C) * self.K - self.C
        self.n_l = self.S * self.K
        self.g_dense = (self.S ** 2 + self.C ** 2) * self.K - self.C ** 2
        self.c_dense = (self.S ** 2 + self.S * self.C) * (self.K - 1)
        self.device = int(device)

    @classmethod
    def from_blocks(cls, S, C_, K, device=0, max_batch=1):
        """Block-input plan (no CSR): G_val / C_val of every solve are the reference's dense layouts
        (`pack_blocks`)."""
        self = cls.__new__(cls)
        self._sizes(S, C_, K, device)
        self.nnzG, self.nnzC = self.g_dense, self.c_dense
        h = C.c_void_p()
        _check(lib().gato_plan_create_blocks(self.S, self.C, self.K, self.device, int(max_batch), C.byref(h)))
        self._h = h
        return self

    def __init__(self, S, C_, K, G_row, G_col, C_row, C_col, device=0, max_batch=1):
        self._sizes(S, C_, K, device)
        gr = np.ascontiguousarray(G_row, np.int32)
        gc = np.ascontiguousarray(G_col, np.int32)
        cr = np.ascontiguousarray(C_row, np.int32)
        cc = np.ascontiguousarray(C_col, np.int32)
        if gr.size != self.n_z + 1 or cr.size != self.n_l + 1:
            raise ValueError(f"G_row/C_row must have {self.n_z + 1}/{self.n_l + 1} entries "
                             f"(got {gr.size}/{cr.size})")
        self.nnzG, self.nnzC = int(gc.size), int(cc.size)
        h = C.c_void_p()
        _check(lib().gato_plan_create(self.S, self.C, self.K, _i(gr), _i(gc), self.nnzG, _i(cr), _i(cc),
                                      self.nnzC, self.device, int(max_batch), C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().gato_plan_destroy(self._h)
            self._h = None

    __del__ = close

    def set_cluster(self, cluster_ctas):
        _check(lib().gato_plan_set_cluster(self._h, int(cluster_ctas)))

    def set_kernel(self, kernel):
        """0 = automatic, 1 = warp-per-knot cluster kernel, 2 = one-CTA-per-system kernel, 3 = one-CTA kernel over a cluster, 4 = over a cooperative grid."""
        _check(lib().gato_plan_set_kernel(self._h, int(kernel)))

    def kernel(self):
        return lib().gato_plan_get_kernel(self._h)

    def set_pipeline(self, systems_per_chunk):
        """Chunk size of the H2D / solve / D2H pipeline of solve_batched: 0 = automatic, < 0 = off."""
        _check(lib().gato_plan_set_pipeline(self._h, int(systems_per_chunk)))

    def set_staging(self, on):
        """One-CTA kernel: TMA staging of the value arrays (default on when it fits); results are identical."""
        _check(lib().gato_plan_set_staging(self._h, int(bool(on))))

    def set_zero_copy(self, on):
        """solve(): staged kernels read the pinned input buffer in place over PCIe (default on); same results."""
        _check(lib().gato_plan_set_zero_copy(self._h, int(bool(on))))

    def set_preconditioner(self, which):
        """PRECOND_STAIR (default, the reference as shipped: iteration counts match it) or PRECOND_BLOCK_JACOBI
        (the reference's SS_PRECON = 0, include/gato_defines.h:9-10: theta^-1 blocks only)."""
        which = {"stair": PRECOND_STAIR, "block_jacobi": PRECOND_BLOCK_JACOBI}.get(which, which)
        _check(lib().gato_plan_set_preconditioner(self._h, int(which)))

    def preconditioner(self):
        return int(lib().gato_plan_get_preconditioner(self._h))

    def kernel_for(self, nsys, warm_start=WARM_OFF):
        """The kernel a call with `nsys` systems runs (automatic selection picks kernel 5 for device-filling batches)."""
        return int(lib().gato_plan_get_kernel_for(self._h, int(nsys), int(warm_start)))

    def set_scaled_variant(self, variant):
        """Kernel 5, second half: 0 = by call size, 1 = latency kernel (4 lanes per knot), 2 = throughput kernel."""
        _check(lib().gato_plan_set_scaled_variant(self._h, int(variant)))

    def set_shared_matrix_path(self, on):
        """Kernel 5, batched calls with ONE matrix pair (ldG == ldC == 0): True (default) = Schur phase and factorisation
        once per call, False = per system."""
        _check(lib().gato_plan_set_shared_matrix_path(self._h, int(bool(on))))

    def set_pcg_form(self, form):
        """Kernel 5, throughput second half: 0 = two reductions per iteration (reference order), 1 = single reduction."""
        _check(lib().gato_plan_set_pcg_form(self._h, int(form)))

    def set_profiling(self, on):
        _check(lib().gato_plan_set_profiling(self._h, int(bool(on))))

    def last_profile(self):
        """Kernel 5: (first-half ms, second-half ms, launches) of the last call made with profiling on."""
        a, b, n = C.c_float(), C.c_float(), C.c_int()
        _check(lib().gato_plan_last_profile(self._h, C.byref(a), C.byref(b), C.byref(n)))
        return a.value, b.value, n.value

    def set_scaled_chunk(self, systems_per_chunk):
        """Kernel 5: systems per pass through its scratch buffer (0 = automatic)."""
        _check(lib().gato_plan_set_scaled_chunk(self._h, int(systems_per_chunk)))

    def scaled_debug(self, G_val, C_val, g, c, rho):
        """Kernel 5, first half only, one system: -> dict(psi [K, S, S] (psi_b = M_b phi_b M_{b-1}^T, block 0 zero),
        M [K, S, S], gamma_hat [K, S], region [K, REG]) decoded from the scratch layout."""
        lay = (C.c_int * 8)()
        _check(lib().gato_scaled_layout(self._h, lay))
        KSCR, REG, oPSI, PSIL, oM, oGAM, VS, HALFP = list(lay)
        S, K = self.S, self.K
        raw = np.zeros(K * KSCR, np.float32)
        G_val = _f32(G_val, (self.nnzG,)); C_val = _f32(C_val, (self.nnzC,))
        g = _f32(g, (self.n_z,)); c = _f32(c, (self.n_l,))
        _check(lib().gato_scaled_debug(self._h, _f(G_val), _f(C_val), _f(g), _f(c), float(rho), _f(raw), raw.size))
        raw = raw.reshape(K, KSCR)
        HU = S // 2
        pos = [(r // HU) * HALFP + r % HU for r in range(S)]
        M = raw[:, oM:oM + S * VS].reshape(K, S, VS)[:, :, pos]
        gam = raw[:, oGAM:oGAM + VS][:, pos]
        psi = np.zeros((K, S, S), np.float32)
        PSIB = PSIL // 2
        for gg in range(2):          # lane g: two row-major HU x HU blocks, [column half g | other half]
            for s_ in range(2):
                o = oPSI + gg * PSIL + s_ * PSIB
                hh = gg if s_ == 0 else 1 - gg
                psi[:, gg * HU:(gg + 1) * HU, hh * HU:(hh + 1) * HU] = raw[:, o:o + HU * HU].reshape(K, HU, HU)
        return dict(psi=psi, M=M, gamma_hat=gam, region=raw[:, :REG])

    def cluster(self):
        a, b = C.c_int(), C.c_int()
        _check(lib().gato_plan_get_cluster(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # -- single system, host buffers -------------------------------------------------------------
    def solve(self, G_val, C_val, g, c, lambda_in=None, exit_tol=1e-6, max_iters=500, warm_start=WARM_OFF,
              rho=0.0):
        """-> (lambda [S*K], dz [(S+C)K-C], info{iters, device_ms})"""
        G_val = _f32(G_val, (self.nnzG,)); C_val = _f32(C_val, (self.nnzC,))
        g = _f32(g, (self.n_z,)); c = _f32(c, (self.n_l,))
        lam_in = _f32(lambda_in, (self.n_l,)) if lambda_in is not None else None
        lam = np.empty(self.n_l, np.float32)
        dz = np.empty(self.n_z, np.float32)
        it, ms = C.c_int(), C.c_float()
        _check(lib().gato_solve(self._h, _f(G_val), _f(C_val), _f(g), _f(c),
                                _f(lam_in) if lam_in is not None else None, _f(lam), _f(dz),
                                exit_tol, int(max_iters), int(warm_start), rho, C.byref(it), C.byref(ms)))
        return lam, dz, dict(iters=it.value, device_ms=ms.value)

    # -- MPC sequence, host buffers ----------------------------------------------------------------
    def solve_sequence(self, G_val, C_val, g, c, rho, lambda0=None, exit_tol=1e-6, max_iters=500,
                       warm_start=WARM_TRUE):
        """`nsteps` consecutive solves; arrays are [nsteps, n] or [n] (shared); `rho` a scalar or
        [nsteps].  With WARM_TRUE step t starts from step t-1's lambda, which stays on the device.
        -> (lambda [nsteps, S*K], dz [nsteps, n_z], iters [nsteps], device_ms)"""
        g = np.ascontiguousarray(g, np.float32)
        if g.ndim != 2 or g.shape[1] != self.n_z:
            raise ValueError(f"g must be [nsteps, {self.n_z}]")
        nsteps = g.shape[0]

        def prep(a, n):
            a = np.ascontiguousarray(a, np.float32)
            if a.ndim == 1:
                a = a[None]
            if a.shape[1] != n or a.shape[0] not in (1, nsteps):
                raise ValueError(f"expected [{nsteps} or 1, {n}], got {a.shape}")
            return a, (0 if (a.shape[0] == 1 and nsteps > 1) else n)

        G_val, ldG = prep(G_val, self.nnzG)
        C_val, ldC = prep(C_val, self.nnzC)
        c, ldc = prep(c, self.n_l)
        rho = np.asarray(rho, np.float32)
        if rho.ndim == 0:
            rho_p, rho0 = None, float(rho)
        else:
            rho = np.ascontiguousarray(rho)
            if rho.shape != (nsteps,):
                raise ValueError(f"rho must be a scalar or [{nsteps}]")
            rho_p, rho0 = _f(rho), 0.0
        lam0 = _f32(lambda0, (self.n_l,)) if lambda0 is not None else None
        lam = np.empty((nsteps, self.n_l), np.float32)
        dz = np.empty((nsteps, self.n_z), np.float32)
        iters = np.empty(nsteps, np.int32)
        ms = C.c_float()
        _check(lib().gato_solve_sequence(self._h, nsteps, _f(G_val), ldG, _f(C_val), ldC, _f(g), self.n_z, _f(c),
                                         ldc, rho_p, rho0, _f(lam0) if lam0 is not None else None, _f(lam),
                                         _f(dz), _i(iters), exit_tol, int(max_iters), int(warm_start),
                                         C.byref(ms)))
        return lam, dz, iters, ms.value

    # -- batch, host buffers ---------------------------------------------------------------------
    def solve_batched(self, G_val, C_val, g, c, lambda_in=None, exit_tol=1e-6, max_iters=500,
                      warm_start=WARM_OFF, rho=0.0, out=None):
        """Arrays are [nsys, n] or [n] (shared by all systems).  -> (lambda, dz, iters, device_ms)
        `out=(lambda, dz, iters)`: preallocated float32/float32/int32 result arrays (e.g. views of
        pinned host memory, which makes the D2H copies asynchronous)."""
        g = np.ascontiguousarray(g, np.float32)
        if g.ndim != 2 or g.shape[1] != self.n_z:
            raise ValueError(f"g must be [nsys, {self.n_z}]")
        nsys = g.shape[0]

        def prep(a, n):
            a = np.ascontiguousarray(a, np.float32)
            if a.ndim == 1:
                a = a[None]
            if a.shape[1] != n or a.shape[0] not in (1, nsys):
                raise ValueError(f"expected [{nsys} or 1, {n}], got {a.shape}")
            return a, (0 if (a.shape[0] == 1 and nsys > 1) else n)

        G_val, ldG = prep(G_val, self.nnzG)
        C_val, ldC = prep(C_val, self.nnzC)
        c, ldc = prep(c, self.n_l)
        lam_in, ldl = (prep(lambda_in, self.n_l) if lambda_in is not None else (None, 0))
        if out is None:
            lam = np.empty((nsys, self.n_l), np.float32)
            dz = np.empty((nsys, self.n_z), np.float32)
            iters = np.empty(nsys, np.int32)
        else:
            lam, dz, iters = out
            for a, shp, dt in ((lam, (nsys, self.n_l), np.float32), (dz, (nsys, self.n_z), np.float32),
                               (iters, (nsys,), np.int32)):
                if a.shape != shp or a.dtype != dt or not a.flags.c_contiguous:
                    raise ValueError(f"out array must be C-contiguous {shp} {np.dtype(dt).name}")
        ms = C.c_float()
        _check(lib().gato_solve_batched(self._h, nsys, _f(G_val), ldG, _f(C_val), ldC, _f(g), self.n_z,
                                        _f(c), ldc, _f(lam_in) if lam_in is not None else None, ldl,
                                        _f(lam), _f(dz), _i(iters), exit_tol, int(max_iters),
                                        int(warm_start), rho, C.byref(ms)))
        return lam, dz, iters, ms.value

    # -- batch, device pointers (ints), asynchronous ----------------------------------------------
    def solve_batched_device(self, nsys, dG, ldG, dC, ldC, dg, ldg, dc, ldc, dlam_out, ddz_out, diters_out,
                             dlam_in=0, ldl=0, exit_tol=1e-6, max_iters=500, warm_start=WARM_OFF, rho=0.0,
                             stream=0):
        _check(lib().gato_solve_batched_device(self._h, int(nsys), dG, ldG, dC, ldC, dg, ldg, dc, ldc,
                                               dlam_in or None, ldl, dlam_out, ddz_out, diters_out, exit_tol,
                                               int(max_iters), int(warm_start), rho, stream or None))

    def noalloc_max_systems(self):
        """Systems one `solve_batched_device_noalloc` call can take (0: kernel 5 does not hold this plan's sizes)."""
        return int(lib().gato_noalloc_max_systems(self._h))

    def solve_batched_device_noalloc(self, nsys, dG, ldG, dC, ldC, dg, ldg, dc, ldc, dlam_out, ddz_out, diters_out,
                                     exit_tol=1e-6, max_iters=500, rho=0.0, stream=0):
        """`gato_solve_batched_device_noalloc`: kernel 5 with its scratch in a fixed device array; the call allocates
        nothing.  Calls that overlap on the GPU race on that array."""
        _check(lib().gato_solve_batched_device_noalloc(self._h, int(nsys), dG, ldG, dC, ldC, dg, ldg, dc, ldc, dlam_out,
                                                       ddz_out, diters_out, exit_tol, int(max_iters), rho, stream or None))

    # -- batch, any objects exposing __cuda_array_interface__ (torch, CuPy, Numba ...), zero copies --------
    def solve_batched_cuda(self, G_val, C_val, g, c, lam_out, dz_out, iters_out, lambda_in=None, exit_tol=1e-6,
                           max_iters=500, warm_start=WARM_OFF, rho=0.0, stream=0):
        """Device arrays in, device arrays out, asynchronous on `stream` (a raw cudaStream_t value).  `g` is
        [nsys, n_z]; G_val / C_val / c / lambda_in are [nsys, n] or [n] (shared by all systems); outputs are
        caller-allocated float32 [nsys, n_l], float32 [nsys, n_z], int32 [nsys].  Nothing is copied or
        synchronised: this is `gato_solve_batched_device` for callers that hold framework tensors."""
        def ptr(a, dtype, name):
            cai = getattr(a, "__cuda_array_interface__", None)
            if cai is None and (hasattr(a, "__dlpack__") or type(a).__name__ == "PyCapsule"):
                cai = _dlpack_as_cai(a, name, self.device)      # DLPack producers and raw "dltensor" capsules
            if cai is None:
                raise TypeError(f"{name} exposes neither __cuda_array_interface__ nor __dlpack__")
            if np.dtype(cai["typestr"]) != np.dtype(dtype):
                raise ValueError(f"{name} must be {np.dtype(dtype).name}, got {np.dtype(cai['typestr']).name}")
            if cai.get("strides") is not None:
                want, acc = [], np.dtype(dtype).itemsize
                for n in reversed(cai["shape"]):
                    want.insert(0, acc); acc *= n
                if tuple(cai["strides"]) != tuple(want):
                    raise ValueError(f"{name} must be C-contiguous")
            return int(cai["data"][0]), tuple(cai["shape"])

        pg, sg = ptr(g, np.float32, "g")
        if len(sg) != 2 or sg[1] != self.n_z:
            raise ValueError(f"g must be [nsys, {self.n_z}]")
        nsys = sg[0]

        def rows(a, n, name):
            pa, sa = ptr(a, np.float32, name)
            if sa == (n,) or sa == (1, n):
                return pa, (0 if nsys > 1 else n)
            if sa != (nsys, n):
                raise ValueError(f"{name} must be [{nsys}, {n}] or [{n}], got {sa}")
            return pa, n

        pG, ldG = rows(G_val, self.nnzG, "G_val")
        pC, ldC = rows(C_val, self.nnzC, "C_val")
        pc, ldc = rows(c, self.n_l, "c")
        pl_, ldl = rows(lambda_in, self.n_l, "lambda_in") if lambda_in is not None else (0, 0)
        po, so = ptr(lam_out, np.float32, "lam_out")
        pz, sz = ptr(dz_out, np.float32, "dz_out")
        pi, si = ptr(iters_out, np.int32, "iters_out")
        if so != (nsys, self.n_l) or sz != (nsys, self.n_z) or si != (nsys,):
            raise ValueError("output arrays must be [nsys, n_l], [nsys, n_z], [nsys]")
        self.solve_batched_device(nsys, pG, ldG, pC, ldC, pg, self.n_z, pc, ldc, po, pz, pi, dlam_in=pl_, ldl=ldl,
                                  exit_tol=exit_tol, max_iters=max_iters, warm_start=warm_start, rho=rho,
                                  stream=stream)

    # -- parity hooks ----------------------------------------------------------------------------
    def unpack(self, G_val, C_val, rho):
        G_val = _f32(G_val, (self.nnzG,)); C_val = _f32(C_val, (self.nnzC,))
        Gd = np.empty(self.g_dense, np.float32)
        Cd = np.empty(max(self.c_dense, 1), np.float32)
        _check(lib().gato_unpack(self._h, _f(G_val), _f(C_val), rho, _f(Gd), _f(Cd)))
        return Gd, Cd[: self.c_dense]

    def solve_debug(self, G_val, C_val, g, c, exit_tol=1e-6, max_iters=500, rho=0.0, dumps=True):
        """Solve + phase time stamps.  `dumps=True` also returns the Schur intermediates in the
        reference's layout (this forces the cluster kernel, the only one that can dump them);
        `dumps=False` profiles whichever kernel the plan selects."""
        G_val = _f32(G_val, (self.nnzG,)); C_val = _f32(C_val, (self.nnzC,))
        g = _f32(g, (self.n_z,)); c = _f32(c, (self.n_l,))
        lam = np.empty(self.n_l, np.float32); dz = np.empty(self.n_z, np.float32)
        nS = 3 * self.S * self.S * self.K
        it = C.c_int()
        clk = np.zeros(16, np.int64)
        if dumps:
            S_ = np.empty(nS, np.float32); P_ = np.empty(nS, np.float32)
            gam = np.empty(self.n_l, np.float32); Gi = np.empty(self.g_dense, np.float32)
            ptrs = (_f(S_), _f(P_), _f(gam), _f(Gi))
        else:
            S_ = P_ = gam = Gi = None
            ptrs = (None, None, None, None)
        _check(lib().gato_solve_debug(self._h, _f(G_val), _f(C_val), _f(g), _f(c), _f(lam), _f(dz), exit_tol,
                                      int(max_iters), rho, C.byref(it), *ptrs,
                                      clk.ctypes.data_as(C.POINTER(C.c_longlong))))
        return dict(lam=lam, dz=dz, iters=it.value, S=S_, Pinv=P_, gamma=gam, Ginv=Gi, clk_ns=clk)


def pack_blocks(Q, R, A, B):
    """Per-knot blocks -> the reference's dense layouts, the value arrays of a
    `Plan.from_blocks` plan.  Q [K,S,S], R [K-1,C,C]; A [K-1,S,S], B [K-1,S,C] are the blocks AS
    STORED in C's block row k+1 (for x+ = Ax + Bu that is -A, -B).  Leading batch dimensions are
    kept.  -> (G_dense [..., (S^2+C^2)K-C^2], C_dense [..., (S^2+SC)(K-1)])"""
    Q, R, A, B = (np.asarray(x, np.float32) for x in (Q, R, A, B))
    K, S = Q.shape[-3], Q.shape[-1]
    Cc = R.shape[-1] if R.size else B.shape[-1]
    lead = Q.shape[:-3]
    cm = lambda x: np.swapaxes(x, -1, -2).reshape(x.shape[:-2] + (-1,))   # noqa: E731  column-major flatten
    G = np.zeros(lead + (K, S * S + Cc * Cc), np.float32)
    G[..., :, : S * S] = cm(Q)
    if K > 1:
        G[..., : K - 1, S * S:] = cm(R)
    G = G.reshape(lead + (-1,))[..., : (S * S + Cc * Cc) * K - Cc * Cc]
    Cd = np.concatenate([cm(A), cm(B)], axis=-1).reshape(lead + (-1,)) if K > 1 else np.zeros(lead + (0,), np.float32)
    return np.ascontiguousarray(G), np.ascontiguousarray(Cd)


# ------------------------------------------------------------------------------------------------
# Host-side mirror of gpu_library.linsys_solve for runtime-chosen sizes
# ------------------------------------------------------------------------------------------------
_plan_cache: dict = {}
last_info: dict = {}


def _cached_plan(S, C_, K, G_row, G_col, C_row, C_col, device=0):
    gr = np.ascontiguousarray(G_row, np.int32); gc = np.ascontiguousarray(G_col, np.int32)
    cr = np.ascontiguousarray(C_row, np.int32); cc = np.ascontiguousarray(C_col, np.int32)
    key = (S, C_, K, device, gr.tobytes(), gc.tobytes(), cr.tobytes(), cc.tobytes())
    pl = _plan_cache.get(key)
    if pl is None:
        if len(_plan_cache) >= 8:
            _plan_cache.pop(next(iter(_plan_cache))).close()
        pl = _plan_cache[key] = Plan(S, C_, K, gr, gc, cr, cc, device=device)
    return pl


def linsys_solve(G_row, G_col, G_val, C_row, C_col, C_val, g_val, c_val, input_lambda, testiters, exit_tol,
                 max_iters, warm_start, rho, *, S, C, K, true_warm_start=False, device=0):
    """Same positional arguments, meaning and return value as the reference's
    `gpu_library.linsys_solve` (gpu_library.cu:85-87, 221-231): -> (list lambda, list dz).
    The compile-time knobs become the keyword-only S, C, K.  `warm_start=True` has no numerical
    effect unless `true_warm_start=True` (the reference ignores it, src/gato_pcg.cuh:300-304)."""
    if testiters < 1:
        raise ValueError("testiters must be >= 1")
    pl = _cached_plan(S, C, K, G_row, G_col, C_row, C_col, device)
    lam_in = np.asarray(input_lambda, np.float32)
    if lam_in.size < pl.n_l:
        raise ValueError(f"input_lambda has {lam_in.size} entries, expected {pl.n_l}")
    warm = WARM_OFF if not warm_start else (WARM_TRUE if true_warm_start else WARM_REFERENCE)
    times = []
    for _ in range(testiters):
        lam, dz, info = pl.solve(G_val, C_val, g_val, c_val, lam_in[: pl.n_l], exit_tol, max_iters, warm, rho)
        times.append(info["device_ms"])
    last_info.clear()
    last_info.update(iters=info["iters"], first_ms=times[0], avg_ms=float(np.mean(times)))
    return [float(x) for x in lam], [float(x) for x in dz]
