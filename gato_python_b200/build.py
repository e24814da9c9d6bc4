"""In-tree build of the native pieces (nvcc, sm_100a only).

* `libgato_b200.so`  -- the C-ABI library (include/gato_b200.h), kernels for (S,C) = (2,1), (14,7)
  plus one optional extra pair.
* `build/S{S}_C{C}_K{K}/gpu_library<ext>.so` -- the pybind11 drop-in module with the reference's
  compile-time knobs (CMakeLists.txt:18: -DSTATE_SIZE -DCONTROL_SIZE -DKNOT_POINTS).

Command line:  python -m gato_python_b200.build [--module S C K]
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libgato_b200.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _lib_sources():
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))
            if f.endswith((".cu", ".cuh", ".inc", ".h")) and f not in ("floors.cu", "gpu_library.cu", "tf32_experiment.cu")]
    srcs.append(os.path.join(ROOT, "include", "gato_b200.h"))
    return srcs


#: (STATE_SIZE, CONTROL_SIZE) pairs every build of the library carries (csrc/gato_capi.cu, GATO_FOR_SIZES)
BUILTIN_SIZES = [(2, 1), (4, 2), (7, 3), (14, 7)]


def build_dev(defines=(), out=None, verbose=False):
    """Development build: (14, 7) kernels only, extra -D defines; load it with GATO_B200_LIB=<path>."""
    target = out or os.path.join(ROOT, "build", "dev", "libgato_b200.so")
    os.makedirs(os.path.dirname(target), exist_ok=True)
    cmd = [_nvcc()] + ARCH + COMMON + ["-DGATO_ONLY_IIWA"] + [f"-D{d}" for d in defines]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    subprocess.run(cmd + [os.path.join(CSRC, "gato_capi.cu"), "-o", target], check=True)
    return target


def build_library(force=False, extra_sc=None, verbose=False, out=None):
    """Compile libgato_b200.so (or, with `out`, a second copy elsewhere -- e.g. one that carries an extra
    (STATE_SIZE, CONTROL_SIZE) pair in every kernel family); returns its path."""
    srcs = _lib_sources()
    target = out or LIB
    if not force and not _stale(target, srcs):
        return target
    os.makedirs(os.path.dirname(target), exist_ok=True)
    cmd = [_nvcc()] + ARCH + COMMON
    if extra_sc:
        cmd += [f"-DGATO_EXTRA_S={int(extra_sc[0])}", f"-DGATO_EXTRA_C={int(extra_sc[1])}"]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [os.path.join(CSRC, "gato_capi.cu"), "-o", target]
    subprocess.run(cmd, check=True)
    return target


def module_dir(S, C, K):
    return os.path.join(ROOT, "build", f"S{S}_C{C}_K{K}")


def build_module(S, C, K, force=False):
    """Compile the pybind11 module `gpu_library` for one (STATE_SIZE, CONTROL_SIZE, KNOT_POINTS);
    returns the directory to put on sys.path (the reference uses PYTHONPATH=$PWD/build,
    install.bash:20)."""
    import pybind11

    out_dir = module_dir(S, C, K)
    os.makedirs(out_dir, exist_ok=True)
    if (S, C) in BUILTIN_SIZES:
        lib, lib_dir, rpath = build_library(), HERE, "$ORIGIN/../../gato_python_b200"
    else:   # any other STATE_SIZE / CONTROL_SIZE: the module gets its own copy of the library with that pair compiled in
        lib = build_library(extra_sc=(S, C), out=os.path.join(out_dir, "libgato_b200.so"))
        lib_dir, rpath = out_dir, "$ORIGIN"
    target = os.path.join(out_dir, "gpu_library" + sysconfig.get_config_var("EXT_SUFFIX"))
    src = os.path.join(CSRC, "gpu_library.cu")
    if not force and not _stale(target, [src, lib]):
        return out_dir
    cmd = [_nvcc()] + ARCH + COMMON + [
        f"-DSTATE_SIZE={S}", f"-DCONTROL_SIZE={C}", f"-DKNOT_POINTS={K}",
        "-I" + sysconfig.get_paths()["include"], "-I" + pybind11.get_include(),
        src, "-o", target, "-L" + lib_dir, "-lgato_b200",
        "-Xlinker", "-rpath", "-Xlinker", rpath,
    ]
    subprocess.run(cmd, check=True)
    return out_dir


def build_floors(force=False):
    """build/floors: micro-benchmarks of the launch / barrier floors (csrc/floors.cu)."""
    src = os.path.join(CSRC, "floors.cu")
    out = os.path.join(ROOT, "build", "floors")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    if force or _stale(out, [src]):
        subprocess.run([_nvcc()] + ARCH + ["-O3", "-std=c++17", "-lineinfo", "-rdc=true", src, "-o", out], check=True)
    return out


def build_tf32_experiment(force=False):
    """build/tf32_experiment: the tensor-core option for the Schur GEMMs, timed and error-checked (csrc/tf32_experiment.cu)."""
    src = os.path.join(CSRC, "tf32_experiment.cu")
    out = os.path.join(ROOT, "build", "tf32_experiment")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    if force or _stale(out, [src]):
        subprocess.run([_nvcc()] + ARCH + ["-O3", "-std=c++17", "-lineinfo", src, "-o", out], check=True)
    return out


DEFAULT_MODULES = [(2, 1, 5), (14, 7, 50), (14, 7, 256)]


def build_all(force=False):
    build_library(force=force)
    build_floors(force=force)
    build_tf32_experiment(force=force)
    return [build_module(S, C, K, force=force) for (S, C, K) in DEFAULT_MODULES]


if __name__ == "__main__":
    if len(sys.argv) >= 2 and sys.argv[1] == "--dev":
        print(build_dev(defines=[a for a in sys.argv[2:] if "=" in a or a.isidentifier()]))
    elif len(sys.argv) >= 5 and sys.argv[1] == "--module":
        print(build_module(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), force=True))
    else:
        for d in build_all(force="--force" in sys.argv):
            print(d)
