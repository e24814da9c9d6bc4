"""Build the UNMODIFIED reference (`/root/reference/gpu_library.cu`) into `oracle/_ref/`
(TEST INFRASTRUCTURE ONLY -- checker and baseline, never the product path).

The reference's own CMake build fails on Python 3.12 because of its vendored pybind11 2.3.dev0
(SURVEY.md 8c), so its single translation unit is compiled directly with the image's pybind11
headers; no reference source is copied or modified.  One module per compile-time configuration:
    oracle/_ref/S{S}_C{C}_K{K}/gpu_library<ext>.so          (-arch=sm_100)
    oracle/_ref/S{S}_C{C}_K{K}/libref_harness.so            (oracle/ref_harness.cu: the reference's src/*.cuh
        #included unchanged, driven on pre-allocated buffers: dumps of its intermediates + kernel-only timing)
`oracle/_ref/` is git-ignored but travels to the GPU box.  /root/reference exists only in the
authoring container: on the GPU box the prebuilt modules are used as they are.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
OUT = os.path.join(HERE, "_ref")
CONFIGS = [(2, 1, 5), (14, 7, 50), (14, 7, 256)]


def ref_dir(S, C, K):
    return os.path.join(OUT, f"S{S}_C{C}_K{K}")


def build_ref(S, C, K, force=False):
    import pybind11

    d = ref_dir(S, C, K)
    target = os.path.join(d, "gpu_library" + sysconfig.get_config_var("EXT_SUFFIX"))
    if os.path.exists(target) and not force:
        return d
    src = os.path.join(REF, "gpu_library.cu")
    if not os.path.exists(src):
        return None   # GPU box: nothing to build from
    os.makedirs(d, exist_ok=True)
    cmd = ["/usr/local/cuda/bin/nvcc", "-O3", "-std=c++17", "-arch=sm_100", "-Xcompiler", "-fPIC", "-shared",
           "-w", f"-DKNOT_POINTS={K}", f"-DSTATE_SIZE={S}", f"-DCONTROL_SIZE={C}",
           "-I" + sysconfig.get_paths()["include"], "-I" + pybind11.get_include(), src, "-o", target]
    subprocess.run(cmd, check=True, cwd=REF)
    return d


def build_harness(S, C, K, force=False):
    """oracle/ref_harness.cu against the unmodified reference headers (-I/root/reference); C ABI, loaded with ctypes
    by tests/refharness.py."""
    d = ref_dir(S, C, K)
    target = os.path.join(d, "libref_harness.so")
    src = os.path.join(HERE, "ref_harness.cu")
    if os.path.exists(target) and not force and os.path.getmtime(target) >= os.path.getmtime(src):
        return target
    if not os.path.exists(os.path.join(REF, "src", "gato_pcg.cuh")):
        return target if os.path.exists(target) else None   # GPU box: nothing to build from
    os.makedirs(d, exist_ok=True)
    cmd = ["/usr/local/cuda/bin/nvcc", "-O3", "-std=c++17", "-arch=sm_100", "-lineinfo", "-Xcompiler", "-fPIC", "-shared",
           "-w", f"-DKNOT_POINTS={K}", f"-DSTATE_SIZE={S}", f"-DCONTROL_SIZE={C}", "-I" + REF, src, "-o", target]
    subprocess.run(cmd, check=True)
    return target


def build_all(force=False):
    out = [build_ref(S, C, K, force) for (S, C, K) in CONFIGS]
    for (S, C, K) in CONFIGS:
        build_harness(S, C, K, force)
    return out


if __name__ == "__main__":
    for d in build_all("--force" in sys.argv):
        print(d)
