"""CPU tests of the C-ABI boundary: the library loads, exports every symbol include/gato_b200.h
declares, validates patterns on the host, and fails loudly (no CPU fallback) without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

from gato_python_b200 import solver
from gato_python_b200.problems import iiwa_like, pendulum5

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "gato_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gato_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(native_lib):
    names = header_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(native_lib, n), f"{n} declared in include/gato_b200.h but not exported"
    assert sorted(solver.SYMBOLS) == names


def test_supported_sizes(native_lib):
    sizes = solver.supported_sizes()
    assert (2, 1) in sizes and (14, 7) in sizes


def _create(S, C, K, prob):
    h = ctypes.c_void_p()
    ip = ctypes.POINTER(ctypes.c_int)
    arrs = [np.ascontiguousarray(prob[k], np.int32) for k in ("G_row", "G_col", "C_row", "C_col")]
    rc = solver.lib().gato_plan_create(S, C, K, arrs[0].ctypes.data_as(ip), arrs[1].ctypes.data_as(ip),
                                       arrs[1].size, arrs[2].ctypes.data_as(ip), arrs[3].ctypes.data_as(ip),
                                       arrs[3].size, 0, 1, ctypes.byref(h))
    return rc, solver.lib().gato_last_error().decode(), h


def test_pattern_validation_happens_on_host(native_lib):
    """Entries the reference would scatter outside their block are rejected."""
    p = pendulum5()
    bad = dict(p, G_col=p["G_col"].copy())
    bad["G_col"][0] = 2            # state row 0 coupled to the control column of knot 0
    rc, msg, _ = _create(2, 1, 5, bad)
    assert rc == -2 and "block-diagonal" in msg
    bad = dict(p, C_col=p["C_col"].copy())
    bad["C_col"][-1] = 99
    rc, msg, _ = _create(2, 1, 5, bad)
    assert rc == -2 and "out of range" in msg
    bad = dict(p, G_row=p["G_row"].copy())
    bad["G_row"][-1] = 13
    rc, msg, _ = _create(2, 1, 5, bad)
    assert rc == -1
    q = iiwa_like(S=3, C=2, K=4)
    rc, msg, _ = _create(3, 2, 4, q)
    assert rc == -3 and "no kernels" in msg


def test_wrong_lengths_raise_valueerror(native_lib):
    p = pendulum5()
    with pytest.raises(ValueError):
        solver.Plan(2, 1, 6, p["G_row"], p["G_col"], p["C_row"], p["C_col"])


@pytest.mark.skipif(solver.device_count() > 0, reason="CUDA device present")
def test_no_cpu_fallback(native_lib):
    """Without a device every compute entry point fails loudly."""
    p = pendulum5()
    rc, msg, _ = _create(2, 1, 5, p)
    assert rc == -4, msg
    with pytest.raises(solver.GatoError):
        solver.linsys_solve(p["G_row"], p["G_col"], p["G_val"], p["C_row"], p["C_col"], p["C_val"], p["g"],
                            p["c"], np.zeros(10), 1, 1e-6, 10, False, 1e-3, S=2, C=1, K=5)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "gato_python_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f
                assert "gato_oracle" not in text, f


def test_dropin_module_builds_and_has_reference_surface():
    """The pybind11 module keeps the reference's name, function and positional-only signature
    (gpu_library.cu:236-239).  Import only: no compute without a GPU."""
    import subprocess
    import sys

    from gato_python_b200 import build
    d = build.build_module(2, 1, 5)
    code = ("import sys; sys.path.insert(0, %r); import gpu_library as g; "
            "assert (g.STATE_SIZE, g.CONTROL_SIZE, g.KNOT_POINTS) == (2, 1, 5); "
            "assert 'arg0' in g.linsys_solve.__doc__ and 'arg13' in g.linsys_solve.__doc__; "
            "import pytest\n"
            "try:\n    g.linsys_solve([0],[0],[0.],[0],[0],[0.],[0.],[0.],[0.],1,1e-6,1,False,0.)\n"
            "except ValueError as e: print('OK', e)\n") % d
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK" in r.stdout, r.stdout + r.stderr
