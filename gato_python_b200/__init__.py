"""gato_python_b200 -- B200-native (sm_100a) KKT linear-system hot path of GATO.

Product code: `csrc/` (CUDA kernels + C ABI, include/gato_b200.h), `solver.py` (ctypes binding and
the host-side mirror of `gpu_library.linsys_solve`), `build.py` (in-tree nvcc build),
`problems.py` (synthetic KKT systems), `sharding.py` (batch sharding over GPUs).
"""
from .solver import (GatoError, MultiPlan, Plan, PRECOND_BLOCK_JACOBI, PRECOND_STAIR, WARM_OFF, WARM_REFERENCE,  # noqa: F401
                     WARM_TRUE, device_count, linsys_solve, supported_sizes)
