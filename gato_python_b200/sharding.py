"""Batch sharding over GPUs: independent KKT systems are split into contiguous
chunks, one process per GPU; there is NO collective on the solve path, only a final gather of
(lambda, dz, iters) to rank 0.  `torch.distributed` is plumbing (rendezvous, barrier, gather)."""
from __future__ import annotations

import numpy as np


def shard_range(nsys: int, rank: int, world: int):
    """Contiguous chunk [lo, hi) of rank `rank`: sizes differ by at most one, earlier ranks larger."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} of {world}")
    base, rem = divmod(nsys, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def solve_sharded(solve_fn, nsys, arrays, group=None, dst=0):
    """Run `solve_fn(chunk_arrays) -> (lam [n, n_l], dz [n, n_z], iters [n])` on this rank's chunk of
    every per-system array in `arrays` (dict name -> [nsys, *] or shared 1-D array) and gather the
    results on rank `dst`.  Returns (lam, dz, iters) on dst, None elsewhere.  Works without
    torch.distributed being initialised (world size 1)."""
    import torch
    import torch.distributed as dist

    on = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if on else 0
    world = dist.get_world_size(group) if on else 1
    lo, hi = shard_range(nsys, rank, world)
    chunk = {k: (v[lo:hi] if getattr(v, "ndim", 1) == 2 and v.shape[0] == nsys else v) for k, v in arrays.items()}
    if hi > lo:
        lam, dz, iters = solve_fn(chunk)
    else:
        lam = dz = iters = None
    if not on or world == 1:
        return lam, dz, iters
    # final host gather: ragged chunks -> gather_object of NumPy arrays (a few KB per system)
    out = [None] * world if rank == dst else None
    dist.gather_object((lo, hi, lam, dz, iters), out, dst=dst, group=group)
    if rank != dst:
        return None
    parts = sorted((p for p in out if p[1] > p[0]), key=lambda p: p[0])
    assert parts[0][0] == 0 and parts[-1][1] == nsys and all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
    return (np.concatenate([p[2] for p in parts]), np.concatenate([p[3] for p in parts]),
            np.concatenate([p[4] for p in parts]))


# ------------------------------------------------------------------------------------------------
# Host placement: one process per GPU should run, and pin its staging memory, on the GPU's own NUMA node
# ------------------------------------------------------------------------------------------------
def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def gpu_numa_node(index: int):
    """NUMA node of CUDA device `index` from its PCI address (sysfs), or None when unknown."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.lower().split(":", 1)
        with open(f"/sys/bus/pci/devices/{dom[-4:]}:{rest}/numa_node") as f:
            node = int(f.read().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def bind_to_gpu_numa(index: int):
    """Restrict this process to the CPUs of GPU `index`'s NUMA node, so that the pinned host buffers it
    allocates afterwards are first-touched there and its H2D/D2H copies do not cross the socket link.
    Returns {"node", "cpus"} or None when the topology is unknown (nothing is changed then)."""
    import os
    node = gpu_numa_node(index)
    if node is None:
        return None
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read()) & os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"node": node, "cpus": len(cpus)}
    except Exception:
        return None
