"""CPU tests of the N>1 path: contiguous batch sharding + final gather over `gloo`, world_size 2.
The solve itself is replaced by the oracle here (checker in a test; the product has no CPU path)."""
import os
import socket
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from gato_python_b200.sharding import shard_range  # noqa: E402


def test_shard_range_partitions():
    for nsys in (0, 1, 7, 8, 8192, 8191):
        for world in (1, 2, 3, 4, 8):
            r = [shard_range(nsys, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == nsys
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _worker(rank, world, port, q):
    import torch.distributed as dist

    from gato_python_b200.problems import iiwa_like_batch
    from gato_python_b200.sharding import solve_sharded
    from oracle import gato_oracle as orc

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b = iiwa_like_batch(5, S=4, C=2, K=6, seed0=10)
    calls = []

    def solve_fn(ch):
        n = ch["g"].shape[0]
        calls.append(n)
        out = [orc.linsys_solve_info(4, 2, 6, b["G_row"], b["G_col"], ch["G_val"][i], b["C_row"], b["C_col"],
                                     ch["C_val"][i], ch["g"][i], ch["c"][i], np.zeros(24), 1, 1e-6, 100, False, 1e-3)
               for i in range(n)]
        return (np.stack([o[0] for o in out]), np.stack([o[1] for o in out]),
                np.array([o[2]["iters"] for o in out], np.int32))

    res = solve_sharded(solve_fn, 5, dict(G_val=b["G_val"], C_val=b["C_val"], g=b["g"], c=b["c"]))
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, calls, None if res is None else [np.asarray(x) for x in res]))


def test_gloo_world2_shard_and_gather():
    import torch.multiprocessing as mp

    from gato_python_b200.problems import iiwa_like_batch
    from oracle import gato_oracle as orc

    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(2):
        rank, calls, res = q.get(timeout=240)
        got[rank] = (calls, res)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert got[0][0] == [3] and got[1][0] == [2]     # 5 systems -> chunks of 3 and 2
    assert got[1][1] is None
    lam, dz, iters = got[0][1]
    b = iiwa_like_batch(5, S=4, C=2, K=6, seed0=10)
    assert lam.shape == (5, 24) and dz.shape == (5, 34) and iters.shape == (5,)
    for i in range(5):
        lo, do, io = orc.linsys_solve_info(4, 2, 6, b["G_row"], b["G_col"], b["G_val"][i], b["C_row"], b["C_col"],
                                           b["C_val"][i], b["g"][i], b["c"][i], np.zeros(24), 1, 1e-6, 100, False, 1e-3)
        assert np.array_equal(lam[i], lo) and np.array_equal(dz[i], do) and iters[i] == io["iters"]


def test_numa_binding_helpers():
    from gato_python_b200.sharding import _parse_cpulist, bind_to_gpu_numa
    assert _parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert _parse_cpulist("") == set()
    import os
    before = os.sched_getaffinity(0)
    r = bind_to_gpu_numa(0)          # no GPU / unknown topology: nothing changes
    assert r is None or (r["cpus"] >= 1 and os.sched_getaffinity(0) <= before)
    os.sched_setaffinity(0, before)
