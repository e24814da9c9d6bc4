"""SciPy CPU baselines of BASELINE.md section 2.2 (TEST INFRASTRUCTURE / REPORTED BASELINE ONLY -- never on the product path).

The reference ships no CPU path; these are the algorithm-matched (C1) and direct (C2) CPU solves of the identical KKT
system, and their batched form over the host cores (C3):

* C1 `schur_cg`   -- form the sparse S = C (G + rho I_stored_diag)^-1 C^T and gamma = C G^-1 g - c (c[0:S] ignored, as the
                     reference does, src/gato_schur.cuh:110-147), run `scipy.sparse.linalg.cg` with the block-tridiagonal
                     stair preconditioner D^-1 - D^-1 O D^-1 (src/gato_schur.cuh:498-649) as a `LinearOperator` and the
                     reference's exit rule |r^T P^-1 r| < exit_tol (src/gato_pcg.cuh:404) through the callback, then
                     dz = G^-1 (g - C^T lambda) (src/gato_schur.cuh:759-867).  FP32 or FP64.
* C2 `kkt_spsolve` -- `scipy.sparse.linalg.spsolve` of [[G + rho I, C^T], [C, 0]] [dz; lambda] = [g; c~] in FP64.
* C3 `batched`     -- `multiprocessing.Pool` over independent systems, BLAS threads pinned to 1.
"""
import os
import time

import numpy as np


class _Stop(Exception):
    pass


def _csr(n_rows, n_cols, row, col, val, dtype):
    from scipy import sparse
    return sparse.csr_matrix((np.asarray(val, dtype), np.asarray(col), np.asarray(row)), shape=(n_rows, n_cols))


def _g_with_rho(S, C, K, G_row, G_col, G_val, rho, dtype):
    n_z = (S + C) * K - C
    G = _csr(n_z, n_z, G_row, G_col, G_val, dtype).tolil()
    rows = np.repeat(np.arange(n_z), np.diff(np.asarray(G_row)))
    stored_diag = rows[np.asarray(G_col)[: len(rows)] == rows]
    for d in stored_diag:                       # rho only where the diagonal entry is stored
        G[d, d] += dtype(rho)
    return G.tocsc()


def schur_cg(S, C, K, G_row, G_col, G_val, C_row, C_col, C_val, g, c, exit_tol, max_iters, rho, dtype=np.float32):
    """C1.  -> (lambda, dz, iters)"""
    from scipy import sparse
    from scipy.sparse.linalg import LinearOperator, cg, splu

    dt = np.dtype(dtype).type
    n_z, n_l = (S + C) * K - C, S * K
    G = _g_with_rho(S, C, K, G_row, G_col, G_val, rho, dt)
    Cm = _csr(n_l, n_z, C_row, C_col, C_val, dt)
    lu = splu(G.astype(np.float64))            # G is block diagonal: its LU is block diagonal too
    Ginv_Ct = sparse.csc_matrix(lu.solve(Cm.T.toarray().astype(np.float64)).astype(dt))
    Sm = (Cm @ Ginv_Ct).tocsr().astype(dt)
    gv = np.asarray(g, dt)
    cc = np.asarray(c, dt).copy()
    cc[:S] = 0
    gamma = (Cm @ lu.solve(gv.astype(np.float64)).astype(dt) - cc).astype(dt)
    # block-tridiagonal stair preconditioner from the S x S blocks of Sm
    Sd = Sm.toarray().reshape(K, S, K, S)
    Dinv = np.stack([np.linalg.inv(Sd[b, :, b, :].astype(np.float64)).astype(dt) for b in range(K)])
    Off = np.stack([Sd[b, :, b - 1, :] for b in range(1, K)]) if K > 1 else np.zeros((0, S, S), dt)

    def stair(r):
        r = np.asarray(r, dt).reshape(K, S)
        t = np.einsum("bij,bj->bi", Dinv, r)
        o = np.zeros_like(t)
        if K > 1:
            o[1:] += np.einsum("bij,bj->bi", Off, t[:-1])
            o[:-1] += np.einsum("bji,bj->bi", Off, t[1:])
        return np.einsum("bij,bj->bi", Dinv, r - o).reshape(-1).astype(dt)

    P = LinearOperator((n_l, n_l), matvec=stair, dtype=dt)
    state = {"it": 0, "lam": None}

    def cb(xk):
        state["it"] += 1
        r = gamma - Sm @ xk
        if abs(float(r @ stair(r))) < exit_tol:
            state["lam"] = xk.copy()
            raise _Stop

    try:
        lam, _ = cg(Sm, gamma, x0=np.zeros(n_l, dt), rtol=0.0, atol=0.0, maxiter=max_iters, M=P, callback=cb)
    except _Stop:
        lam = state["lam"]
    dz = lu.solve((gv - Cm.T @ lam).astype(np.float64)).astype(dt)
    return np.asarray(lam, dt), dz, state["it"] - 1      # the reference counts the exit iteration 0-based


def kkt_spsolve(S, C, K, G_row, G_col, G_val, C_row, C_col, C_val, g, c, rho):
    """C2 (FP64).  -> (lambda, dz)"""
    from scipy import sparse
    from scipy.sparse.linalg import spsolve

    n_z, n_l = (S + C) * K - C, S * K
    G = _g_with_rho(S, C, K, G_row, G_col, G_val, rho, np.float64)
    Cm = _csr(n_l, n_z, C_row, C_col, C_val, np.float64)
    cc = np.asarray(c, np.float64).copy()
    cc[:S] = 0
    KKT = sparse.bmat([[G, Cm.T], [Cm, None]], format="csc")
    x = spsolve(KKT, np.concatenate([np.asarray(g, np.float64), cc]))
    return x[n_z:], x[:n_z]


def _worker(job):
    which, first, n, seed = job
    os.environ["OMP_NUM_THREADS"] = "1"
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    from gato_python_b200.problems import iiwa_like_batch_fast

    b = iiwa_like_batch_fast(n, 14, 7, 50, seed=seed, first=first)
    t0 = time.perf_counter()
    its = []
    for i in range(n):
        a = (b["S"], b["C"], b["K"], b["G_row"], b["G_col"], b["G_val"][i], b["C_row"], b["C_col"], b["C_val"][i],
             b["g"][i], b["c"][i])
        if which == "schur_cg":
            its.append(schur_cg(*a, b["exit_tol"], b["max_iters"], b["rho"])[2])
        else:
            kkt_spsolve(*a, b["rho"])
    return time.perf_counter() - t0, its


def batched(which, per_core, cores=None, seed=1000):
    """C3: `per_core` IIWA-50 systems of the benchmark generator per process.  -> dict(value solves/s, cores, ...)"""
    import multiprocessing as mp

    cores = cores or os.cpu_count() or 1
    with mp.get_context("spawn").Pool(cores) as pool:
        res = pool.map(_worker, [(which, i * per_core, per_core, seed) for i in range(cores)])
    busy = max(r[0] for r in res)
    its = [i for r in res for i in r[1]]
    return {"value": per_core * cores / busy, "unit": "solves/s", "cores": cores, "solver": which,
            "systems": per_core * cores, "slowest_worker_s": busy, "mean_iters": float(np.mean(its)) if its else None}
