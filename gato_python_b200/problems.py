"""Synthetic KKT systems for tests and benchmarks.

* `pendulum5()`  -- the literal arrays of the reference's only known-answer test
  (`/root/reference/test_pendulum_5.py:9-24`, same data as `README.md:40-55`).
* `iiwa_like(S, C, K, seed)` -- IIWA-shaped generator: block-diagonal cost G (diag or dense SPD
  blocks) and block-bidiagonal dynamics Jacobian C = [I; -A_k | -B_k | I], CSR with a STRUCTURAL
  sparsity pattern that does not depend on the seed, so a batch of systems shares one pattern.
* `iiwa_like_batch(...)` -- value arrays [nsys, nnz] over that shared pattern.

Only NumPy is used here (no SciPy) so it runs anywhere the product runs.
"""
from __future__ import annotations

import numpy as np


def pendulum5():
    G_row = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14]
    G_col = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13]
    G_val = [1., 1., 0.1, 1., 1., 0.1, 1., 1., 0.1, 1., 1., 0.1, 100., 100.]
    C_row = [0, 1, 2, 5, 9, 12, 16, 19, 23, 26, 30]
    C_col = [0, 1, 0, 1, 3, 0, 1, 2, 4, 3, 4, 6, 3, 4, 5, 7, 6, 7, 9, 6, 7, 8, 10, 9, 10, 12, 9, 10,
             11, 13]
    C_val = [1., 1., -1., -0.1, 1., 0.981, -1., -0.1, 1., -1., -0.1, 1., 0.981, -1., -0.1, 1., -1.,
             -0.1, 1., 0.981, -1., -0.1, 1., -1., -0.1, 1., 0.981, -1., -0.1, 1.]
    g_val = [-3.1416, 0., 0., -3.1416, 0., 0., -3.1416, 0., 0., -3.1416, 0., 0., -314.159, 0.]
    c_val = [0] * 10
    return dict(S=2, C=1, K=5,
                G_row=np.array(G_row, np.int32), G_col=np.array(G_col, np.int32),
                G_val=np.array(G_val, np.float32),
                C_row=np.array(C_row, np.int32), C_col=np.array(C_col, np.int32),
                C_val=np.array(C_val, np.float32),
                g=np.array(g_val, np.float32), c=np.array(c_val, np.float32),
                exit_tol=1e-6, max_iters=10, rho=1e-3)


# ------------------------------------------------------------------------------------------------
def _pattern(S, C, K, dense_G, dense_AB):
    """Structural CSR pattern (row_ptr, col, and for each nz a (kind, k, r, c) source tag)."""
    SC = S + C
    h = S // 2
    # which entries of A (S x S) / B (S x C) are structurally non-zero
    if dense_AB:
        Amask = np.ones((S, S), bool)
        Bmask = np.ones((S, C), bool)
    else:
        Amask = np.eye(S, dtype=bool)
        Amask[:h, h:2 * h] |= np.eye(h, dtype=bool)   # position += dt * velocity
        Amask[h:, :] = True                            # velocity rows dense
        Bmask = np.zeros((S, C), bool)
        Bmask[h:, :] = True
    # ---- G
    g_rows, g_cols, g_src = [], [], []
    for k in range(K):
        for r in range(S):
            cols = range(S) if dense_G else [r]
            for c in cols:
                g_rows.append(k * SC + r); g_cols.append(k * SC + c); g_src.append((0, k, r, c))
        if k < K - 1:
            for r in range(C):
                cols = range(C) if dense_G else [r]
                for c in cols:
                    g_rows.append(k * SC + S + r); g_cols.append(k * SC + S + c)
                    g_src.append((1, k, r, c))
    # ---- C
    c_rows, c_cols, c_src = [], [], []
    for r in range(S):
        c_rows.append(r); c_cols.append(r); c_src.append((4, 0, r, r))          # +I (x0 row)
    for k in range(K - 1):
        for r in range(S):
            row = (k + 1) * S + r
            for c in range(S):
                if Amask[r, c]:
                    c_rows.append(row); c_cols.append(k * SC + c); c_src.append((2, k, r, c))
            for c in range(C):
                if Bmask[r, c]:
                    c_rows.append(row); c_cols.append(k * SC + S + c); c_src.append((3, k, r, c))
            c_rows.append(row); c_cols.append((k + 1) * SC + r); c_src.append((4, k + 1, r, r))

    def csr(rows, cols, n_rows):
        rows = np.asarray(rows); cols = np.asarray(cols)
        # rows are already sorted, columns ascending within a row by construction
        rp = np.zeros(n_rows + 1, np.int32)
        np.add.at(rp, rows + 1, 1)
        return np.cumsum(rp).astype(np.int32), cols.astype(np.int32)

    G_row, G_col = csr(g_rows, g_cols, SC * K - C)
    C_row, C_col = csr(c_rows, c_cols, S * K)
    return G_row, G_col, np.array(g_src, np.int32), C_row, C_col, np.array(c_src, np.int32), Amask, Bmask


_PRESETS = {
    # (q_pos, q_vel, r, q_N)
    "stiff": (10.0, 0.1, 1e-4, 100.0),
    "mild": (10.0, 1.0, 0.01, 10.0),
}


def _blocks(S, C, K, rng, preset, dt, dense_G, Amask, Bmask):
    h = S // 2
    q_pos, q_vel, rr, q_N = _PRESETS[preset]
    qd = np.array([q_pos] * h + [q_vel] * (S - h))
    Q = np.tile(np.diag(qd), (K, 1, 1))
    Q[K - 1] *= q_N
    R = np.tile(rr * np.eye(C), (K - 1, 1, 1))
    if dense_G:
        M = 0.3 * rng.standard_normal((K, S, S))
        Q = Q + M @ M.transpose(0, 2, 1)
        M = 0.03 * rng.standard_normal((K - 1, C, C))
        R = R + M @ M.transpose(0, 2, 1)
    A = np.tile(np.eye(S), (K - 1, 1, 1))
    A[:, :h, h:2 * h] += dt * np.eye(h)
    A[:, h:, :] += dt * rng.standard_normal((K - 1, S - h, S))
    B = np.zeros((K - 1, S, C))
    eyeB = np.zeros((S - h, C)); n = min(S - h, C); eyeB[:n, :n] = np.eye(n)
    B[:, h:, :] = dt * (2.0 * eyeB + 0.3 * rng.standard_normal((K - 1, S - h, C)))
    if Amask.all():
        A[:, :h, :] += dt * 0.1 * rng.standard_normal((K - 1, h, S))
        B[:, :h, :] += dt * 0.1 * rng.standard_normal((K - 1, h, C))
    g = rng.standard_normal((S + C) * K - C)
    c = 0.01 * rng.standard_normal(S * K)
    c[:S] = 0.0
    return Q, R, A, B, g, c


def _values(g_src, c_src, Q, R, A, B):
    gv = np.empty(len(g_src), np.float32)
    kind, k, r, c = g_src.T
    m = kind == 0; gv[m] = Q[k[m], r[m], c[m]]
    m = kind == 1; gv[m] = R[k[m], r[m], c[m]]
    cv = np.empty(len(c_src), np.float32)
    kind, k, r, c = c_src.T
    m = kind == 2; cv[m] = -A[k[m], r[m], c[m]]
    m = kind == 3; cv[m] = -B[k[m], r[m], c[m]]
    m = kind == 4; cv[m] = 1.0
    return gv, cv


def iiwa_like(S=14, C=7, K=50, seed=0, preset="stiff", dt=0.01, dense_G=False, dense_AB=False,
              rho=1e-3, exit_tol=1e-6, max_iters=500):
    G_row, G_col, g_src, C_row, C_col, c_src, Amask, Bmask = _pattern(S, C, K, dense_G, dense_AB)
    rng = np.random.default_rng(seed)
    Q, R, A, B, g, c = _blocks(S, C, K, rng, preset, dt, dense_G, Amask, Bmask)
    gv, cv = _values(g_src, c_src, Q, R, A, B)
    return dict(S=S, C=C, K=K, G_row=G_row, G_col=G_col, G_val=gv, C_row=C_row, C_col=C_col,
                C_val=cv, g=g.astype(np.float32), c=c.astype(np.float32),
                exit_tol=exit_tol, max_iters=max_iters, rho=rho)


def iiwa_like_batch(nsys, S=14, C=7, K=50, seed0=1000, preset="stiff", dt=0.01, dense_G=False,
                    dense_AB=False, rho=1e-3, exit_tol=1e-6, max_iters=500, shared_matrix=False):
    """Batch over one shared sparsity pattern.  Returns value arrays of shape [nsys, *].
    `shared_matrix=True`: only g and c vary between systems (config 4's 'randomised g/c')."""
    G_row, G_col, g_src, C_row, C_col, c_src, Amask, Bmask = _pattern(S, C, K, dense_G, dense_AB)
    GV = np.empty((nsys, len(G_col)), np.float32)
    CV = np.empty((nsys, len(C_col)), np.float32)
    gg = np.empty((nsys, (S + C) * K - C), np.float32)
    cc = np.empty((nsys, S * K), np.float32)
    base = None
    for i in range(nsys):
        rng = np.random.default_rng(seed0 + i)
        Q, R, A, B, g, c = _blocks(S, C, K, rng, preset, dt, dense_G, Amask, Bmask)
        if shared_matrix:
            if base is None:
                base = _values(g_src, c_src, Q, R, A, B)
            GV[i], CV[i] = base
        else:
            GV[i], CV[i] = _values(g_src, c_src, Q, R, A, B)
        gg[i], cc[i] = g, c
    return dict(S=S, C=C, K=K, nsys=nsys, G_row=G_row, G_col=G_col, G_val=GV, C_row=C_row,
                C_col=C_col, C_val=CV, g=gg, c=cc, exit_tol=exit_tol, max_iters=max_iters, rho=rho)


def iiwa_like_batch_fast(nsys, S=14, C=7, K=50, seed=1000, preset="stiff", dt=0.01, rho=1e-3, exit_tol=1e-6,
                         max_iters=500, first=0):
    """Vectorised batch generator for benchmarks: the same model as `iiwa_like` (diagonal G, structural
    A/B sparsity) drawn for all systems at once.  System i depends only on (seed, first + i), so a
    shard [first, first + nsys) of a larger batch is reproducible on its own."""
    G_row, G_col, g_src, C_row, C_col, c_src, Amask, Bmask = _pattern(S, C, K, False, False)
    h = S // 2
    q_pos, q_vel, rr, q_N = _PRESETS[preset]
    n_z, n_l = (S + C) * K - C, S * K
    GV = np.empty((nsys, len(G_col)), np.float32)
    CV = np.empty((nsys, len(C_col)), np.float32)
    gg = np.empty((nsys, n_z), np.float32)
    cc = np.empty((nsys, n_l), np.float32)
    kind, k, r, c = g_src.T
    gdiag = np.where(kind == 0, np.where(r < h, q_pos, q_vel) * np.where(k == K - 1, q_N, 1.0), rr)
    GV[:] = gdiag.astype(np.float32)
    ck, kk, cr, ccol = c_src.T
    mA, mB, mI = ck == 2, ck == 3, ck == 4
    chunk = 256
    for s0 in range(0, nsys, chunk):
        n = min(chunk, nsys - s0)
        rng = np.random.default_rng([seed, first + s0])
        A = np.tile(np.eye(S), (n, K - 1, 1, 1))
        A[:, :, :h, h:2 * h] += dt * np.eye(h)
        A[:, :, h:, :] += dt * rng.standard_normal((n, K - 1, S - h, S))
        B = np.zeros((n, K - 1, S, C))
        eyeB = np.zeros((S - h, C)); m = min(S - h, C); eyeB[:m, :m] = np.eye(m)
        B[:, :, h:, :] = dt * (2.0 * eyeB + 0.3 * rng.standard_normal((n, K - 1, S - h, C)))
        cv = CV[s0:s0 + n]
        cv[:, mA] = -A[:, kk[mA], cr[mA], ccol[mA]]
        cv[:, mB] = -B[:, kk[mB], cr[mB], ccol[mB]]
        cv[:, mI] = 1.0
        gg[s0:s0 + n] = rng.standard_normal((n, n_z))
        cx = 0.01 * rng.standard_normal((n, n_l)); cx[:, :S] = 0.0
        cc[s0:s0 + n] = cx
    return dict(S=S, C=C, K=K, nsys=nsys, G_row=G_row, G_col=G_col, G_val=GV, C_row=C_row, C_col=C_col,
                C_val=CV, g=gg, c=cc, exit_tol=exit_tol, max_iters=max_iters, rho=rho)
