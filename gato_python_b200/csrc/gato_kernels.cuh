// gato_kernels.cuh -- sm_100a kernels for GATO's KKT linear-system hot path.
//
// One fused, persistent kernel per solve: CSR values -> per-knot dense blocks -> Schur complement
// S, gamma and the symmetric-stair preconditioner -> PCG -> dz recovery, all without leaving the
// chip.  A KKT system is owned by ONE thread-block cluster; each CTA owns `kpc` consecutive knot
// points, one warp per knot.  During PCG every lane keeps half a block-row (3S/2 values) of S and
// of P^-1 in REGISTERS; iterates live in shared memory; neighbour-knot blocks and the per-knot dot
// partials are exchanged through distributed shared memory (DSMEM) and `barrier.cluster`.
//
// What this replaces in the reference (/root/reference):
//   gato_convert_kkt_format          src/gato_schur.cuh:674-756
//   gato_form_schur_jacobi(+_inner)  src/gato_schur.cuh:13-494
//   gato_form_ss(+_inner)            src/gato_schur.cuh:496-671
//   parallelPCG(+_inner)             src/gato_pcg.cuh:270-470
//   gato_compute_dz(+_inner)         src/gato_schur.cuh:758-879
// Differences by design: every Q_k is inverted once; stair right blocks are
// transposes of the neighbour's left block; true signs are used internally (the reference stores
// -S, -Pinv, -gamma); reductions are order-deterministic (no atomics); the last state rows of dz
// use the intended value instead of the reference's out-of-bounds read.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>

namespace gato {
namespace cg = cooperative_groups;

constexpr int kMapDiagBit = 1 << 30;
constexpr int kMapMask = kMapDiagBit - 1;

struct SolveParams {
    // --- plan (sparsity pattern, shared by every system of a batch) ---
    const int* gmap;    // [nnzG] destination inside the knot's [Q|R] block, bit30 = diagonal, <0 = drop
    const int* cmap;    // [nnzC] destination inside the stage's [A|B] block, <0 = drop
    const int* gstart;  // [K+1]  nz range of knot k in G
    const int* cstart;  // [K+1]  nz range of stage k in C (rows (k+1)S..(k+2)S); stage K-1 is empty
    // --- per-system values (element strides between systems; 0 = shared by all systems) ---
    const float* G_val; long long ldG;
    const float* C_val; long long ldC;
    const float* g;     long long ldg;
    const float* c;     long long ldc;
    const float* lambda_in; long long ldl;  // only read when warm == 2
    // --- outputs ---
    float* lambda_out; long long ldlo;   // [nsys][S*K]
    float* dz_out;     long long lddz;   // [nsys][(S+C)K-C]
    int*   iters_out;                    // [nsys]
    float* eta_out;                      // [nsys] last |r^T P^-1 r| (may be null)
    // --- optional dumps of the Schur phase, reference layout + reference (negated) signs ---
    float* dbg_S; float* dbg_P; float* dbg_gamma; float* dbg_Ginv;
    long long* dbg_clk;                  // [16] globaltimer ns at phase boundaries (system 0, rank 0), or null
    int K, nsys, max_iters, warm;        // warm: 0/1 = reference behaviour (lambda_in ignored), 2 = true warm start
    int precond;                         // 0 = symmetric stair (the reference as shipped), 1 = block-Jacobi only (SS_PRECON=0)
    float rho, exit_tol;
};

template <int S, int C>
struct Layout {
    static constexpr int SS = S * S, SC = S * C, CC = C * C;
    // per-knot region (floats).  [A|B] and [Q|R] are contiguous, matching the reference's
    // per-stage dense layouts so the unpack maps index them directly.
    static constexpr int oA = 0;
    static constexpr int oB = oA + SS;
    static constexpr int oQi = oB + SC;
    static constexpr int oRi = oQi + SS;
    static constexpr int oq = oRi + CC;
    static constexpr int or_ = oq + S;
    static constexpr int PREV = or_ + C;       // [0, PREV) is what block row b+1 reads from stage b
    static constexpr int oPhi = PREV;
    static constexpr int oTh = oPhi + SS;
    static constexpr int oThi = oTh + SS;
    static constexpr int oPL = oThi + SS;      // BR (S*C) aliases the head of PL
    static constexpr int oGam = oPL + SS;
    static constexpr int KN = (oGam + S + 3) & ~3;
    static constexpr int H0 = (3 * S + 1) / 2;  // half block-row length held per lane
    static constexpr int MD = S > C ? S : C;
    static constexpr int CH = (S + 1) / 2;      // column half for S-wide outputs
    static constexpr int CHC = (C + 1) / 2;     // column half for C-wide outputs

    __host__ __device__ static constexpr int cta_floats(int kpc, int cl) {
        return kpc * KN + 2 * (kpc + 2) * S + 4 * S + (kpc + 1) * S + 2 * cl * kpc + PREV + 8;
    }
};

__device__ __forceinline__ void stamp(const SolveParams& p, int slot) {
    if (p.dbg_clk != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
        long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        p.dbg_clk[slot] = t;
    }
}

__device__ __forceinline__ float half_sum16(float v) {
    // butterfly inside each 16-lane half; every lane of the half ends with the same bits
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

__device__ __forceinline__ float slots_sum(const float* red, int n, int lane) {
    // fixed-order sum of the per-knot partials; identical bits in every lane, warp and CTA
    float s = 0.f;
    for (int i = lane; i < n; i += 32) s += red[i];
    s += __shfl_xor_sync(0xffffffffu, s, 16);
    s += __shfl_xor_sync(0xffffffffu, s, 8);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    return s;
}

// In-register Gauss-Jordan inversion without pivoting (the reference's choice,
// src/gato_utils.cuh:468-586; SPD blocks).  Lane (half, row) owns row `row` of a D x D matrix;
// the two 16-lane halves of a warp invert two different matrices in lock-step.
template <int MD>
__device__ __forceinline__ void gj_invert_rows(float (&a)[MD], int D, int row, int lane) {
#pragma unroll
    for (int pv = 0; pv < MD; ++pv) {
        float prow[MD];
#pragma unroll
        for (int j = 0; j < MD; ++j) prow[j] = __shfl_sync(0xffffffffu, a[j], (lane & 16) | pv);
        if (pv < D) {
            const float inv = 1.0f / prow[pv];
            if (row == pv) {
#pragma unroll
                for (int j = 0; j < MD; ++j) a[j] = prow[j] * inv;
                a[pv] = inv;
            } else {
                const float fi = a[pv] * inv;
#pragma unroll
                for (int j = 0; j < MD; ++j) a[j] = fmaf(-fi, prow[j], a[j]);
                a[pv] = -fi;
            }
        }
    }
}

// Scatter one stage's CSR values into its dense block through the plan's map.  Loads are issued
// in batches of 4 x 32 before any dependent store so the whole stage costs ~one memory latency.
__device__ __forceinline__ void scatter_stage(float* dst, const int* __restrict__ map,
                                              const float* __restrict__ val, int start, int end,
                                              float rho, int lane) {
    for (int t0 = start; t0 < end; t0 += 128) {
        int m[4];
        float v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + u * 32 + lane;
            const bool ok = t < end;
            m[u] = ok ? __ldg(map + t) : -1;
            v[u] = ok ? __ldg(val + t) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (m[u] >= 0) dst[m[u] & kMapMask] = v[u] + ((m[u] & kMapDiagBit) ? rho : 0.0f);
    }
}

template <int S, int C, int MAXW>
__global__ void __launch_bounds__(MAXW * 32, 1) gato_fused_solve(const SolveParams p) {
    using L = Layout<S, C>;
    constexpr int SS = L::SS, H0 = L::H0, MD = L::MD, CH = L::CH, CHC = L::CHC;
    extern __shared__ __align__(16) float smem[];

    cg::cluster_group cluster = cg::this_cluster();
    const int CL = (int)cluster.dim_blocks().x;
    const int rank = (int)cluster.block_rank();
    const int sys = blockIdx.x / CL;
    const int kpc = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = lane & 15, h = lane >> 4;
    const int K = p.K;
    const int b = rank * kpc + warp;        // global knot / block row owned by this warp
    const bool active = b < K;
    const bool rowS = r < S;                // lane maps to a state row
    const bool own = active && h == 0 && rowS;

    // ---- shared memory carve-up -------------------------------------------------------------
    float* kn = smem + warp * L::KN;
    float* xr = smem + kpc * L::KN;               // r with one-knot halos      [(kpc+2)S]
    float* xp = xr + (kpc + 2) * S;               // p with one-knot halos      [(kpc+2)S]
    float* hu = xp + (kpc + 2) * S;               // received boundary upsilon  [2][S] (0 left, 1 right)
    float* ht = hu + 2 * S;                       // received boundary r~       [2][S]
    float* lamh = ht + 2 * S;                     // lambda with right halo     [(kpc+1)S]
    float* redv = lamh + (kpc + 1) * S;           // p.Sp partials              [CL*kpc]
    float* rede = redv + CL * kpc;                // r.Pr partials              [CL*kpc]
    float* prevbuf = rede + CL * kpc;             // staged copy of stage b0-1  [PREV]
    const int nred = CL * kpc;
    const int nfloats = L::cta_floats(kpc, CL);

    stamp(p, 0);
    for (int i = threadIdx.x; i < nfloats; i += blockDim.x) smem[i] = 0.f;
    __syncthreads();
    stamp(p, 1);

    const float* Gv = p.G_val + (long long)sys * p.ldG;
    const float* Cv = p.C_val + (long long)sys * p.ldC;
    const float* gv = p.g + (long long)sys * p.ldg;
    const float* cv = p.c + (long long)sys * p.ldc;

    // ==== 1. unpack this knot's stage: [Q|R] -> kn+oQi, [A|B] -> kn+oA, q, r =================
    if (active) {
        scatter_stage(kn + L::oQi, p.gmap, Gv, __ldg(p.gstart + b), __ldg(p.gstart + b + 1), p.rho, lane);
        scatter_stage(kn + L::oA, p.cmap, Cv, __ldg(p.cstart + b), __ldg(p.cstart + b + 1), 0.0f, lane);
        if (lane < S) kn[L::oq + lane] = __ldg(gv + b * (S + C) + lane);
        if (lane < C && b < K - 1) kn[L::or_ + lane] = __ldg(gv + b * (S + C) + S + lane);
    }
    __syncwarp();
    stamp(p, 2);

    // ==== 2. invert Q_b (half 0) and R_b (half 1) in registers ================================
    if (active) {
        const int D = h == 0 ? S : (b < K - 1 ? C : 0);
        const bool valid = r < D;
        float a[MD];
        const float* src = kn + (h == 0 ? L::oQi : L::oRi);
        const int ld = h == 0 ? S : C;
#pragma unroll
        for (int j = 0; j < MD; ++j) a[j] = (valid && j < D) ? src[j * ld + r] : (j == r ? 1.f : 0.f);
        if (b == 0 && h == 0 && rowS) {   // block row 0 keeps raw Q0 as theta_0^-1 (src/gato_schur.cuh:75-81)
#pragma unroll
            for (int j = 0; j < S; ++j) kn[L::oThi + j * S + r] = a[j];
        }
        gj_invert_rows<MD>(a, D, r, lane);
        if (valid) {
            float* dst = kn + (h == 0 ? L::oQi : L::oRi);
#pragma unroll
            for (int j = 0; j < MD; ++j)
                if (j < D) dst[j * ld + r] = a[j];
        }
    }
    stamp(p, 3);
    cluster.sync();   // ---- (B1) every stage's A, B, Q^-1, R^-1, q, r are in shared memory

    // ==== 3. block row b: phi, gamma, theta, theta^-1 ==========================================
    float gam_r = 0.f;   // gamma_b[r] (true sign), held by lanes (h=0, r<S)
    if (active) {
        if (b == 0) {
            if (h == 0 && rowS) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < S; ++k) {
                    const float qi = kn[L::oQi + k * S + r];
                    kn[L::oTh + k * S + r] = qi;                    // theta_0 = Q0^-1
                    acc = fmaf(qi, kn[L::oq + k], acc);
                }
                gam_r = acc;                                        // c[0:S] is ignored
                kn[L::oGam + r] = acc;
            }
        } else {
            const float* prev;
            if (warp > 0) {
                prev = kn - L::KN;
            } else {   // stage b-1 lives in the previous CTA: stage it through DSMEM once
                const float* rem = cluster.map_shared_rank(smem + (kpc - 1) * L::KN, rank - 1);
                for (int i = lane; i < L::PREV; i += 32) prevbuf[i] = rem[i];
                __syncwarp();
                prev = prevbuf;
            }
            const float* Ap = prev + L::oA;
            const float* Bp = prev + L::oB;
            const float* Qip = prev + L::oQi;
            const float* Rip = prev + L::oRi;
            float* phi = kn + L::oPhi;
            float* BR = kn + L::oPL;
            if (rowS) {
                float x[S];
#pragma unroll
                for (int k = 0; k < S; ++k) x[k] = Ap[k * S + r];
#pragma unroll
                for (int cc = 0; cc < CH; ++cc) {          // phi = A Q^-1            (:277-285)
                    const int c0 = h * CH + cc;
                    if (c0 < S) {
                        float acc = 0.f;
#pragma unroll
                        for (int k = 0; k < S; ++k) acc = fmaf(x[k], Qip[c0 * S + k], acc);
                        phi[c0 * S + r] = acc;
                    }
                }
                float y[C];
#pragma unroll
                for (int k = 0; k < C; ++k) y[k] = Bp[k * S + r];
#pragma unroll
                for (int cc = 0; cc < CHC; ++cc) {         // BR = B R^-1             (:293-301)
                    const int c0 = h * CHC + cc;
                    if (c0 < C) {
                        float acc = 0.f;
#pragma unroll
                        for (int k = 0; k < C; ++k) acc = fmaf(y[k], Rip[c0 * C + k], acc);
                        BR[c0 * S + r] = acc;
                    }
                }
            }
            __syncwarp();
            if (rowS) {
                float x[S], y[C];
#pragma unroll
                for (int k = 0; k < S; ++k) x[k] = phi[k * S + r];
#pragma unroll
                for (int k = 0; k < C; ++k) y[k] = BR[k * S + r];
                if (h == 0) {                              // gamma                   (:306-339)
                    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
                    for (int k = 0; k < S; ++k) {
                        a0 = fmaf(kn[L::oQi + k * S + r], kn[L::oq + k], a0);
                        a1 = fmaf(x[k], prev[L::oq + k], a1);
                    }
#pragma unroll
                    for (int k = 0; k < C; ++k) a2 = fmaf(y[k], prev[L::or_ + k], a2);
                    gam_r = (a0 - __ldg(cv + b * S + r)) + (a2 + a1);
                    kn[L::oGam + r] = gam_r;
                }
#pragma unroll
                for (int cc = 0; cc < CH; ++cc) {          // theta = phi A^T + Q_b^-1 + BR B^T  (:342-385)
                    const int c0 = h * CH + cc;
                    if (c0 < S) {
                        float acc = 0.f;
#pragma unroll
                        for (int k = 0; k < S; ++k) acc = fmaf(x[k], Ap[k * S + c0], acc);
                        acc += kn[L::oQi + c0 * S + r];
                        float acc2 = 0.f;
#pragma unroll
                        for (int k = 0; k < C; ++k) acc2 = fmaf(y[k], Bp[k * S + c0], acc2);
                        kn[L::oTh + c0 * S + r] = acc + acc2;
                    }
                }
            }
            __syncwarp();
            {   // theta^-1 (half 0 does the work; half 1 idles on an identity)          (:409-411)
                const int D = h == 0 ? S : 0;
                float a[MD];
#pragma unroll
                for (int j = 0; j < MD; ++j)
                    a[j] = (r < D && j < D) ? kn[L::oTh + j * S + r] : (j == r ? 1.f : 0.f);
                gj_invert_rows<MD>(a, D, r, lane);
                if (r < D) {
#pragma unroll
                    for (int j = 0; j < S; ++j) kn[L::oThi + j * S + r] = a[j];
                }
            }
        }
    }
    stamp(p, 4);
    cluster.sync();   // ---- (B2) phi_b, theta_b^-1, gamma_b of every block row are in shared memory

    // ==== 4. stair left block: PL_b = theta_b^-1 phi_b theta_{b-1}^-1  (P^-1[b,left] = -PL_b) ==
    if (active && b > 0) {
        const float* thip;
        if (warp > 0) {
            thip = kn - L::KN + L::oThi;
        } else {
            const float* rem = cluster.map_shared_rank(smem + (kpc - 1) * L::KN + L::oThi, rank - 1);
            for (int i = lane; i < SS; i += 32) prevbuf[i] = rem[i];
            __syncwarp();
            thip = prevbuf;
        }
        {   // all 32 lanes run the shuffles; lanes without a state row only guard their stores
            const int rs = rowS ? r : 0;
            float x[S];
#pragma unroll
            for (int k = 0; k < S; ++k) x[k] = kn[L::oThi + k * S + rs];
            float tm[CH];                                  // T = theta_b^-1 phi_b, my column half
#pragma unroll
            for (int cc = 0; cc < CH; ++cc) {
                const int c0 = h * CH + cc;
                float acc = 0.f;
                if (c0 < S) {
#pragma unroll
                    for (int k = 0; k < S; ++k) acc = fmaf(x[k], kn[L::oPhi + c0 * S + k], acc);
                }
                tm[cc] = acc;
            }
            float trow[2 * CH];                            // full row r of T
#pragma unroll
            for (int cc = 0; cc < CH; ++cc) {
                const float other = __shfl_xor_sync(0xffffffffu, tm[cc], 16);
                trow[cc] = h == 0 ? tm[cc] : other;
                trow[CH + cc] = h == 0 ? other : tm[cc];
            }
#pragma unroll
            for (int cc = 0; cc < CH; ++cc) {
                const int c0 = h * CH + cc;
                if (c0 < S && rowS) {
                    float acc = 0.f;
#pragma unroll
                    for (int k = 0; k < S; ++k) acc = fmaf(trow[k], thip[c0 * S + k], acc);
                    kn[L::oPL + c0 * S + r] = acc;
                }
            }
        }
    } else if (active) {   // b == 0: PL region held BR scratch for nobody; make sure it is zero
        for (int i = lane; i < SS; i += 32) kn[L::oPL + i] = 0.f;
    }
    stamp(p, 5);
    cluster.sync();   // ---- (B3) every PL_b is in shared memory

    // ==== 5. pull this lane's half block-rows of S and P^-1 into registers =====================
    float Srow[H0], Prow[H0];
    {
        const bool hasNext = active && (b + 1 < K);
        const float* nxt = kn + L::KN;
        if (warp == kpc - 1 && hasNext) nxt = cluster.map_shared_rank(smem, rank + 1);
#pragma unroll
        for (int k = 0; k < H0; ++k) {
            const int ci = h * H0 + k;
            const int blk = ci / S, col = ci - blk * S;
            float sv = 0.f, pv = 0.f;
            if (active && rowS) {
                if (blk == 0) {
                    sv = kn[L::oPhi + col * S + r];
                    pv = p.precond ? 0.f : -kn[L::oPL + col * S + r];      // block-Jacobi: no off-diagonal blocks
                } else if (blk == 1) {
                    sv = kn[L::oTh + col * S + r];
                    pv = kn[L::oThi + col * S + r];
                } else if (blk == 2 && hasNext) {
                    sv = nxt[L::oPhi + r * S + col];        // S[b,right] = phi_{b+1}^T
                    pv = p.precond ? 0.f : -nxt[L::oPL + r * S + col];     // P^-1[b,right] = -PL_{b+1}^T
                }
            }
            Srow[k] = sv;
            Prow[k] = pv;
        }
        if (p.dbg_S != nullptr && active && rowS) {   // reference layout, reference (negated) signs
            float* dS = p.dbg_S + ((long long)sys * K + b) * 3 * SS;
            float* dP = p.dbg_P + ((long long)sys * K + b) * 3 * SS;
#pragma unroll
            for (int k = 0; k < H0; ++k) {
                const int ci = h * H0 + k;
                if (ci < 3 * S) {
                    dS[ci * S + r] = -Srow[k];
                    dP[ci * S + r] = -Prow[k];
                }
            }
            if (h == 0) p.dbg_gamma[((long long)sys * K + b) * S + r] = -gam_r;
        }
        if (p.dbg_Ginv != nullptr && active) {
            float* dG = p.dbg_Ginv + (long long)sys * ((SS + L::CC) * K - L::CC) + (long long)b * (SS + L::CC);
            const int n = b < K - 1 ? SS + L::CC : SS;
            for (int i = lane; i < n; i += 32) dG[i] = kn[L::oQi + i];
        }
    }

    // matvec of this lane's half-row against [x_{b-1} | x_b | x_{b+1}]; both halves get the sum
    auto matvec = [&](const float (&m)[H0], const float* x) -> float {
        const float* xc = x + warp * S + h * H0;
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int k = 0; k < H0; ++k) {
            // odd 3S: the second half is one element short; never read past the row
            const float xv = ((3 * S) % 2 == 0 || h * H0 + k < 3 * S) ? xc[k] : 0.f;
            if (k & 1) a1 = fmaf(m[k], xv, a1); else a0 = fmaf(m[k], xv, a0);
        }
        float acc = a0 + a1;
        acc += __shfl_xor_sync(0xffffffffu, acc, 16);
        return acc;
    };

    // DSMEM targets for the partial broadcast and the boundary pushes
    float* r_redv = cluster.map_shared_rank(redv, lane < CL ? lane : 0);
    float* r_rede = cluster.map_shared_rank(rede, lane < CL ? lane : 0);
    const bool pushL = active && warp == 0 && rank > 0 && h == 0 && rowS;           // my first knot -> left CTA's right halo
    const bool pushR = active && warp == kpc - 1 && rank + 1 < CL && (b + 1 < K) && h == 0 && rowS;
    float* nbL = pushL ? cluster.map_shared_rank(smem, rank - 1) : smem;
    float* nbR = pushR ? cluster.map_shared_rank(smem, rank + 1) : smem;
    const int o_xr = kpc * L::KN, o_hu = o_xr + 2 * (kpc + 2) * S, o_ht = o_hu + 2 * S,
              o_lamh = o_ht + 2 * S;
    const bool haloL = warp == 0 && h == 1 && rowS;          // lanes that maintain the halos
    const bool haloR = warp == kpc - 1 && h == 1 && rowS;
    const int slot = rank * kpc + warp;

    stamp(p, 6);
    // ==== 6. PCG (src/gato_pcg.cuh:270-439) =====================================================
    float lam = 0.f, rr = 0.f, pp = 0.f;
    if (p.warm == 2) {   // opt-in true warm start: r0 = gamma - S lambda0
        const float* l0 = p.lambda_in + (long long)sys * p.ldl;
        if (own) { lam = __ldg(l0 + b * S + r); xp[(warp + 1) * S + r] = lam; }
        if (haloL && rank > 0 && rank * kpc - 1 < K) xp[r] = __ldg(l0 + (rank * kpc - 1) * S + r);
        if (haloR && (rank + 1) * kpc < K) xp[(kpc + 1) * S + r] = __ldg(l0 + (rank + 1) * kpc * S + r);
        __syncthreads();
        const float sl = matvec(Srow, xp);
        rr = gam_r - sl;
        __syncthreads();
    } else {
        rr = gam_r;
    }
    if (own) xr[(warp + 1) * S + r] = rr;
    if (pushL) nbL[o_xr + (kpc + 1) * S + r] = rr;
    if (pushR) nbR[o_xr + r] = rr;
    cluster.sync();

    float rt = matvec(Prow, xr);
    {
        const float part = half_sum16(own ? rr * rt : 0.f);
        const float part0 = __shfl_sync(0xffffffffu, part, 0);
        if (lane < CL) r_rede[slot] = part0;
        if (pushL) nbL[o_ht + S + r] = rt;
        if (pushR) nbR[o_ht + r] = rt;
    }
    cluster.sync();
    float eta = slots_sum(rede, nred, lane);
    pp = rt;
    if (own) xp[(warp + 1) * S + r] = pp;
    if (haloL) xp[r] = ht[r];
    if (haloR) xp[(kpc + 1) * S + r] = ht[S + r];
    __syncthreads();

    int iters = p.max_iters;
    float eta_last = eta;
    stamp(p, 7);
    for (int it = 0; it < p.max_iters; ++it) {
        const float ups = matvec(Srow, xp);
        {
            const float part = half_sum16(own ? pp * ups : 0.f);
            const float part0 = __shfl_sync(0xffffffffu, part, 0);
            if (lane < CL) r_redv[slot] = part0;
            if (pushL) nbL[o_hu + S + r] = ups;
            if (pushR) nbR[o_hu + r] = ups;
        }
        cluster.sync();
        const float v = slots_sum(redv, nred, lane);
        const float alpha = eta / v;
        lam = fmaf(alpha, pp, lam);
        rr = fmaf(-alpha, ups, rr);
        if (own) xr[(warp + 1) * S + r] = rr;
        if (haloL) xr[r] = fmaf(-alpha, hu[r], xr[r]);
        if (haloR) xr[(kpc + 1) * S + r] = fmaf(-alpha, hu[S + r], xr[(kpc + 1) * S + r]);
        __syncthreads();
        rt = matvec(Prow, xr);
        {
            const float part = half_sum16(own ? rr * rt : 0.f);
            const float part0 = __shfl_sync(0xffffffffu, part, 0);
            if (lane < CL) r_rede[slot] = part0;
            if (pushL) nbL[o_ht + S + r] = rt;
            if (pushR) nbR[o_ht + r] = rt;
        }
        cluster.sync();
        const float eta_new = slots_sum(rede, nred, lane);
        eta_last = eta_new;
        if (fabsf(eta_new) < p.exit_tol) { iters = it; break; }
        const float beta = eta_new / eta;
        pp = fmaf(beta, pp, rt);
        if (own) xp[(warp + 1) * S + r] = pp;
        if (haloL) xp[r] = fmaf(beta, xp[r], ht[r]);
        if (haloR) xp[(kpc + 1) * S + r] = fmaf(beta, xp[(kpc + 1) * S + r], ht[S + r]);
        eta = eta_new;
        __syncthreads();
    }

    stamp(p, 8);
    // ==== 7. dz = G^-1 (g - C^T lambda)  (src/gato_schur.cuh:759-867) ===========================
    if (own) lamh[warp * S + r] = lam;
    if (pushL) nbL[o_lamh + kpc * S + r] = lam;
    cluster.sync();
    if (active) {
        const float* ln = lamh + (warp + 1) * S;     // lambda_{b+1}, zero past the horizon
        float* tb = kn + L::oTh;                     // theta is in registers now: reuse as scratch
        if (h == 0) {
            if (rowS) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < S; ++k) acc = fmaf(kn[L::oA + r * S + k], ln[k], acc);
                tb[r] = kn[L::oq + r] - (lam + acc);
            }
        } else if (r < C) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < S; ++k) acc = fmaf(kn[L::oB + r * S + k], ln[k], acc);
            tb[S + r] = kn[L::or_ + r] - acc;
        }
        __syncwarp();
        float* dz = p.dz_out + (long long)sys * p.lddz + (long long)b * (S + C);
        if (h == 0) {
            if (rowS) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < S; ++k) acc = fmaf(kn[L::oQi + k * S + r], tb[k], acc);
                dz[r] = acc;
                p.lambda_out[(long long)sys * p.ldlo + b * S + r] = lam;
            }
        } else if (r < C && b < K - 1) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < C; ++k) acc = fmaf(kn[L::oRi + k * C + r], tb[S + k], acc);
            dz[S + r] = acc;
        }
    }
    stamp(p, 9);
    if (rank == 0 && threadIdx.x == 0) {
        p.iters_out[sys] = iters;
        if (p.eta_out != nullptr) p.eta_out[sys] = fabsf(eta_last);
    }
}

// ------------------------------------------------------------------------------------------------
// Standalone unpack (parity with gato_convert_kkt_format, src/gato_schur.cuh:745-756): writes the
// reference's dense G / C layouts to global memory through the same maps the fused kernel uses.
// One warp per knot; outputs must be pre-zeroed (the reference callocs them, gpu_library.cu:36-37).
// ------------------------------------------------------------------------------------------------
template <int S, int C>
__global__ void gato_unpack_kernel(const int* __restrict__ gmap, const int* __restrict__ cmap,
                                   const int* __restrict__ gstart, const int* __restrict__ cstart,
                                   const float* __restrict__ G_val, const float* __restrict__ C_val,
                                   float* __restrict__ G_dense, float* __restrict__ C_dense, int K,
                                   float rho) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= K) return;
    scatter_stage(G_dense + (long long)warp * (S * S + C * C), gmap, G_val, gstart[warp],
                  gstart[warp + 1], rho, lane);
    if (warp < K - 1)
        scatter_stage(C_dense + (long long)warp * (S * S + S * C), cmap, C_val, cstart[warp],
                      cstart[warp + 1], 0.0f, lane);
}

}  // namespace gato
