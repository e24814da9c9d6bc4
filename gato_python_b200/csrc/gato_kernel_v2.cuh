// gato_kernel_v2.cuh -- "one system per SM" fused KKT solve for sm_100a.
//
// One CTA owns one KKT system for its whole life (CSR values -> dense blocks -> Schur complement ->
// PCG -> dz), so a batch runs one system per SM with block barriers only (no cluster barrier, no
// DSMEM).  The register file is the resident store of the iteration matrices: what makes a whole
// IIWA-50 system fit in one SM's 64 K registers is that the symmetric-stair preconditioner
//     P^-1 = D^-1 - D^-1 O D^-1          (src/gato_schur.cuh:498-649, SS_PRECON)
// is APPLIED in factored form  P^-1 r = D^-1 (r - O (D^-1 r))  with D^-1 = diag(theta_b^-1) and
// O = the off-diagonal blocks of S, which are in registers anyway for the S matvec.  Per knot that
// is 3*S^2 (S) + S^2 (theta^-1) = 784 floats instead of 6*S^2 = 1176, and the four S x S GEMMs per
// knot of gato_form_ss disappear.
//
// Thread mapping: LPK lanes per knot (4 for S = 14: 8 knots per warp, 7 warps for 50 knots, at most
// two warps per SM sub-partition so each thread may use ~250 registers).  During PCG a lane (g, h)
// holds R matrix rows (row group g) times one column group h of H, so each word of an iterate
// delivered from shared memory feeds R FMAs; partial rows are combined across the H lanes with a
// reduce-scatter shuffle that leaves every lane with the rows of the iterates it owns.  Iterates
// live in shared memory in a split-pad layout (each half of a knot's S entries padded to a
// multiple of 4) so all vector reads are 128-bit and both column groups see the same static pads.
//
// Replaces in the reference: gato_convert_kkt_format (src/gato_schur.cuh:674-756),
// gato_form_schur_jacobi (:13-494), gato_form_ss (:496-671, by the factored application),
// parallelPCG (src/gato_pcg.cuh:270-470) and gato_compute_dz (src/gato_schur.cuh:758-879).
#pragma once
#include "gato_kernels.cuh"

namespace gato {

template <int S, int C>
struct Layout2 {
    static constexpr int SS = S * S, SC = S * C, CC = C * C;
    static constexpr int LPK = (S > 8) ? 4 : 8;                // lanes per knot
    static constexpr int KPW = 32 / LPK;                       // knots per warp
    static constexpr int H = (S > 8 && S % 2 == 0) ? 2 : 1;    // column groups
    static constexpr int G = LPK / H;                          // row groups
    static constexpr int R = (S + G - 1) / G;                  // matrix rows per lane
    static constexpr int OWN = (R + H - 1) / H;                // iterate entries owned per lane (max)
    static constexpr int OWNC = (C + LPK - 1) / LPK;           // control rows per lane (Schur / dz phase)
    static constexpr int HALFU = S / H;                        // useful entries per column half
    static constexpr int HALFP = (HALFU + 3) & ~3;             // padded entries per column half
    static constexpr int VS = HALFP * H;                       // padded knot stride of the iterates
    static_assert(H == 1 || R == HALFU, "row groups and column halves must split at the same row");
    static_assert(C <= VS, "control vector must fit one iterate slot");
    // per-knot shared-memory region (floats); [A|B] and [Q|R] contiguous like the reference layout
    static constexpr int oA = 0;
    static constexpr int oB = oA + SS;
    static constexpr int oQ = oB + SC;
    static constexpr int oR = oQ + SS;
    static constexpr int oq = oR + CC;
    static constexpr int or_ = oq + S;
    static constexpr int oPhi = (or_ + C + 3) & ~3;
    static constexpr int oTmp = oPhi + SS;
    static constexpr int KN = (oTmp + SS + 3) & ~3;

    // position of row r inside a knot's iterate slot, and back
    __host__ __device__ static constexpr int pos(int r) { return (r / HALFU) * HALFP + (r % HALFU); }
    __host__ __device__ static constexpr bool pos_valid(int p) { return (p % HALFP) < HALFU; }
    __host__ __device__ static constexpr int pos_row(int p) { return (p / HALFP) * HALFU + (p % HALFP); }
    // iterate rows owned by lane-in-knot `lik` (= g*H + h): row group g, sub-range h of it
    __host__ __device__ static constexpr int own_row(int lik, int j) { return (lik / H) * R + (lik % H) * OWN + j; }
    __host__ __device__ static constexpr bool own_valid(int lik, int j) {
        return (lik % H) * OWN + j < R && own_row(lik, j) < S;
    }
    // owner of row r under that mapping
    __host__ __device__ static constexpr int row_lik(int r) { return (r / R) * H + (r % R) / OWN; }
    __host__ __device__ static constexpr int row_reg(int r) { return (r % R) % OWN; }
    __host__ __device__ static constexpr int warps(int K) { return (K + KPW - 1) / KPW; }
    // nt = threads per CTA (>= 32 * warps(K))
    __host__ __device__ static constexpr int smem_floats(int K, int nt) {
        const int nkp = (nt / 32) * KPW;
        return K * KN + (2 * (nkp + 2) + 2 * nkp) * VS + 64;
    }
};

// row maps for the in-register Gauss-Jordan: which lane-in-knot / register holds row r
template <int S, int C>
struct MapState {
    using L = Layout2<S, C>;
    static constexpr int D = S, N = L::OWN;
    __device__ static constexpr int lik(int r) { return L::row_lik(r); }
    __device__ static constexpr int reg(int r) { return L::row_reg(r); }
    __device__ static constexpr int row(int lik_, int i) { return L::own_valid(lik_, i) ? L::own_row(lik_, i) : -1; }
};
template <int S, int C>
struct MapControl {
    using L = Layout2<S, C>;
    static constexpr int D = C, N = L::OWNC;
    __device__ static constexpr int lik(int r) { return r / N; }
    __device__ static constexpr int reg(int r) { return r % N; }
    __device__ static constexpr int row(int lik_, int i) { return lik_ * N + i < C ? lik_ * N + i : -1; }
};

__device__ __forceinline__ float warp_sum32(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

// fixed-order sum of the per-warp partials (16-byte broadcast loads; same bits in every thread)
__device__ __forceinline__ float block_partials_sum(const float* red, int nw4) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (int i = 0; i < nw4; ++i) {
        const float4 v = reinterpret_cast<const float4*>(red)[i];
        s0 += v.x; s1 += v.y; s2 += v.z; s3 += v.w;
    }
    return (s0 + s1) + (s2 + s3);
}

// Gauss-Jordan without pivoting (src/gato_utils.cuh:468-586) on a D x D matrix whose rows are spread
// over the LPK lanes of a knot as M::row(lik, i).  Must be called by the whole warp.
template <class M, int LPK>
__device__ __forceinline__ void gj_invert_knot(float (&a)[M::N][M::D], int lik, int lane) {
    constexpr int D = M::D, N = M::N;
    const int base = lane & ~(LPK - 1);
#pragma unroll
    for (int pv = 0; pv < D; ++pv) {
        float prow[D];
#pragma unroll
        for (int j = 0; j < D; ++j) prow[j] = __shfl_sync(0xffffffffu, a[M::reg(pv)][j], base + M::lik(pv));
        const float inv = 1.0f / prow[pv];
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if (M::row(lik, i) == pv) {
#pragma unroll
                for (int j = 0; j < D; ++j) a[i][j] = prow[j] * inv;
                a[i][pv] = inv;
            } else {
                const float fi = a[i][pv] * inv;
#pragma unroll
                for (int j = 0; j < D; ++j) a[i][j] = fmaf(-fi, prow[j], a[i][j]);
                a[i][pv] = -fi;
            }
        }
    }
}

struct SolveParams2 {
    SolveParams p;        // same arguments as the cluster kernel; gmap/cmap are unused here
    const int* gdst;      // [nnzG] absolute shared-memory float offset of each G non-zero (bit30 = diagonal)
    const int* cdst;      // [nnzC] absolute shared-memory float offset of each C non-zero, <0 = dropped
    int nnzG, nnzC;
    int g_diag;           // pattern of G is purely diagonal: Q_k^-1, R_k^-1 are reciprocals
};

// NT = threads per CTA (a multiple of 32, >= LPK * K).  The register file is split over the four SM
// sub-partitions (16 K registers each, warps dealt round-robin), so the cap per thread is
// 16384 / (32 * ceil(warps / 4)); __launch_bounds__ derives exactly that from NT.
template <int S, int C, int NT>
__global__ void __launch_bounds__(NT, 1) gato_fused_solve_v2(const SolveParams2 q) {
    using L = Layout2<S, C>;
    using MS = MapState<S, C>;
    using MC = MapControl<S, C>;
    constexpr int LPK = L::LPK, OWN = L::OWN, OWNC = L::OWNC, H = L::H, R = L::R, VS = L::VS,
                  HALFP = L::HALFP, HALFU = L::HALFU, KN = L::KN;
    constexpr int NW = NT / 32, NKP = NW * L::KPW, NW4 = (NW + 3) / 4;
    extern __shared__ __align__(16) float smem[];
    const SolveParams& p = q.p;

    const int K = p.K;
    const int sys = blockIdx.x;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int lik = lane % LPK;                     // lane in knot
    const int b = warp * L::KPW + lane / LPK;       // knot / block row owned by this lane group
    const bool active = b < K;
    const int h = lik % H, g = lik / H;             // column group, row group (matrix mapping)

    // ---- shared memory carve-up ----------------------------------------------------------------
    float* kn = smem + (active ? b : 0) * KN;          // this knot's region
    float* xp = smem + K * KN;                         // p  with one-knot halos  [(NKP+2) VS]
    float* xt = xp + (NKP + 2) * VS;                   // t  with one-knot halos  [(NKP+2) VS]
    float* xr = xt + (NKP + 2) * VS;                   // r  (own knot only)      [NKP VS]
    float* xz = xr + NKP * VS;                         // z  (own knot only)      [NKP VS]
    float* redA = xz + NKP * VS;                       // per-warp partials of p.Sp    [32]
    float* redB = redA + 32;                           // per-warp partials of r.P^-1r [32]
    const int nfloats = L::smem_floats(K, NT);

    // rows of the iterates (and of the Schur-phase row work) owned by this lane
    int orow[OWN], po[OWN];
    bool ov[OWN];
#pragma unroll
    for (int j = 0; j < OWN; ++j) {
        const bool valid = L::own_valid(lik, j);
        orow[j] = valid ? L::own_row(lik, j) : 0;
        ov[j] = active && valid;
        po[j] = L::pos(orow[j]);
    }

    stamp(p, 0);
    {   // zero everything: sparse patterns leave structural zeros, halos and pads must read as 0
        float4* s4 = reinterpret_cast<float4*>(smem);
        for (int i = tid; i < nfloats / 4; i += NT) s4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    stamp(p, 1);

    const float* Gv = p.G_val + (long long)sys * p.ldG;
    const float* Cv = p.C_val + (long long)sys * p.ldC;
    const float* gv = p.g + (long long)sys * p.ldg;
    const float* cv = p.c + (long long)sys * p.ldc;

    // ==== 1. unpack: coalesced reads of the CSR values, scatter into the dense per-knot blocks ====
    for (int t = tid; t < q.nnzC; t += NT) {
        const int m = __ldg(q.cdst + t);
        const float v = __ldg(Cv + t);
        if (m >= 0) smem[m] = v;
    }
    for (int t = tid; t < q.nnzG; t += NT) {
        const int m = __ldg(q.gdst + t);
        const float v = __ldg(Gv + t);
        smem[m & kMapMask] = v + ((m & kMapDiagBit) ? p.rho : 0.0f);
    }
    {   // g -> per-knot q | r (contiguous in the region)
        const int n_z = (S + C) * K - C;
        for (int t = tid; t < n_z; t += NT) {
            const int k = t / (S + C), o = t - k * (S + C);
            smem[k * KN + L::oq + o] = __ldg(gv + t);
        }
    }
    __syncthreads();
    stamp(p, 2);

    // ==== 2. Q_b^-1 and R_b^-1 (rows spread over the knot's lanes) ================================
    // The shuffles inside the elimination need the whole warp: lanes without a matrix carry an
    // identity through it and simply do not store.
    {
        float a[OWN][S];
#pragma unroll
        for (int i = 0; i < OWN; ++i) {
#pragma unroll
            for (int j = 0; j < S; ++j)
                a[i][j] = ov[i] ? kn[L::oQ + j * S + orow[i]] : (MS::row(lik, i) == j ? 1.f : 0.f);
        }
        if (b == 0) {   // block row 0 keeps raw Q0 as theta_0^-1 (src/gato_schur.cuh:75-81)
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (ov[i]) {
#pragma unroll
                    for (int j = 0; j < S; ++j) kn[L::oTmp + j * S + orow[i]] = a[i][j];
                }
        }
        if (q.g_diag) {
#pragma unroll
            for (int i = 0; i < OWN; ++i)
#pragma unroll
                for (int j = 0; j < S; ++j)
                    if (MS::row(lik, i) == j) a[i][j] = 1.0f / a[i][j];
        } else {
            gj_invert_knot<MS, LPK>(a, lik, lane);
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < OWN; ++i)
            if (ov[i]) {
#pragma unroll
                for (int j = 0; j < S; ++j) kn[L::oQ + j * S + orow[i]] = a[i][j];
            }
        const bool hasR = active && b < K - 1;
        float c2[OWNC][C];
#pragma unroll
        for (int i = 0; i < OWNC; ++i) {
            const int row = MC::row(lik, i);
#pragma unroll
            for (int j = 0; j < C; ++j)
                c2[i][j] = (hasR && row >= 0) ? kn[L::oR + j * C + row] : (j == row ? 1.f : 0.f);
        }
        if (q.g_diag) {
#pragma unroll
            for (int i = 0; i < OWNC; ++i)
#pragma unroll
                for (int j = 0; j < C; ++j)
                    if (MC::row(lik, i) == j) c2[i][j] = 1.0f / c2[i][j];
        } else {
            gj_invert_knot<MC, LPK>(c2, lik, lane);
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < OWNC; ++i) {
            const int row = MC::row(lik, i);
            if (hasR && row >= 0) {
#pragma unroll
                for (int j = 0; j < C; ++j) kn[L::oR + j * C + row] = c2[i][j];
            }
        }
    }
    __syncthreads();   // every stage's A, B, Q^-1, R^-1, q, r are in shared memory
    stamp(p, 3);

    // ==== 3. block row b in the own-row mapping: phi_b, gamma_b, theta_b, theta_b^-1 ==============
    //   phi_b = A Q^-1, BR = B R^-1 (stage b-1), theta_b = phi A^T + Q_b^-1 + BR B^T,
    //   gamma_b = Q_b^-1 q_b - c_b + phi q_{b-1} + BR r_{b-1}          (src/gato_schur.cuh:277-385)
    float gam[OWN];
    float thi[OWN][S];
    {
        float th[OWN][S];
#pragma unroll
        for (int i = 0; i < OWN; ++i) {
            gam[i] = 0.f;
#pragma unroll
            for (int j = 0; j < S; ++j) th[i][j] = (MS::row(lik, i) == j) ? 1.f : 0.f;
        }
        if (active) {
            if (b == 0) {
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (ov[i]) {
                        float acc = 0.f;
#pragma unroll
                        for (int k = 0; k < S; ++k) {
                            const float qi = kn[L::oQ + k * S + orow[i]];
                            th[i][k] = qi;                               // theta_0 = Q0^-1 (:26-149)
                            acc = fmaf(qi, kn[L::oq + k], acc);
                        }
                        gam[i] = acc;                                    // c[0:S] is ignored
                    }
            } else {
                const float* prev = kn - KN;
                const float* Ap = prev + L::oA;
                const float* Bp = prev + L::oB;
                const float* Qip = prev + L::oQ;
                const float* Rip = prev + L::oR;
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (ov[i]) {
                        const int row = orow[i];
                        float x[S], y[C], phi[S], br[C];
#pragma unroll
                        for (int k = 0; k < S; ++k) x[k] = Ap[k * S + row];
#pragma unroll
                        for (int k = 0; k < C; ++k) y[k] = Bp[k * S + row];
#pragma unroll
                        for (int c0 = 0; c0 < S; ++c0) {
                            float acc = 0.f;
#pragma unroll
                            for (int k = 0; k < S; ++k) acc = fmaf(x[k], Qip[c0 * S + k], acc);
                            phi[c0] = acc;
                            kn[L::oPhi + c0 * S + row] = acc;
                        }
#pragma unroll
                        for (int c0 = 0; c0 < C; ++c0) {
                            float acc = 0.f;
#pragma unroll
                            for (int k = 0; k < C; ++k) acc = fmaf(y[k], Rip[c0 * C + k], acc);
                            br[c0] = acc;
                        }
                        float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
                        for (int k = 0; k < S; ++k) {
                            a0 = fmaf(kn[L::oQ + k * S + row], kn[L::oq + k], a0);
                            a1 = fmaf(phi[k], prev[L::oq + k], a1);
                        }
#pragma unroll
                        for (int k = 0; k < C; ++k) a2 = fmaf(br[k], prev[L::or_ + k], a2);
                        gam[i] = (a0 - __ldg(cv + b * S + row)) + (a2 + a1);
#pragma unroll
                        for (int c0 = 0; c0 < S; ++c0) {
                            float acc = 0.f;
#pragma unroll
                            for (int k = 0; k < S; ++k) acc = fmaf(phi[k], Ap[k * S + c0], acc);
                            acc += kn[L::oQ + c0 * S + row];
                            float acc2 = 0.f;
#pragma unroll
                            for (int k = 0; k < C; ++k) acc2 = fmaf(br[k], Bp[k * S + c0], acc2);
                            th[i][c0] = acc + acc2;
                        }
                    }
            }
        }
#pragma unroll
        for (int i = 0; i < OWN; ++i)
#pragma unroll
            for (int j = 0; j < S; ++j) thi[i][j] = th[i][j];
        // theta_b rows -> shared (tmp), to be re-read in the matrix mapping below
        if (active && b > 0) {
            // (block row 0's tmp still holds raw Q0; its theta_0 = Q0^-1 is re-read from the Q slot)
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (ov[i]) {
#pragma unroll
                    for (int j = 0; j < S; ++j) kn[L::oTmp + j * S + orow[i]] = th[i][j];
                }
        }
    }
    gj_invert_knot<MS, LPK>(thi, lik, lane);             // warp-uniform; idle lanes invert an identity
    if (active && b == 0) {
#pragma unroll
        for (int i = 0; i < OWN; ++i)
            if (ov[i]) {
#pragma unroll
                for (int j = 0; j < S; ++j) thi[i][j] = kn[L::oTmp + j * S + orow[i]];   // raw Q0
            }
    }
    stamp(p, 4);
    __syncthreads();   // phi_b of every block row (and theta_b in tmp) is in shared memory

    // ==== 4. iteration matrices into registers, matrix mapping ====================================
    // Lane (g, h) holds rows g*R .. g*R+R-1 of [S_left | S_main | S_right] restricted to its column
    // group, and the same rows of theta^-1.  Local column order (H == 2): [side knot (VS positions:
    // the left knot for h = 0, the right knot for h = 1) | own half of the main block (HALFP)];
    // (H == 1): [left | main | right], VS positions each.  Pad positions are never touched.
    constexpr int NLOC = (H == 2) ? VS + HALFP : 3 * VS;
    constexpr int MAIN0 = VS;                      // first local column of the main block (both layouts)
    constexpr int NMAIN = (H == 2) ? HALFP : VS;   // local columns of the main block held by a lane
    float Sm[R][NLOC], Tm[R][HALFP];
    {
        const float* mainsrc = kn + (b == 0 ? L::oQ : L::oTmp);   // theta_0 = Q0^-1 sits in the Q slot
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int rr = g * R + i;
            const bool rv = active && rr < S;
            const int rs = rv ? rr : 0;
#pragma unroll
            for (int l = 0; l < NMAIN; ++l) {
                float v = 0.f;
                if (l < HALFU) {
                    const int col = h * HALFU + l;        // H == 1: h = 0, HALFU = S
                    if (rv) v = mainsrc[col * S + rs];
                }
                Sm[i][MAIN0 + l] = v;
            }
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < OWN; ++i)
            if (ov[i]) {
#pragma unroll
                for (int j = 0; j < S; ++j) kn[L::oTmp + j * S + orow[i]] = thi[i][j];
            }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int rr = g * R + i;
            const bool rv = active && rr < S;
            const int rs = rv ? rr : 0;
#pragma unroll
            for (int l = 0; l < HALFP; ++l) {
                float v = 0.f;
                if (l < HALFU) {
                    const int col = h * HALFU + l;
                    if (rv) v = kn[L::oTmp + col * S + rs];
                }
                Tm[i][l] = v;
            }
        }
        const bool hasL = active && b > 0, hasR = active && b + 1 < K;
        const float* phiL = kn + L::oPhi;            // phi_b      : S[b,left]  = phi_b
        const float* phiR = kn + KN + L::oPhi;       // phi_{b+1}  : S[b,right] = phi_{b+1}^T
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int rr = g * R + i;
            const bool rv = active && rr < S;
            const int rs = rv ? rr : 0;
            if (H == 2) {
#pragma unroll
                for (int l = 0; l < VS; ++l) {
                    float v = 0.f;
                    if (L::pos_valid(l)) {
                        const int col = L::pos_row(l);
                        if (h == 0) { if (rv && hasL) v = phiL[col * S + rs]; }
                        else        { if (rv && hasR) v = phiR[rs * S + col]; }
                    }
                    Sm[i][l] = v;
                }
            } else {
#pragma unroll
                for (int l = 0; l < VS; ++l) {
                    float vl = 0.f, vr = 0.f;
                    if (l < S) {
                        if (rv && hasL) vl = phiL[l * S + rs];
                        if (rv && hasR) vr = phiR[rs * S + l];
                    }
                    Sm[i][l] = vl;
                    Sm[i][2 * VS + l] = vr;
                }
            }
        }
    }
    stamp(p, 5);
    // ---- iterate plumbing ------------------------------------------------------------------------
    auto store_own = [&](float* xk, const float (&v)[OWN]) {   // xk = base of this knot's slot
#pragma unroll
        for (int j = 0; j < OWN; ++j)
            if (ov[j]) xk[po[j]] = v[j];
    };
    auto scatter = [&](const float (&acc)[R], float (&out)[OWN]) {   // combine the H column groups
        if (H == 2) {
#pragma unroll
            for (int j = 0; j < OWN; ++j) {   // h = 0 keeps rows [0, OWN), h = 1 keeps rows [OWN, R)
                const float hi = (OWN + j < R) ? acc[OWN + j] : 0.f;
                const float send = h ? acc[j] : hi;
                const float mine = h ? hi : acc[j];
                out[j] = mine + __shfl_xor_sync(0xffffffffu, send, 1);
            }
        } else {
#pragma unroll
            for (int j = 0; j < OWN; ++j) out[j] = acc[j];
        }
    };
    // rows of [S_left | S_main | S_right] against x (halo layout: knot k at slot k + 1)
    auto apply_S = [&](const float* x, float (&out)[OWN]) {
        float acc[R];
#pragma unroll
        for (int i = 0; i < R; ++i) acc[i] = 0.f;
        if (H == 2) {
            const float4* side = reinterpret_cast<const float4*>(x + (b + 2 * h) * VS);
            const float4* mainh = reinterpret_cast<const float4*>(x + (b + 1) * VS + h * HALFP);
#pragma unroll
            for (int ci = 0; ci < VS / 4; ++ci) {
                const float4 v = side[ci];
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (L::pos_valid(4 * ci + u)) {
#pragma unroll
                        for (int i = 0; i < R; ++i) acc[i] = fmaf(Sm[i][4 * ci + u], e[u], acc[i]);
                    }
            }
#pragma unroll
            for (int ci = 0; ci < HALFP / 4; ++ci) {
                const float4 v = mainh[ci];
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (4 * ci + u < HALFU) {
#pragma unroll
                        for (int i = 0; i < R; ++i) acc[i] = fmaf(Sm[i][VS + 4 * ci + u], e[u], acc[i]);
                    }
            }
        } else {
            const float4* x4 = reinterpret_cast<const float4*>(x + b * VS);
#pragma unroll
            for (int ci = 0; ci < 3 * VS / 4; ++ci) {
                const float4 v = x4[ci];
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if ((4 * ci + u) % VS < S) {
#pragma unroll
                        for (int i = 0; i < R; ++i) acc[i] = fmaf(Sm[i][4 * ci + u], e[u], acc[i]);
                    }
            }
        }
        scatter(acc, out);
    };
    // off-diagonal blocks only: O x = S_left x_{b-1} + S_right x_{b+1}
    auto apply_O = [&](const float* x, float (&out)[OWN]) {
        float acc[R];
#pragma unroll
        for (int i = 0; i < R; ++i) acc[i] = 0.f;
        if (H == 2) {
            const float4* side = reinterpret_cast<const float4*>(x + (b + 2 * h) * VS);
#pragma unroll
            for (int ci = 0; ci < VS / 4; ++ci) {
                const float4 v = side[ci];
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (L::pos_valid(4 * ci + u)) {
#pragma unroll
                        for (int i = 0; i < R; ++i) acc[i] = fmaf(Sm[i][4 * ci + u], e[u], acc[i]);
                    }
            }
        } else {
            const float4* xl = reinterpret_cast<const float4*>(x + b * VS);
            const float4* xr4 = reinterpret_cast<const float4*>(x + (b + 2) * VS);
#pragma unroll
            for (int ci = 0; ci < VS / 4; ++ci) {
                const float4 v = xl[ci], w = xr4[ci];
                const float e[4] = {v.x, v.y, v.z, v.w}, f[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (4 * ci + u < S) {
#pragma unroll
                        for (int i = 0; i < R; ++i) {
                            acc[i] = fmaf(Sm[i][4 * ci + u], e[u], acc[i]);
                            acc[i] = fmaf(Sm[i][2 * VS + 4 * ci + u], f[u], acc[i]);
                        }
                    }
            }
        }
        scatter(acc, out);
    };
    // theta_b^-1 against this knot's slot xk (no halo)
    auto apply_T = [&](const float* xk, float (&out)[OWN]) {
        float acc[R];
#pragma unroll
        for (int i = 0; i < R; ++i) acc[i] = 0.f;
        const float4* x4 = reinterpret_cast<const float4*>(xk + h * HALFP);
#pragma unroll
        for (int ci = 0; ci < HALFP / 4; ++ci) {
            const float4 v = x4[ci];
            const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (4 * ci + u < HALFU) {
#pragma unroll
                    for (int i = 0; i < R; ++i) acc[i] = fmaf(Tm[i][4 * ci + u], e[u], acc[i]);
                }
        }
        scatter(acc, out);
    };
    float* xr_k = xr + b * VS;
    float* xz_k = xz + b * VS;
    float* xt_k = xt + (b + 1) * VS;
    float* xp_k = xp + (b + 1) * VS;
    // r~ = P^-1 r = theta^-1 (r - O theta^-1 r); contains one block barrier
    auto precond = [&](const float (&r_)[OWN], float (&out)[OWN]) {
        float t[OWN], o[OWN], z[OWN];
        store_own(xr_k, r_);
        __syncwarp();
        apply_T(xr_k, t);
        store_own(xt_k, t);
        __syncthreads();
        apply_O(xt, o);
#pragma unroll
        for (int j = 0; j < OWN; ++j) z[j] = r_[j] - o[j];
        store_own(xz_k, z);
        __syncwarp();
        apply_T(xz_k, out);
    };

    stamp(p, 6);
    // ==== 5. PCG (src/gato_pcg.cuh:270-439; reductions are order-deterministic, no atomics) =======
    float lam[OWN], rr[OWN], pp[OWN], rt[OWN], ups[OWN];
#pragma unroll
    for (int j = 0; j < OWN; ++j) { lam[j] = 0.f; rr[j] = ov[j] ? gam[j] : 0.f; }
    if (p.warm == 2) {   // opt-in true warm start: r0 = gamma - S lambda0
        const float* l0 = p.lambda_in + (long long)sys * p.ldl;
#pragma unroll
        for (int j = 0; j < OWN; ++j)
            if (ov[j]) lam[j] = __ldg(l0 + b * S + orow[j]);
        store_own(xp_k, lam);
        __syncthreads();
        apply_S(xp, ups);
#pragma unroll
        for (int j = 0; j < OWN; ++j) rr[j] = ov[j] ? gam[j] - ups[j] : 0.f;
        __syncthreads();
    }
    precond(rr, rt);
    {
        float part = 0.f;
#pragma unroll
        for (int j = 0; j < OWN; ++j) part = fmaf(rr[j], rt[j], part);
        part = warp_sum32(part);
        if (lane == 0) redB[warp] = part;
    }
#pragma unroll
    for (int j = 0; j < OWN; ++j) pp[j] = rt[j];
    store_own(xp_k, pp);
    __syncthreads();
    float eta = block_partials_sum(redB, NW4);
    int iters = p.max_iters;
    float eta_last = eta;
    stamp(p, 7);
    for (int it = 0; it < p.max_iters; ++it) {
        apply_S(xp, ups);
        {
            float part = 0.f;
#pragma unroll
            for (int j = 0; j < OWN; ++j) part = fmaf(pp[j], ups[j], part);
            part = warp_sum32(part);
            if (lane == 0) redA[warp] = part;
        }
        __syncthreads();
        const float v = block_partials_sum(redA, NW4);
        const float alpha = eta / v;
#pragma unroll
        for (int j = 0; j < OWN; ++j) {
            lam[j] = fmaf(alpha, pp[j], lam[j]);
            rr[j] = fmaf(-alpha, ups[j], rr[j]);
        }
        precond(rr, rt);
        {
            float part = 0.f;
#pragma unroll
            for (int j = 0; j < OWN; ++j) part = fmaf(rr[j], rt[j], part);
            part = warp_sum32(part);
            if (lane == 0) redB[warp] = part;
        }
        __syncthreads();
        const float eta_new = block_partials_sum(redB, NW4);
        eta_last = eta_new;
        if (fabsf(eta_new) < p.exit_tol) { iters = it; break; }
        const float beta = eta_new / eta;
#pragma unroll
        for (int j = 0; j < OWN; ++j) pp[j] = fmaf(beta, pp[j], rt[j]);
        store_own(xp_k, pp);
        eta = eta_new;
        __syncthreads();
    }
    stamp(p, 8);

    // ==== 6. dz = G^-1 (g - C^T lambda)  (src/gato_schur.cuh:759-867) =============================
    store_own(xp_k, lam);
    if (active) {
#pragma unroll
        for (int j = 0; j < OWN; ++j)
            if (ov[j]) p.lambda_out[(long long)sys * p.ldlo + b * S + orow[j]] = lam[j];
    }
    __syncthreads();
    if (active) {
        const float* ln = xp + (b + 2) * VS;        // lambda_{b+1}; zero past the horizon
#pragma unroll
        for (int j = 0; j < OWN; ++j) {
            const int row = orow[j];
            if (ov[j]) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < S; ++k) acc = fmaf(kn[L::oA + row * S + k], ln[L::pos(k)], acc);
                xz_k[po[j]] = kn[L::oq + row] - (lam[j] + acc);
            }
        }
#pragma unroll
        for (int j = 0; j < OWNC; ++j) {
            const int cr = MC::row(lik, j) < 0 ? C : MC::row(lik, j);
            if (cr < C) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < S; ++k) acc = fmaf(kn[L::oB + cr * S + k], ln[L::pos(k)], acc);
                xr_k[cr] = kn[L::or_ + cr] - acc;
            }
        }
    }
    __syncwarp();
    if (active) {
        float* dz = p.dz_out + (long long)sys * p.lddz + (long long)b * (S + C);
#pragma unroll
        for (int j = 0; j < OWN; ++j) {
            const int row = orow[j];
            if (ov[j]) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < S; ++k) acc = fmaf(kn[L::oQ + k * S + row], xz_k[L::pos(k)], acc);
                dz[row] = acc;
            }
        }
        if (b < K - 1) {
#pragma unroll
            for (int j = 0; j < OWNC; ++j) {
                const int cr = MC::row(lik, j) < 0 ? C : MC::row(lik, j);
                if (cr < C) {
                    float acc = 0.f;
#pragma unroll
                    for (int k = 0; k < C; ++k) acc = fmaf(kn[L::oR + k * C + cr], xr_k[k], acc);
                    dz[S + cr] = acc;
                }
            }
        }
    }
    stamp(p, 9);
    if (tid == 0) {
        p.iters_out[sys] = iters;
        if (p.eta_out != nullptr) p.eta_out[sys] = fabsf(eta_last);
    }
}

}  // namespace gato
