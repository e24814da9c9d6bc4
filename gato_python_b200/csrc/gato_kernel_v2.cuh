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
    // Per-knot shared-memory region (floats).  Every index that runs over state rows uses the
    // split-pad position pos(r) with leading dimension VS, so a lane's own rows are one aligned
    // float4 and a whole column / row is VS/4 float4:
    //   A   [k < S][VS]   A[r,k]   at k*VS + pos(r)        B    [k < C][VS]  B[r,k] at k*VS + pos(r)
    //   Q   [r < S][VS]   Q[r,c]   at r*VS + pos(c)  (raw Q, then Q^-1, row-major)
    //   R   [r < C][LDC]  R[r,c]   at r*LDC + c      (raw R, then R^-1, row-major)
    //   q   [VS] at pos(r)         r [LDC]
    //   phi [k < S][VS]   phi[r,k] at k*VS + pos(r)  (block row 0 stashes raw Q0 here, row-major)
    static constexpr int LDC = (C + 3) & ~3;
    static constexpr int oA = 0;
    static constexpr int oB = oA + S * VS;
    static constexpr int oQ = oB + C * VS;
    static constexpr int oR = oQ + S * VS;
    static constexpr int oq = oR + C * LDC;
    static constexpr int or_ = oq + VS;
    static constexpr int oPhi = or_ + LDC;
    static constexpr int KN0 = oPhi + S * VS;
    // consecutive knots of a warp start 4 banks apart: KPW 16-byte accesses at the same in-knot
    // offset hit distinct banks
    static constexpr int KN = ((KN0 - 4 + 31) / 32) * 32 + 4;
    static constexpr bool VEC_OWN = (OWN == 4);   // own rows = one aligned float4 of positions

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
    // cluster variant: kpc own knots + 2 halo knot regions per CTA, two partial arrays with one entry per
    // warp of every CTA of a (<= 16-CTA) cluster
    __host__ __device__ static constexpr int red_cl(int nt) { return (16 * (nt / 32) + 3) & ~3; }
    __host__ __device__ static constexpr int smem_floats_cl(int kpc, int nt) {
        const int nkp = (nt / 32) * KPW;
        return (kpc + 2) * KN + (2 * (nkp + 2) + 2 * nkp) * VS + 2 * red_cl(nt) + 2 * VS;
    }
};

// Kernel 5: what its first half (Schur + congruence, FP64) hands to its second half (PCG + dz) through global memory,
// per knot, in floats.  [0, REG): verbatim copy of the knot's shared-memory region A | B | Q^-1 | R^-1 | q | r (the dz
// phase reads it with the same offsets; with a diagonal cost pattern only the diagonal of Q^-1 is copied); PSI:
// psi_b = M_b phi_b M_{b-1}^T in lane-ready order of the second half (2 lanes x PSIL floats: lane g holds rows g S/2 ..
// as two row-major S/2 x S/2 blocks, the column half g first); M: M_b row-major [S][VS] in pos layout (the final
// lambda = M^T lambda^); GAM: gamma^_b in pos layout.
template <int S, int C>
struct ScaledLayout {
    using L = Layout2<S, C>;
    static constexpr int REG = L::oPhi;
    static constexpr int PSIB = (L::R * L::HALFU + 3) & ~3;   // one R x HALFU block, row-major, padded to 16 bytes
    static constexpr int PSIL = 2 * PSIB;                     // lane g of the second half: [block of its column half g | other half]
    static constexpr int oPSI = REG;
    static constexpr int oM = oPSI + 2 * PSIL;
    static constexpr int oGAM = oM + S * L::VS;
    static constexpr int KSCR = oGAM + L::VS;
    // shared-matrix extras per knot: phi_b as in shared memory ([c][VS], rows in pos layout) | M_b in FP64 [S][S]
    static constexpr int xPHI = 0, xM64 = S * L::VS, KSHX = xM64 + 2 * S * S;
    // FP64 scratch inside a knot's (dead) shared-memory region: M64 [S][S] doubles, then T64 [S][2 HT] doubles
    static constexpr int HT = (L::HALFU + 1) & ~1;
    static constexpr int oM64 = 0, oT64 = 2 * S * S;
    static_assert(L::H == 2 && L::LPK == 4, "kernel 5 is written for the 2 x 2 lane split (even S > 8)");
    static_assert(oT64 + 2 * S * 2 * HT <= L::KN0, "FP64 scratch must fit the knot region");
    static_assert(REG % 4 == 0 && KSCR % 4 == 0, "16-byte aligned sections");
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

// Sum over the warp in THREE dependent shuffle steps instead of five (4-ary, 4-ary, 2-ary: seven shuffles): the PCG
// loops wait on the dependent chain of their reductions, not on shuffle issue slots.  Every lane ends with the same
// bits (each level adds the same pair sums, and a + b == b + a exactly).
__device__ __forceinline__ float warp_sum32(float v) {
    const float a1 = __shfl_xor_sync(0xffffffffu, v, 1), a2 = __shfl_xor_sync(0xffffffffu, v, 2),
                a3 = __shfl_xor_sync(0xffffffffu, v, 3);
    v = (v + a1) + (a2 + a3);
    const float b1 = __shfl_xor_sync(0xffffffffu, v, 4), b2 = __shfl_xor_sync(0xffffffffu, v, 8),
                b3 = __shfl_xor_sync(0xffffffffu, v, 12);
    v = (v + b1) + (b2 + b3);
    return v + __shfl_xor_sync(0xffffffffu, v, 16);
}

// fixed-order sum of the per-warp partials (16-byte broadcast loads; same bits in every thread)
__device__ __forceinline__ float block_partials_sum(const float* red, int nw4) {
    const float4 v0 = reinterpret_cast<const float4*>(red)[0];   // nw4 >= 1
    float s0 = v0.x, s1 = v0.y, s2 = v0.z, s3 = v0.w;
    for (int i = 1; i < nw4; ++i) {
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
        float inv;   // reciprocal seed + one Newton step (the IEEE division subroutine sits on the pivot chain)
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(prow[pv]));
        inv = fmaf(fmaf(-prow[pv], inv, 1.0f), inv, inv);
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

// eta / v without the IEEE division subroutine: reciprocal seed + one residual correction (the
// quotient is correctly rounded except in rare last-bit cases; operands are far from the range ends)
__device__ __forceinline__ float rcp_seed(float b) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
    return r;
}
// r = rcp_seed(b), formed by the caller before a becomes known (takes the MUFU off the dependent chain)
__device__ __forceinline__ float fast_div_seeded(float a, float b, float r) {
    const float qv = a * r;
    return fmaf(fmaf(-b, qv, a), r, qv);
}
__device__ __forceinline__ float fast_div(float a, float b) { return fast_div_seeded(a, b, rcp_seed(b)); }

struct SolveParams2 {
    SolveParams p;        // same arguments as the cluster kernel; gmap/cmap are unused here
    const int* gdst;      // [nnzG] absolute shared-memory float offset of each G non-zero (bit30 = diagonal)
    const int* cdst;      // [nnzC] absolute shared-memory float offset of each C non-zero, <0 = dropped
    int nnzG, nnzC;
    int g_diag;           // pattern of G is purely diagonal: Q_k^-1, R_k^-1 are reciprocals
    int stage_off;        // > 0: float offset of the TMA staging area in shared memory (values of C, G and g
                          // arrive by cp.async.bulk, the NEXT system's while this one iterates); 0: direct loads
    int tail_off;         // float offset of the 48-byte control tail: mbarrier (staging; cluster variant: reduction A)
                          // | current system | staging parity | mbarriers B, C of the cluster variant | their parities
    int kpc;              // cluster variant: knots owned per CTA (local slots 1..kpc; 0 and kpc+1 are halos)
    int ncta;             // grid variant: CTAs per system
    float* gbuf;          // grid variant: global scratch, [2][256] per-CTA dot partials | [ncta][3][2][VS] boundary slices
    float* scr;           // kernel 5 (scaled system): per-system scratch between its two kernels (ScaledLayout)
    long long scr_ld;     // floats between consecutive systems in scr
    // kernel 5 with matrices shared by all systems of a call (ldG == ldC == 0): the first half runs ONCE and also leaves
    // phi_b and M_b (FP64) per knot in shx (ScaledLayout::KSHX floats per knot); gato_gamma_scaled then forms gamma^ of
    // every system into gam, and the second half reads psi / M / the regions from the one scratch (scr_ld = 0), gamma^
    // from gam and q, r from the caller's g.
    // one-CTA kernels with staged values: the non-zeros of C in an order whose every 32 consecutive entries hit 32
    // distinct banks on both sides of the scatter -- one word per entry, (shared-memory destination << kPermShift) | index
    // into C_val, kPermPad = padding; null: scatter in CSR order through cdst
    const int* cperm;
    int ncperm;
    float* shx;
    float* gam;           // [nsys][K VS] gamma^ in pos layout, or null: gamma^ sits in the scratch
    long long gam_ld;
};

constexpr int kPermPad = (int)0x80000000;   // padding entry of that map: negative (skipped by the store), index bits zero
constexpr int kPermShift = 14;   // bank-sorted scatter map: low bits = index into C_val (< 16384), high bits = destination (< 2^17)

// ---- TMA bulk copy (cp.async.bulk, global -> shared, completion on an mbarrier) ---------------------
__device__ __forceinline__ uint32_t smem_u32(const void* ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// ---- cluster variant: remote stores that signal the RECEIVER's mbarrier (st.async, complete_tx) ---------
// A waiter that sees the phase complete (acquire at cluster scope) sees the data: no cluster-wide barrier.
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, int cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void st_async_f32(uint32_t raddr, float v, uint32_t rmbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f32 [%0], %1, [%2];" ::"r"(raddr),
                 "f"(v), "r"(rmbar) : "memory");
}
__device__ __forceinline__ void st_async_v4(uint32_t raddr, float a, float b, float c, float d, uint32_t rmbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(
                     raddr), "f"(a), "f"(b), "f"(c), "f"(d), "r"(rmbar) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAITC_%=:\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONEC_%=;\n\t"
        "bra WAITC_%=;\n\t"
        "DONEC_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// A bulk copy needs 16-byte aligned addresses and sizes; rows of a [nsys][n] float array are only
// 4-byte aligned, so the copy covers the enclosing 16-byte words (at most 12 bytes either side, inside
// words that hold row data) and the consumer skips `lead` floats.
struct StageSpan { const char* src; uint32_t bytes; int lead; };
__device__ __forceinline__ StageSpan stage_span(const float* row, int n) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(row), a0 = a & ~(uintptr_t)15;
    StageSpan sp;
    sp.src = reinterpret_cast<const char*>(a0);
    sp.lead = (int)((a - a0) >> 2);
    sp.bytes = n > 0 ? (uint32_t)(((a - a0) + 4u * (uint32_t)n + 15u) & ~15u) : 0u;
    return sp;
}
__host__ __device__ constexpr int stage_len(int n) { return (n + 8) & ~3; }   // floats reserved for a row of n


// NT = threads per CTA (a multiple of 32, >= LPK * K).  The register file is split over the four SM
// sub-partitions (16 K registers each, warps dealt round-robin), so the cap per thread is
// 16384 / (32 * ceil(warps / 4)); __launch_bounds__ derives exactly that from NT.
//
// CL = true: the same kernel over ONE THREAD-BLOCK CLUSTER per system, for horizons that do not fit one
// SM's registers (IIWA: K > 56).  CTA c of the cluster owns q.kpc consecutive knots as local slots
// 1..kpc; slot 0 / slot kpc+1 are HALO knots (the last knot of CTA c-1 / the first of CTA c+1) whose lanes
// redo just enough of the Schur phase (Q^-1, R^-1 of the left halo; phi of the right halo) that no CTA
// ever READS a neighbour's shared memory.  During PCG the boundary knots push their slices of p and
// theta^-1 r into the neighbour CTA's halo slot through DSMEM, every warp pushes its dot-product partial
// to every CTA, and the four block barriers per iteration become cluster barriers
// (barrier.cluster arrive.release / wait.acquire); all CTAs sum the same partials in the same order, so
// they agree bit for bit on alpha, beta and the exit test.
//
// GRID = true (implies CL): the cluster variant's decomposition over a COOPERATIVE GRID, for horizons beyond one
// 16-CTA cluster (the job of the reference's grid-stride parallelPCG_fixed, src/gato_pcg.cuh:17-268).  The CTAs are
// not cluster mates, so boundary slices and per-CTA dot partials travel through a small global scratch buffer and
// the synchronisation points are grid.sync(); every CTA reduces the same partials in the same order.
//
// SCALED = true: FIRST HALF OF KERNEL 5 (gato_kernel_v5.cuh).  The same unpack and Schur phases, then -- instead of
// the PCG loop -- the congruence  S^ = M S M^T  with M_b = inverse Cholesky factor of theta_b, formed in FP64, whose
// diagonal blocks are the identity: only psi_b = M_b phi_b M_{b-1}^T, gamma^ = M gamma and M leave the SM (to the
// per-system scratch q.scr), together with a verbatim copy of the knot regions for the dz phase of the second half.
// Development builds (-DGATO_S_TRACE): thread 0 of CTA 0 accumulates clock64() differences between the phase boundaries
// of every system it takes (static shared memory) and leaves the sums 2 KB behind the work counter of kernel 5.
// 1: fetch the first round of the bank-sorted scatter map before shared memory is cleared (first half of kernel 5 only).
// Measured three ways -- 24 (index, destination) pairs per thread: the 48 live registers spill, 1.38 vs 1.32 ms per 8192
// systems; 22 packed words per thread, plain loads: ptxas sinks them back to their use; the same as volatile loads: no
// spills and no gain (the scatter of C stays at ~3 750 cycles per system: it is not waiting for the map) -- off.
#ifndef GATO_V5_INV_STEPS
#define GATO_V5_INV_STEPS 2   // Newton steps on the reciprocal of a pivot in the FP64 elimination of kernel 5's first half (1: 2^-46, and 6 685 instead of 7 000 cycles for that phase: 0.2 % of the kernel -- not worth the margin)
#endif
#ifndef GATO_RANK1_UNROLL
#define GATO_RANK1_UNROLL 7   // k steps of phi A^T in flight (2: 7 730, 7: 7 350, 14: 7 200 cycles for the phase, but 14 spills)
#endif
#ifndef GATO_PERM_UP
#define GATO_PERM_UP 22
#endif
#ifndef GATO_PERM_PREFETCH
#define GATO_PERM_PREFETCH 0
#endif
#ifdef GATO_S_TRACE
#define S_STAMP(i) do { if (SCALED && blockIdx.x == 0 && threadIdx.x == 0) { const long long t_ = clock64(); s_tr[1 + (i)] += t_ - s_tr[0]; s_tr[0] = t_; } } while (0)
#else
#define S_STAMP(i) do { } while (0)
#endif

// XW (first half of kernel 5 only): one warp MORE than the layout's NT threads.  Its lanes own no knot; while the other
// warps form theta and eliminate, it copies the knot regions to the scratch -- a job bound by the SM's store path, which
// costs the warp that does it ~3 400 cycles per system whatever else that warp could be doing (gato_kernel_v5_schur.inc).
template <int S, int C, int NT, bool CL = false, bool GRID = false, bool SCALED = false, bool XW = false>
__global__ void __launch_bounds__(XW ? NT + 32 : NT, 1) gato_fused_solve_v2(const SolveParams2 q) {
    static_assert(!GRID || CL, "the grid variant uses the cluster variant's decomposition");
    static_assert(!SCALED || !CL, "the scaled-system variant is a one-CTA kernel");
    using L = Layout2<S, C>;
    using MS = MapState<S, C>;
    using MC = MapControl<S, C>;
    constexpr int LPK = L::LPK, OWN = L::OWN, OWNC = L::OWNC, H = L::H, R = L::R, VS = L::VS,
                  HALFP = L::HALFP, HALFU = L::HALFU, KN = L::KN, LDC = L::LDC;
    constexpr bool VEC = L::VEC_OWN;
    constexpr int NW = NT / 32, NKP = NW * L::KPW, NW4 = (NW + 3) / 4;
    constexpr int NTT = XW ? NT + 32 : NT;    // threads launched (loop strides); NT stays the layout's thread count
    static_assert(!XW || (SCALED && !CL), "the extra warp belongs to the first half of kernel 5");
    extern __shared__ __align__(16) float smem[];
    const SolveParams& p = q.p;
    const int K = p.K;
    // The persistent loop carries NO register across a system: its state (current system, mbarrier parity)
    // lives in the control tail of shared memory, so the PCG loop keeps the whole register file.
    // TMA staging of the value arrays (the batched path): one thread issues three bulk copies per system;
    // the copy for the NEXT system of this persistent CTA is in flight during the PCG loop.
    const bool staged = q.stage_off > 0;
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + q.tail_off);
    volatile int* ctl = reinterpret_cast<volatile int*>(smem + q.tail_off + 2);
    auto issue_stage = [&](int s_) {   // one thread only
        float* stC_ = smem + q.stage_off;
        float* stG_ = stC_ + stage_len(q.nnzC);
        float* stg_ = stG_ + stage_len(q.nnzG);
        const StageSpan c_ = stage_span(p.C_val + (long long)s_ * p.ldC, q.nnzC);
        const StageSpan g_ = stage_span(p.G_val + (long long)s_ * p.ldG, q.nnzG);
        const StageSpan v_ = stage_span(p.g + (long long)s_ * p.ldg, (S + C) * K - C);
        mbar_expect_tx(mbar, c_.bytes + g_.bytes + v_.bytes);
        if (c_.bytes) bulk_g2s(stC_, c_.src, c_.bytes, mbar);
        if (g_.bytes) bulk_g2s(stG_, g_.src, g_.bytes, mbar);
        if (v_.bytes) bulk_g2s(stg_, v_.src, v_.bytes, mbar);
    };
    int crank = 0, ncta = 1;
    if constexpr (GRID) {
        ncta = q.ncta;
        crank = (int)blockIdx.x % ncta;
    } else if constexpr (CL) {
        cg::cluster_group cluster = cg::this_cluster();
        crank = (int)cluster.block_rank();
        ncta = (int)cluster.dim_blocks().x;
    }
    // cluster variant: the three per-iteration synchronisation points are transaction barriers of the RECEIVING
    // CTA (A: dot partials of p.Sp, B: theta^-1 r slices of the boundary knots, C: partials of r.r~ and r~ slices)
    uint64_t* mbA = mbar;
    uint64_t* mbB = reinterpret_cast<uint64_t*>(smem + q.tail_off + 4);
    uint64_t* mbC = reinterpret_cast<uint64_t*>(smem + q.tail_off + 6);
    volatile int* parity_word = reinterpret_cast<volatile int*>(smem + q.tail_off + 8);
    if (threadIdx.x == 0) {
        if constexpr (CL && !GRID) {
            mbar_init(mbA, 1);
            mbar_init(mbB, 1);
            mbar_init(mbC, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            parity_word[0] = 0;
        }
        ctl[0] = CL ? blockIdx.x / ncta : blockIdx.x;
        ctl[1] = 0;
        if (staged) {
            mbar_init(mbar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            if ((int)blockIdx.x < p.nsys) issue_stage(blockIdx.x);
        }
    }
    __syncthreads();

    // persistent: CTA i solves systems i, i + gridDim.x, ...
#ifdef GATO_S_TRACE
    __shared__ long long s_tr[24];
    if (threadIdx.x == 0) { for (int i = 0; i < 24; ++i) s_tr[i] = 0; s_tr[0] = clock64(); }
#endif
#pragma unroll 1
    while (true) {
        int sys = ctl[0];
        if (sys >= p.nsys) break;
        // Everything below is recomputed per system from an opaque zero, so that the compiler cannot
        // hoist per-thread address arithmetic out of the persistent loop and keep it live (spilled)
        // across the PCG loop: each system compiles like a straight-line kernel.
        int opaque;
        asm volatile("mov.u32 %0, 0;" : "=r"(opaque));

        const int tid = (int)threadIdx.x + opaque, warp = tid >> 5, lane = tid & 31;
        const int lik = lane % LPK;                     // lane in knot
        const int slot = warp * L::KPW + lane / LPK;    // local knot slot of this lane group
        const int k0 = CL ? crank * q.kpc : 0;          // first knot owned by this CTA
        const int b = CL ? k0 - 1 + slot : slot;        // GLOBAL knot / block row of this lane group
        const bool active = CL ? (slot >= 1 && slot <= q.kpc && b < K) : (b < K);   // owns block row b
        const bool lhalo = CL && slot == 0 && b >= 0;           // redoes Q^-1, R^-1 of the previous CTA's last knot
        const bool rhalo = CL && slot == q.kpc + 1 && b < K;    // redoes phi of the next CTA's first knot
        const int nreg = CL ? q.kpc + 2 : K;            // knot regions in this CTA's shared memory
        // iterate slot of the LEFT neighbour (own = xb + 1); lanes without a block row read slots 0..2 (their
        // matrices are zero, the products must only stay finite)
        const int xb = CL ? (active ? slot - 1 : 0) : slot;
        const int h = lik % H, g = lik / H;             // column group, row group (matrix mapping)

        // ---- shared memory carve-up ----------------------------------------------------------------
        float* kn = smem + ((active || lhalo || rhalo) ? slot : 0) * KN;   // this knot's region
        float* xp = smem + nreg * KN;                      // p  with one-knot halos  [(NKP+2) VS]
        float* xt = xp + (NKP + 2) * VS;                   // t  with one-knot halos  [(NKP+2) VS]
        float* xr = xt + (NKP + 2) * VS;                   // r  (own knot only)      [NKP VS]
        float* xz = xr + NKP * VS;                         // z  (own knot only)      [NKP VS]
        constexpr int NRED = CL ? L::red_cl(NT) : 32;
        float* redA = xz + NKP * VS;                       // per-warp partials of p.Sp    [NRED]
        float* redB = redA + NRED;                         // per-warp partials of r.P^-1r [NRED]
        float* xq = redB + NRED;                           // cluster variant: r~ of the two knots across the CTA
                                                           // boundaries, [0] left, [1] right  [2 VS]
        const int nfloats = CL ? L::smem_floats_cl(q.kpc, NT) : L::smem_floats(K, NT);
        const int nred4 = CL ? (ncta * NW + 3) / 4 : NW4;  // float4 words of partials to sum
        // rows of the iterates (and of the Schur-phase row work) owned by this lane: OWN consecutive
        // positions starting at po0
        const int po0 = L::own_valid(lik, 0) ? L::pos(L::own_row(lik, 0)) : 0;
        int orow[OWN];
        bool ovr[OWN], ov[OWN];
#pragma unroll
        for (int j = 0; j < OWN; ++j) {
            ovr[j] = L::own_valid(lik, j);
            orow[j] = ovr[j] ? L::own_row(lik, j) : 0;
            ov[j] = active && ovr[j];
        }
        auto load_own = [&](const float* base, float (&v)[OWN]) {    // base = a VS-long slot in pos layout
            if constexpr (VEC) {
                const float4 t = *reinterpret_cast<const float4*>(base + po0);
                const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                for (int j = 0; j < OWN; ++j) v[j] = e[j];           // pad positions hold zeros
            } else {
#pragma unroll
                for (int j = 0; j < OWN; ++j) v[j] = ovr[j] ? base[po0 + j] : 0.f;
            }
        };
        auto store_own_if = [&](bool en, float* base, const float (&v)[OWN]) {   // entries of invalid rows are zero
            if (!en) return;
            if constexpr (VEC) {
                *reinterpret_cast<float4*>(base + po0) = make_float4(v[0], v[1], v[2], v[OWN - 1]);
            } else {
#pragma unroll
                for (int j = 0; j < OWN; ++j)
                    if (ovr[j]) base[po0 + j] = v[j];
            }
        };
        auto store_own = [&](float* base, const float (&v)[OWN]) { store_own_if(active, base, v); };
        // every thread of the system passes here: block barrier, or cluster barrier over all its CTAs
        auto sync_all = [&]() {
            if constexpr (GRID) cg::this_grid().sync();
            else if constexpr (CL) cg::this_cluster().sync();
            else __syncthreads();
        };
        // grid variant: boundary slices in global memory, [cta][kind: 0 p, 1 theta^-1 r, 2 r~][side at the receiver][VS]
        auto gslice = [&](int cta, int kind, int side) { return q.gbuf + 512 + ((cta * 3 + kind) * 2 + side) * VS; };
        // cluster variant: the boundary knots also write their slice of an iterate into the neighbour CTA's
        // halo slot (my left neighbour's slot kpc+1 / my right neighbour's slot 0) through DSMEM
        const bool pushL = CL && active && slot == 1 && crank > 0;
        const bool pushR = CL && active && slot == q.kpc && b + 1 < K;
        auto push_halo = [&](float* arr, const float (&v)[OWN]) {   // arr = xp or xt (slot 0 of the halo layout)
            if constexpr (GRID) {
                const int kind = arr == xp ? 0 : (arr == xt ? 1 : 2);
#pragma unroll
                for (int side = 0; side < 2; ++side) {   // my slice lands in my left neighbour's RIGHT halo and vice versa
                    if (!(side ? pushR : pushL)) continue;
                    float* dst = gslice(crank + (side ? 1 : -1), kind, side ? 0 : 1) + po0;
                    if constexpr (VEC) {
                        __stcg(reinterpret_cast<float4*>(dst), make_float4(v[0], v[1], v[2], v[OWN - 1]));
                    } else {
#pragma unroll
                        for (int j = 0; j < OWN; ++j)
                            if (ovr[j]) __stcg(dst + j, v[j]);
                    }
                }
            } else if constexpr (CL) {
                if (pushL || pushR) {
                    cg::cluster_group cluster = cg::this_cluster();
#pragma unroll
                    for (int side = 0; side < 2; ++side) {
                        if (!(side ? pushR : pushL)) continue;
                        float* rem = cluster.map_shared_rank(arr + (side ? 0 : q.kpc + 1) * VS, crank + (side ? 1 : -1));
                        if constexpr (VEC) {
                            *reinterpret_cast<float4*>(rem + po0) = make_float4(v[0], v[1], v[2], v[OWN - 1]);
                        } else {
#pragma unroll
                            for (int j = 0; j < OWN; ++j)
                                if (ovr[j]) rem[po0 + j] = v[j];
                        }
                    }
                }
            }
        };
        // grid variant, after the grid barrier: the boundary knots copy their neighbours' slices into the halo slots
        // (left halo = slot 0, right halo = slot after the last own knot; xq: [0] left, [1] right)
        auto pull_halo = [&](float* arr) {
            if constexpr (GRID) {
                const int kind = arr == xp ? 0 : (arr == xt ? 1 : 2);
#pragma unroll
                for (int side = 0; side < 2; ++side) {
                    if (!(side ? pushR : pushL)) continue;
                    const float* src = gslice(crank, kind, side) + po0;
                    float* dst = (kind == 2 ? arr + side * VS : arr + (side ? slot + 1 : 0) * VS) + po0;
                    if constexpr (VEC) {
                        *reinterpret_cast<float4*>(dst) = __ldcg(reinterpret_cast<const float4*>(src));
                    } else {
#pragma unroll
                        for (int j = 0; j < OWN; ++j)
                            if (ovr[j]) dst[j] = __ldcg(src + j);
                    }
                }
                __syncwarp();
            }
        };
        // grid variant: a dot product over the whole system.  Warp partials -> one partial per CTA in global memory
        // -> grid barrier -> every CTA adds the ncta partials in the same fixed order (same bits everywhere).
        auto grid_dot = [&](float warp_part, int which) -> float {
            float total = 0.f;
            if constexpr (GRID) {
                if (lane == 0) redA[warp] = warp_part;
                __syncthreads();
                if (tid == 0) {
                    float sc = 0.f;
                    for (int w = 0; w < NW; ++w) sc += redA[w];
                    __stcg(q.gbuf + which * 256 + crank, sc);
                }
                cg::this_grid().sync();
                if (warp == 0) {
                    float sl = 0.f;
                    for (int i = lane; i < ncta; i += 32) sl += __ldcg(q.gbuf + which * 256 + i);
                    sl = warp_sum32(sl);
                    if (lane == 0) redA[NRED - 1] = sl;
                }
                __syncthreads();
                total = redA[NRED - 1];
            }
            return total;
        };
        // the same two pushes as remote stores that complete on the receiver's transaction barrier `mb`
        constexpr uint32_t SIDE_BYTES = VEC ? LPK * 16u : S * 4u;     // one boundary knot's slice of an iterate
        const uint32_t nbr_bytes = CL ? ((crank > 0 ? SIDE_BYTES : 0u) + (crank + 1 < ncta ? SIDE_BYTES : 0u)) : 0u;
        auto push_slice_async = [&](float* left_dst, float* right_dst, const float (&v)[OWN], uint64_t* mb) {
            // left_dst / right_dst: where the slice lands in my LEFT / RIGHT neighbour's shared memory (same offsets
            // in every CTA), e.g. its halo slot kpc+1 / 0 of xt
            if constexpr (CL && !GRID) {
#pragma unroll
                for (int side = 0; side < 2; ++side) {
                    if (!(side ? pushR : pushL)) continue;
                    const int cta = crank + (side ? 1 : -1);
                    const uint32_t ra = mapa_u32(smem_u32((side ? right_dst : left_dst) + po0), cta);
                    const uint32_t rm = mapa_u32(smem_u32(mb), cta);
                    if constexpr (VEC) {
                        st_async_v4(ra, v[0], v[1], v[2], v[OWN - 1], rm);
                    } else {
#pragma unroll
                        for (int j = 0; j < OWN; ++j)
                            if (ovr[j]) st_async_f32(ra + 4u * j, v[j], rm);
                    }
                }
            }
        };
        auto put_partial_async = [&](float* red, float part, uint64_t* mb) {
            if constexpr (CL && !GRID) {
                if (lane < ncta)
                    st_async_f32(mapa_u32(smem_u32(red + crank * NW + warp), lane), part, mapa_u32(smem_u32(mb), lane));
            }
        };
        uint32_t par = (CL && !GRID) ? (uint32_t)parity_word[0] : 0u;   // bit 0/1/2: phase parity of mbA / mbB / mbC
        // per-warp partial of a dot product -> the partial array of every CTA of the system
        auto put_partial = [&](float* red, float part) {
            if constexpr (GRID) {
                // (grid variant reduces through grid_dot)
            } else if constexpr (CL) {
                if (lane < ncta) cg::this_cluster().map_shared_rank(red, lane)[crank * NW + warp] = part;
            } else {
                if (lane == 0) red[warp] = part;
            }
        };


        // The scatter map of C in bank-sorted order (one-CTA kernels, staged values): its first round is fetched now and
        // arrives while shared memory is cleared.
        constexpr int UP = GATO_PERM_UP;   // IIWA-50: 9 184 entries over 224 threads in two rounds of 22
        const bool permuted = !CL && staged && q.cperm != nullptr;
        int pe[UP], pf[UP];      // two rounds in flight: the next one's map loads are issued before this one's scatter
        auto load_perm = [&](int t0, int (&dst)[UP]) {
#pragma unroll
            for (int u = 0; u < UP; ++u) {
                const int t = t0 + u * NTT;
                if constexpr (GATO_PERM_PREFETCH && SCALED) {   // volatile: ptxas otherwise sinks the loads to their use
                    dst[u] = kPermPad;
                    if (t < q.ncperm) asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(dst[u]) : "l"(q.cperm + t));
                } else {
                    dst[u] = t < q.ncperm ? __ldg(q.cperm + t) : kPermPad;
                }
            }
        };
        constexpr bool PREFETCH = GATO_PERM_PREFETCH && SCALED;   // (the full one-CTA kernel has no registers to park it in: it spills)
        if (PREFETCH && permuted) load_perm(tid, pe);
        stamp(p, 0); S_STAMP(0);
        {   // zero everything: sparse patterns leave structural zeros, halos and pads must read as 0.
            // A persistent CTA's later systems with a purely diagonal G write exactly the words its first system
            // wrote in the knot regions (same pattern, reciprocals in place): only the iterates need clearing.
            const bool keep_regions = !CL && !SCALED && q.g_diag && sys >= (int)gridDim.x;   // (kernel 5 reuses the regions as FP64 scratch)
            float4* s4 = reinterpret_cast<float4*>(smem);
            for (int i = (keep_regions ? nreg * KN / 4 : 0) + tid; i < nfloats / 4; i += NTT)
                s4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        sync_all();   // cluster variant: no CTA pushes into a halo slot that its owner has yet to clear
        stamp(p, 1); S_STAMP(1);

        const float* cv = p.c + (long long)sys * p.ldc;
        // value sources: the staging area once this system's bulk copies have landed, else global memory
        const float* Gv = p.G_val + (long long)sys * p.ldG;
        const float* Cv = p.C_val + (long long)sys * p.ldC;
        const float* gv = p.g + (long long)sys * p.ldg;
        if (staged) {
            const float* stC = smem + q.stage_off + opaque;
            const float* stG = stC + stage_len(q.nnzC);
            const float* stg = stG + stage_len(q.nnzG);
            Cv = stC + stage_span(Cv, q.nnzC).lead;
            Gv = stG + stage_span(Gv, q.nnzG).lead;
            gv = stg + stage_span(gv, (S + C) * K - C).lead;
            mbar_wait(mbar, (uint32_t)ctl[1]);
        }
        S_STAMP(10);

        // ==== 1. unpack: coalesced reads of the CSR values, scatter into the dense per-knot blocks ====
        // One CTA per SM means memory latency is hidden only by the loads each thread keeps in flight:
        // every round issues UB (map, value) pairs per thread before the first dependent store.
        {
            // cluster variant: only the non-zeros of this CTA's knots (left halo .. last own knot); the maps hold
            // GLOBAL knot offsets, so the local destination is the map entry minus (k0 - 1) regions
            const int kA = CL ? max(k0 - 1, 0) : 0, kB = CL ? min(k0 + q.kpc, K) : K;   // knots [kA, kB)
            const int cA = CL ? __ldg(p.cstart + kA) : 0, cB = CL ? __ldg(p.cstart + kB) : q.nnzC;
            const int gA = CL ? __ldg(p.gstart + kA) : 0, gB = CL ? __ldg(p.gstart + kB) : q.nnzG;
            const int zA = CL ? kA * (S + C) : 0, zB = CL ? min(kB * (S + C), (S + C) * K - C) : (S + C) * K - C;
            const int shift = CL ? (k0 - 1) * KN : 0;
            constexpr int UB = 16;
            if (permuted) {
                // the staged copy of C_val through a pointer the compiler knows to be shared memory: a generic load of 32
                // scattered words is split per 128-byte line before it reaches the banks (measured: 2.3 x the whole phase)
                const float* CvS = smem + q.stage_off + stage_span(p.C_val + (long long)sys * p.ldC, q.nnzC).lead;
                // 32-bit shared-space addresses by hand (ld.shared / st.shared with one LEA each): through float pointers the
                // same two accesses compiled to 19 instructions per entry (64-bit index arithmetic, selects for the padding)
                const uint32_t src32 = smem_u32(CvS), dst32 = smem_u32(smem);
                auto scatter = [&](const int (&e)[UP]) {
                    float v[UP];
#pragma unroll
                    for (int u = 0; u < UP; ++u)   // (a padding entry is kPermPad: sign bit set, index 0 -- read, never stored)
                        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v[u]) : "r"(src32 + 4u * (uint32_t)(e[u] & ((1 << kPermShift) - 1))));
#pragma unroll
                    for (int u = 0; u < UP; ++u)
                        if (e[u] >= 0) asm volatile("st.shared.f32 [%0], %1;" ::"r"(dst32 + 4u * (uint32_t)(e[u] >> kPermShift)), "f"(v[u]) : "memory");
                };
                if (!PREFETCH) load_perm(tid, pe);
#pragma unroll 1
                for (int t0 = tid; t0 < q.ncperm; t0 += 2 * NTT * UP) {
                    load_perm(t0 + NTT * UP, pf);
                    scatter(pe);
                    if (t0 + 2 * NTT * UP < q.ncperm) load_perm(t0 + 2 * NTT * UP, pe);
                    scatter(pf);
                }
            } else
#pragma unroll 1
            for (int t0 = cA + tid; t0 < cB; t0 += NTT * UB) {
                int m[UB];
                float v[UB];
#pragma unroll
                for (int u = 0; u < UB; ++u) {
                    const int t = t0 + u * NTT;
                    m[u] = -1;
                    v[u] = 0.f;
                    if (t < cB) { m[u] = __ldg(q.cdst + t); v[u] = Cv[t]; }
                }
#pragma unroll
                for (int u = 0; u < UB; ++u)
                    if (m[u] >= 0) smem[m[u] - shift] = v[u];
            }
            S_STAMP(12);
            constexpr int UG = 8;
#pragma unroll 1
            for (int t0 = tid; t0 < max(gB - gA, zB - zA); t0 += NTT * UG) {
                int m[UG];
                float v[UG], w[UG];
#pragma unroll
                for (int u = 0; u < UG; ++u) {
                    const int t = t0 + u * NTT;
                    m[u] = -1;
                    v[u] = w[u] = 0.f;
                    if (gA + t < gB) { m[u] = __ldg(q.gdst + gA + t); v[u] = Gv[gA + t]; }
                    if (zA + t < zB) w[u] = gv[zA + t];
                }
#pragma unroll
                for (int u = 0; u < UG; ++u) {
                    const int t = zA + t0 + u * NTT;
                    if (m[u] >= 0) smem[(m[u] & kMapMask) - shift] = v[u] + ((m[u] & kMapDiagBit) ? p.rho : 0.0f);
                    if (t < zB) {   // g -> per-knot q (pos layout) | r  (a plan map of these destinations was tried: its load, on
                                    // the store's critical path, cost more than the division: 3 355 vs 1 774 cycles for this loop)
                        const int k = t / (S + C), o = t - k * (S + C);
                        smem[k * KN - shift + (o < S ? L::oq + L::pos(o) : L::or_ + (o - S))] = w[u];
                    }
                }
            }
        }
        S_STAMP(13);
        __syncthreads();
        if (!CL && staged && tid == 0 && sys + (int)gridDim.x < p.nsys) {
            // the staging area has been consumed (barrier above): fetch the next system's values now, so
            // they arrive while this system iterates.  Generic-proxy reads -> async-proxy writes.
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue_stage(sys + (int)gridDim.x);
        }
        stamp(p, 2); S_STAMP(2);

        // ---- Schur-phase helpers (own-row mapping) -----------------------------------------------------
        // row `row` of a row-major [.][VS] matrix in pos layout -> registers indexed by true column
        auto load_row = [&](const float* rowbase, float (&v)[S]) {
#pragma unroll
            for (int ci = 0; ci < VS / 4; ++ci) {
                const float4 t = reinterpret_cast<const float4*>(rowbase)[ci];
                const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (L::pos_valid(4 * ci + u)) v[L::pos_row(4 * ci + u)] = e[u];
            }
        };
        auto store_row = [&](float* rowbase, const float (&v)[S]) {
#pragma unroll
            for (int ci = 0; ci < VS / 4; ++ci) {
                float e[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) e[u] = L::pos_valid(4 * ci + u) ? v[L::pos_row(4 * ci + u)] : 0.f;
                reinterpret_cast<float4*>(rowbase)[ci] = make_float4(e[0], e[1], e[2], e[3]);
            }
        };
        // out[i][c] += sum_k X[row_i, k] * Y[c, k]:  X = [k][VS] with rows in pos layout (own rows are
        // one aligned load), Y = [k][VS] with its first index in pos layout (a whole column per k)
        auto rank1_S = [&](float (&out)[OWN][S], const float* X, const float* Y, int nk) {
            constexpr int UR = GATO_RANK1_UNROLL;
#pragma unroll UR
            for (int k = 0; k < nk; ++k) {
                float xv[OWN], yv[S];
                load_own(X + k * VS, xv);
                load_row(Y + k * VS, yv);
#pragma unroll
                for (int i = 0; i < OWN; ++i)
#pragma unroll
                    for (int c = 0; c < S; ++c) out[i][c] = fmaf(xv[i], yv[c], out[i][c]);
            }
        };
        // same with a C-wide output; Y = [k][LDC] plain
        auto rank1_C = [&](float (&out)[OWN][C], const float* X, const float* Y, int nk) {
#pragma unroll 1
            for (int k = 0; k < nk; ++k) {
                float xv[OWN];
                load_own(X + k * VS, xv);
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const float y = Y[k * LDC + c];
#pragma unroll
                    for (int i = 0; i < OWN; ++i) out[i][c] = fmaf(xv[i], y, out[i][c]);
                }
            }
        };

        // ==== 2./3. two passes through ONE copy of the elimination code: pass 0 inverts Q_b (and R_b),
        //      pass 1 forms block row b (phi_b, gamma_b, theta_b) and inverts theta_b.
        //   phi_b = A Q^-1, BR = B R^-1 (stage b-1), theta_b = phi A^T + Q_b^-1 + BR B^T,
        //   gamma_b = Q_b^-1 q_b - c_b + phi q_{b-1} + BR r_{b-1}          (src/gato_schur.cuh:277-385)
        float a[OWN][S];       // rows being inverted; the shuffles need the whole warp, idle lanes carry an identity
        float th[OWN][S];      // theta_b rows (kept for the S main block)
        float gam[OWN];
#pragma unroll
        for (int i = 0; i < OWN; ++i) {
            gam[i] = 0.f;
#pragma unroll
            for (int j = 0; j < S; ++j) th[i][j] = 0.f;
        }
#pragma unroll 1
        for (int pass = 0; pass < 2; ++pass) {
            if constexpr (XW) {
                if (pass == 1 && warp == NW) {   // the extra warp: every knot's A | B | Q^-1 | R^-1 | q | r is final (pass 0 inverted Q and R)
                    using SLX = ScaledLayout<S, C>;
                    constexpr int W4 = SLX::REG / 4, Q0 = L::oQ / 4, Q1 = L::oR / 4;
                    float* scrx = q.scr + (long long)sys * q.scr_ld;
                    const bool skipQ = q.g_diag != 0;
#pragma unroll 2
                    for (int k = 0; k < K; ++k) {
                        const float4* src = reinterpret_cast<const float4*>(smem + k * KN);
                        float4* dst = reinterpret_cast<float4*>(scrx + (long long)k * SLX::KSCR);
#pragma unroll
                        for (int o0 = 0; o0 < W4; o0 += 32) {
                            const int o = o0 + lane;
                            if (o < W4 && !(skipQ && o >= Q0 && o < Q1)) dst[o] = src[o];
                        }
                        if (skipQ && lane < S)     // of a diagonal Q^-1 only the diagonal
                            scrx[(long long)k * SLX::KSCR + L::oQ + lane * VS + L::pos(lane)] = smem[k * KN + L::oQ + lane * VS + L::pos(lane)];
                    }
                }
            }
            // lanes with nothing to invert in this pass carry an identity through the warp-wide elimination
            const bool inv = pass == 0 ? (active || lhalo) : active;
#pragma unroll
            for (int i = 0; i < OWN; ++i)
#pragma unroll
                for (int j = 0; j < S; ++j) a[i][j] = (!inv && MS::row(lik, i) == j) ? 1.f : 0.f;
            if (pass == 0) {
                if (active && b == 0) {   // block row 0 keeps raw Q0 as theta_0^-1 (src/gato_schur.cuh:75-81): stash it
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (ovr[i]) {
#pragma unroll
                            for (int ci = 0; ci < VS / 4; ++ci)
                                reinterpret_cast<float4*>(kn + L::oPhi + orow[i] * VS)[ci] =
                                    reinterpret_cast<const float4*>(kn + L::oQ + orow[i] * VS)[ci];
                        }
                }
                if (q.g_diag) {   // diagonal cost blocks: the inverses are reciprocals (same bits as the elimination)
                    if (inv) {
#pragma unroll
                        for (int i = 0; i < OWN; ++i)
                            if (ovr[i]) {
                                float* d = kn + L::oQ + orow[i] * VS + po0 + i;
                                *d = 1.0f / *d;
                            }
                        if (b < K - 1) {
#pragma unroll
                            for (int i = 0; i < OWNC; ++i) {
                                const int row = MC::row(lik, i);
                                if (row >= 0) {
                                    float* d = kn + L::oR + row * LDC + row;
                                    *d = 1.0f / *d;
                                }
                            }
                        }
                    }
                    __syncthreads();
                    stamp(p, 3); S_STAMP(3);
                    continue;
                }
                if (inv) {
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (ovr[i]) load_row(kn + L::oQ + orow[i] * VS, a[i]);
                }
            } else if (active || rhalo) {
                if (b == 0) {
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (ovr[i]) {
                            float qv[S];
                            load_row(kn + L::oQ + orow[i] * VS, a[i]);     // theta_0 = Q0^-1 (:26-149)
                            load_row(kn + L::oq, qv);
                            float acc = 0.f;
#pragma unroll
                            for (int k = 0; k < S; ++k) acc = fmaf(a[i][k], qv[k], acc);
                            gam[i] = acc;                                  // c[0:S] is ignored
                        }
                } else {
                    const float* prev = kn - KN;
                    float phi[OWN][S], br[OWN][C];
#pragma unroll
                    for (int i = 0; i < OWN; ++i) {
#pragma unroll
                        for (int c = 0; c < S; ++c) phi[i][c] = 0.f;
#pragma unroll
                        for (int c = 0; c < C; ++c) br[i][c] = 0.f;
                    }
                    if (q.g_diag) {
#pragma unroll
                        for (int c = 0; c < S; ++c) {
                            float xv[OWN];
                            load_own(prev + L::oA + c * VS, xv);
                            const float qi = prev[L::oQ + c * VS + L::pos(c)];
#pragma unroll
                            for (int i = 0; i < OWN; ++i) phi[i][c] = xv[i] * qi;
                        }
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            float xv[OWN];
                            load_own(prev + L::oB + c * VS, xv);
                            const float ri = prev[L::oR + c * LDC + c];
#pragma unroll
                            for (int i = 0; i < OWN; ++i) br[i][c] = xv[i] * ri;
                        }
                    } else {
                        rank1_S(phi, prev + L::oA, prev + L::oQ, S);
                        rank1_C(br, prev + L::oB, prev + L::oR, C);
                    }
#pragma unroll
                    for (int c = 0; c < S; ++c) {
                        float col[OWN];
#pragma unroll
                        for (int i = 0; i < OWN; ++i) col[i] = phi[i][c];
                        store_own_if(true, kn + L::oPhi + c * VS, col);
                    }
                    if (active) {   // a right-halo knot stops here: only its phi is needed (S_right of the last own knot)
                        float qv[S], qp[S];
                        load_row(kn + L::oq, qv);
                        load_row(prev + L::oq, qp);
#pragma unroll
                        for (int i = 0; i < OWN; ++i)
                            if (ovr[i]) {
                                float qrow[S];
                                load_row(kn + L::oQ + orow[i] * VS, qrow);
                                float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
                                for (int k = 0; k < S; ++k) {
                                    a0 = fmaf(qrow[k], qv[k], a0);
                                    a1 = fmaf(phi[i][k], qp[k], a1);
                                    a[i][k] = 0.f;
                                }
#pragma unroll
                                for (int k = 0; k < C; ++k) a2 = fmaf(br[i][k], prev[L::or_ + k], a2);
                                gam[i] = (a0 - __ldg(cv + b * S + orow[i])) + (a2 + a1);
                            }
                    }
                    if (active) rank1_S(a, kn + L::oPhi, prev + L::oA, S);   // phi A^T
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (active && ovr[i]) {
                            float qrow[S];
                            load_row(kn + L::oQ + orow[i] * VS, qrow);
#pragma unroll
                            for (int c = 0; c < S; ++c) a[i][c] += qrow[c];   // + Q_b^-1
                        }
                    if (active) {
#pragma unroll
                        for (int k = 0; k < C; ++k) {                          // + BR B^T
                            float yv[S];
                            load_row(prev + L::oB + k * VS, yv);
#pragma unroll
                            for (int i = 0; i < OWN; ++i)
#pragma unroll
                                for (int c = 0; c < S; ++c) a[i][c] = fmaf(br[i][k], yv[c], a[i][c]);
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < OWN; ++i)
#pragma unroll
                    for (int j = 0; j < S; ++j) th[i][j] = a[i][j];
            }
            if (!(SCALED && pass == 1)) gj_invert_knot<MS, LPK>(a, lik, lane);   // kernel 5 never needs theta^-1
            if (pass == 0) {
                if (inv) {
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (ovr[i]) store_row(kn + L::oQ + orow[i] * VS, a[i]);
                }
                const bool hasR = inv && b < K - 1;
                float c2[OWNC][C];
#pragma unroll
                for (int i = 0; i < OWNC; ++i) {
                    const int row = MC::row(lik, i);
#pragma unroll
                    for (int j = 0; j < C; ++j)
                        c2[i][j] = (hasR && row >= 0) ? kn[L::oR + row * LDC + j] : (!hasR && j == row ? 1.f : 0.f);
                }
                gj_invert_knot<MC, LPK>(c2, lik, lane);
#pragma unroll
                for (int i = 0; i < OWNC; ++i) {
                    const int row = MC::row(lik, i);
                    if (hasR && row >= 0) {
#pragma unroll
                        for (int j = 0; j < C; ++j) kn[L::oR + row * LDC + j] = c2[i][j];
                    }
                }
                __syncthreads();   // every stage's A, B, Q^-1, R^-1, q, r are in shared memory
                stamp(p, 3); S_STAMP(3);
            }
        }
        // a = theta_b^-1 rows; block row 0 uses raw Q0 instead (the reference's Pinv[0] = Q0)
        if (active && b == 0) {
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (ovr[i]) load_row(kn + L::oPhi + orow[i] * VS, a[i]);
        }
        stamp(p, 4); S_STAMP(4);
        __syncthreads();   // phi_b of every block row is in shared memory

        if constexpr (SCALED) {
            constexpr bool kXW = XW;
#include "gato_kernel_v5_schur.inc"
        } else {
        // ==== 4. iteration matrices into registers, matrix mapping ====================================
        // Lane (g, h) holds rows g*R .. g*R+R-1 of [S_left | S_main | S_right] restricted to its column
        // group, and the same rows of theta^-1.  Local column order (H == 2): [side knot (VS positions:
        // the left knot for h = 0, the right knot for h = 1) | own half of the main block (HALFP)];
        // (H == 1): [left | main | right], VS positions each.  Pad positions are never touched.
        constexpr int NLOC = (H == 2) ? VS + HALFP : 3 * VS;
        constexpr int MAIN0 = VS;                      // first local column of the main block (both layouts)
        float Sm[R][NLOC], Tm[R][HALFP];
        // register row i of an h = 1 lane holds matrix row g*R + hrow(i) (h = 0: g*R + i)
        auto hrow = [](int i) { return (i + OWN < R) ? i + OWN : (i >= OWN ? i - OWN : i); };
        {
            // own-row mapping (OWN rows x all columns) -> matrix mapping (R rows x own column half):
            // the two lanes of a row group swap the quadrants they do not need
            auto to_matrix = [&](const float (&own)[OWN][S], float (&mm)[R][HALFP]) {
#pragma unroll
                for (int i = 0; i < R; ++i)
#pragma unroll
                    for (int l = 0; l < HALFP; ++l) mm[i][l] = 0.f;
                if constexpr (H == 2) {
#pragma unroll
                    for (int j = 0; j < OWN; ++j)
#pragma unroll
                        for (int l = 0; l < HALFU; ++l) {
                            const float send = h ? own[j][l] : own[j][HALFU + l];
                            const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
                            mm[j][l] = h ? recv : own[j][l];                                   // rows [0, OWN)
                            if (OWN + j < R) mm[OWN + j][l] = h ? own[j][HALFU + l] : recv;    // rows [OWN, R)
                        }
                    // h = 1 keeps its register rows i and OWN + i swapped (hrow), so that the row pair a lane keeps /
                    // sends in `scatter` sits in the same registers for both column groups (no selects in the loop)
#pragma unroll
                    for (int i = 0; i + OWN < R; ++i)
#pragma unroll
                        for (int l = 0; l < HALFU; ++l) {
                            const float lo = mm[i][l], hi = mm[OWN + i][l];
                            mm[i][l] = h ? hi : lo;
                            mm[OWN + i][l] = h ? lo : hi;
                        }
                } else {
#pragma unroll
                    for (int j = 0; j < OWN; ++j)
#pragma unroll
                        for (int l = 0; l < S; ++l) mm[j][l] = own[j][l];
                }
            };
#pragma unroll
            for (int i = 0; i < R; ++i)
#pragma unroll
                for (int l = 0; l < NLOC; ++l) Sm[i][l] = 0.f;
            {
                float mm[R][HALFP];
                to_matrix(th, mm);
#pragma unroll
                for (int i = 0; i < R; ++i)
#pragma unroll
                    for (int l = 0; l < HALFP; ++l) Sm[i][MAIN0 + l] = mm[i][l];
            }
            to_matrix(a, Tm);

            const bool hasL = active && b > 0, hasR = active && b + 1 < K;
            const float* phiL = kn + L::oPhi;            // phi_b      : S[b,left]  = phi_b
            const float* phiR = kn + KN + L::oPhi;       // phi_{b+1}  : S[b,right] = phi_{b+1}^T
            if constexpr (H == 2) {
                if (h == 0) {
                    if (hasL) {   // phi_b[rr, c] at c*VS + pos(rr): the R rows of group g are contiguous
#pragma unroll
                        for (int c = 0; c < S; ++c) {
#pragma unroll
                            for (int ci = 0; ci < HALFP / 4; ++ci) {
                                const float4 t = reinterpret_cast<const float4*>(phiL + c * VS + g * HALFP)[ci];
                                const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                                for (int u = 0; u < 4; ++u)
                                    if (4 * ci + u < R) Sm[4 * ci + u][L::pos(c)] = e[u];
                            }
                        }
                    }
                } else if (hasR) {   // phi_{b+1}[c, rr] at rr*VS + pos(c): a whole row of S[b,right] is contiguous
#pragma unroll
                    for (int i = 0; i < R; ++i) {
#pragma unroll
                        for (int ci = 0; ci < VS / 4; ++ci) {
                            const float4 t = reinterpret_cast<const float4*>(phiR + (g * R + hrow(i)) * VS)[ci];
                            const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                if (L::pos_valid(4 * ci + u)) Sm[i][4 * ci + u] = e[u];
                        }
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < R; ++i) {
                    const int rr = g * R + i;
                    const bool rv = rr < S;
#pragma unroll
                    for (int l = 0; l < S; ++l) {
                        if (rv && hasL) Sm[i][l] = phiL[l * VS + rr];
                        if (rv && hasR) Sm[i][2 * VS + l] = phiR[rr * VS + l];
                    }
                }
            }
        }
        stamp(p, 5);

        // ---- iterate plumbing ------------------------------------------------------------------------
        auto scatter = [&](const float (&acc)[R], float (&out)[OWN]) {   // combine the H column groups
            if (H == 2) {
#pragma unroll
                for (int j = 0; j < OWN; ++j) {   // h = 0 keeps rows [0, OWN), h = 1 keeps rows [OWN, R)
                    if (OWN + j < R) {            // registers j / OWN + j = the row kept / sent (see hrow)
                        out[j] = acc[j] + __shfl_xor_sync(0xffffffffu, acc[OWN + j], 1);
                    } else {                      // a row only h = 0 keeps: h = 1 holds its other half in register j
                        const float recv = __shfl_xor_sync(0xffffffffu, acc[j], 1);
                        out[j] = h ? 0.f : acc[j] + recv;
                    }
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
                const float4* side = reinterpret_cast<const float4*>(x + (xb + 2 * h) * VS);
                const float4* mainh = reinterpret_cast<const float4*>(x + (xb + 1) * VS + h * HALFP);
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
                const float4* x4 = reinterpret_cast<const float4*>(x + xb * VS);
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
                const float4* side = reinterpret_cast<const float4*>(x + (xb + 2 * h) * VS);
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
                const float4* xl = reinterpret_cast<const float4*>(x + xb * VS);
                const float4* xr4 = reinterpret_cast<const float4*>(x + (xb + 2) * VS);
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
        float* xr_k = xr + slot * VS;
        float* xz_k = xz + slot * VS;
        float* xt_k = xt + (xb + 1) * VS;
        float* xp_k = xp + (xb + 1) * VS;
        // r~ = P^-1 r = theta^-1 (r - O theta^-1 r); contains one block barrier.  Block-Jacobi (p.precond = 1, the
        // reference's SS_PRECON = 0, src/gato_schur.cuh:407-429 without :496): r~ = theta^-1 r, the knot's own
        // lanes only -- no exchange, no barrier.
        auto precond = [&](const float (&r_)[OWN], float (&out)[OWN]) {
            float t[OWN], o[OWN], z[OWN];
            store_own(xr_k, r_);
            __syncwarp();
            if (p.precond) {
                apply_T(xr_k, out);
                return;
            }
            apply_T(xr_k, t);
            store_own(xt_k, t);
            if constexpr (GRID) {
                push_halo(xt, t);
                sync_all();
                pull_halo(xt);
            } else if constexpr (CL) {
                push_slice_async(xt + (q.kpc + 1) * VS, xt, t, mbB);
                __syncthreads();                                   // this CTA's own slices
                if (tid == 0) mbar_expect_tx(mbB, nbr_bytes);      // + the neighbours' boundary slices
                mbar_wait_cluster(mbB, (par >> 1) & 1u);
                par ^= 2u;
            } else {
                __syncthreads();
            }
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
            push_halo(xp, lam);
            sync_all();
            pull_halo(xp);
            apply_S(xp, ups);
#pragma unroll
            for (int j = 0; j < OWN; ++j) rr[j] = ov[j] ? gam[j] - ups[j] : 0.f;
            sync_all();
        }
        precond(rr, rt);
        float wpart = 0.f;   // this warp's partial of the dot product in flight (the grid variant reduces it later)
        {
            float part = 0.f;
#pragma unroll
            for (int j = 0; j < OWN; ++j) part = fmaf(rr[j], rt[j], part);
            part = warp_sum32(part);
            put_partial(redB, part);
            wpart = part;
        }
#pragma unroll
        for (int j = 0; j < OWN; ++j) pp[j] = rt[j];
        store_own(xp_k, pp);
        push_halo(xp, pp);
        float eta;
        if constexpr (GRID) {
            eta = grid_dot(wpart, 1);            // its grid barrier also publishes the p slices
            pull_halo(xp);
        } else {
            sync_all();
            eta = block_partials_sum(redB, nred4);
        }
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
                if constexpr (CL && !GRID) put_partial_async(redA, part, mbA);
                if constexpr (!CL) put_partial(redA, part);
                wpart = part;
            }
            float v;
            if constexpr (GRID) {
                v = grid_dot(wpart, 0);
            } else if constexpr (CL) {
                if (tid == 0) mbar_expect_tx(mbA, (uint32_t)(ncta * NW) * 4u);
                mbar_wait_cluster(mbA, par & 1u);
                par ^= 1u;
                v = block_partials_sum(redA, nred4);
            } else {
                __syncthreads();
                v = block_partials_sum(redA, nred4);
            }
            const float alpha = fast_div(eta, v);
#pragma unroll
            for (int j = 0; j < OWN; ++j) {
                lam[j] = fmaf(alpha, pp[j], lam[j]);
                rr[j] = fmaf(-alpha, ups[j], rr[j]);
            }
            const float eta_seed = rcp_seed(eta);   // for beta = eta_new / eta: eta is known before the reduction below
            precond(rr, rt);
            // cluster variant: r~ of the boundary knots travels with the reduction (the neighbour CTA then forms p_new
            // of its halo knot itself -- same fmaf, same bits -- so the p exchange needs only a block barrier)
            push_slice_async(xq + VS, xq, rt, mbC);
            if constexpr (GRID) push_halo(xq, rt);
            {
                float part = 0.f;
#pragma unroll
                for (int j = 0; j < OWN; ++j) part = fmaf(rr[j], rt[j], part);
                part = warp_sum32(part);
                if constexpr (CL && !GRID) put_partial_async(redB, part, mbC);
                if constexpr (!CL) put_partial(redB, part);
                wpart = part;
            }
            float eta_new;
            if constexpr (GRID) {
                eta_new = grid_dot(wpart, 1);
                pull_halo(xq);
            } else if constexpr (CL) {
                if (tid == 0) mbar_expect_tx(mbC, (uint32_t)(ncta * NW) * 4u + nbr_bytes);
                mbar_wait_cluster(mbC, (par >> 2) & 1u);
                par ^= 4u;
                eta_new = block_partials_sum(redB, nred4);
            } else {
                __syncthreads();
                eta_new = block_partials_sum(redB, nred4);
            }
            eta_last = eta_new;
            if (fabsf(eta_new) < p.exit_tol) { iters = it; break; }
            const float beta = fast_div_seeded(eta_new, eta, eta_seed);
#pragma unroll
            for (int j = 0; j < OWN; ++j) pp[j] = fmaf(beta, pp[j], rt[j]);
            store_own(xp_k, pp);
            eta = eta_new;
            if constexpr (CL) {
#pragma unroll
                for (int side = 0; side < 2; ++side) {
                    if (!(side ? pushR : pushL)) continue;
                    float* ph = xp + (side ? slot + 1 : 0) * VS;      // halo slot holding the neighbour's p
                    const float* qh = xq + (side ? VS : 0);           // its r~, pushed before the barrier above
#pragma unroll
                    for (int j = 0; j < OWN; ++j)
                        if (VEC || ovr[j]) ph[po0 + j] = fmaf(beta, ph[po0 + j], qh[po0 + j]);
                }
                __syncthreads();
            } else {
                sync_all();
            }
        }
        stamp(p, 8);

        sys = ctl[0];   // re-read: nothing about the loop is kept in registers across the PCG loop
        // ==== 6. dz = G^-1 (g - C^T lambda)  (src/gato_schur.cuh:759-867) =============================
        store_own(xp_k, lam);
        push_halo(xp, lam);
        if (active) {
#pragma unroll
            for (int j = 0; j < OWN; ++j)
                if (ov[j]) p.lambda_out[(long long)sys * p.ldlo + b * S + orow[j]] = lam[j];
        }
        sync_all();
        pull_halo(xp);
        if (active) {
            float ln[S];                                    // lambda_{b+1}; zero past the horizon
            load_row(xp + (xb + 2) * VS, ln);
            float tb[OWN];
#pragma unroll
            for (int j = 0; j < OWN; ++j) {
                tb[j] = 0.f;
                if (ovr[j]) {   // (A^T lambda_{b+1})[row] = sum_k A[k,row] lambda_{b+1}[k]: column `row` of A is contiguous
                    float col[S];
                    load_row(kn + L::oA + orow[j] * VS, col);
                    float acc = 0.f;
#pragma unroll
                    for (int k = 0; k < S; ++k) acc = fmaf(col[k], ln[k], acc);
                    tb[j] = kn[L::oq + po0 + j] - (lam[j] + acc);
                }
            }
            store_own(xz_k, tb);
#pragma unroll
            for (int j = 0; j < OWNC; ++j) {
                const int cr = MC::row(lik, j);
                if (cr >= 0) {
                    float col[S];
                    load_row(kn + L::oB + cr * VS, col);
                    float acc = 0.f;
#pragma unroll
                    for (int k = 0; k < S; ++k) acc = fmaf(col[k], ln[k], acc);
                    xr_k[cr] = kn[L::or_ + cr] - acc;
                }
            }
        }
        __syncwarp();
        if (active) {
            float* dz = p.dz_out + (long long)sys * p.lddz + (long long)b * (S + C);
            float tbv[S];
            load_row(xz_k, tbv);
#pragma unroll
            for (int j = 0; j < OWN; ++j)
                if (ovr[j]) {
                    float qrow[S];
                    load_row(kn + L::oQ + orow[j] * VS, qrow);
                    float acc = 0.f;
#pragma unroll
                    for (int k = 0; k < S; ++k) acc = fmaf(qrow[k], tbv[k], acc);
                    dz[orow[j]] = acc;
                }
            if (b < K - 1) {
#pragma unroll
                for (int j = 0; j < OWNC; ++j) {
                    const int cr = MC::row(lik, j);
                    if (cr >= 0) {
                        float acc = 0.f;
#pragma unroll
                        for (int k = 0; k < C; ++k) acc = fmaf(kn[L::oR + cr * LDC + k], xr_k[k], acc);
                        dz[S + cr] = acc;
                    }
                }
            }
        }
        stamp(p, 9);
        if constexpr (CL && !GRID) {
            if (tid == 0) parity_word[0] = (int)par;
        }
        if (tid == 0 && crank == 0) {
            p.iters_out[sys] = iters;
            if (p.eta_out != nullptr) p.eta_out[sys] = fabsf(eta_last);
        }
        }   // !SCALED
        sys = ctl[0];
        __syncthreads();   // dz has read the knot regions and the iterates: the next system may clear them
        if (tid == 0) {
            ctl[0] = sys + (int)gridDim.x / ncta;
            ctl[1] ^= 1;
        }
        __syncthreads();
        S_STAMP(11);
    }
#ifdef GATO_S_TRACE
    if (SCALED && blockIdx.x == 0 && threadIdx.x == 0)
        for (int i = 0; i < 24; ++i) (reinterpret_cast<long long*>(q.gbuf) + 256)[i] = s_tr[i];
#endif
}

}  // namespace gato
