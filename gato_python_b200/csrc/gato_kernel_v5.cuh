// gato_kernel_v5.cuh -- SECOND HALF OF KERNEL 5: PCG on the scaled system S^ = I + O^ and the dz recovery, sm_100a.
//
// The first half (gato_fused_solve_v2<.., SCALED = true>, gato_kernel_v5_schur.inc) leaves per knot, in global scratch
// (ScaledLayout): psi_b = M_b phi_b M_{b-1}^T in lane-ready order, gamma^_b, M_b and a copy of the knot region.
// This kernel keeps ONE S/2 x S/2 block of psi_b per lane in registers (49 words for S = 14, against 196 matrix words
// per lane in the one-CTA kernel) and uses every word twice per product,
//     u_b     += psi_b   x_{b-1}      (rows of this knot)
//     w_{b-1} += psi_b^T x_b          (rows of the LEFT knot: O^ is symmetric, S^[b-1, b] = psi_b^T)
// so a 224-thread CTA needs < 128 registers per thread and ~14 KB of shared memory: TWO systems are resident per SM
// and one system's reduction chains and barriers hide behind the other's arithmetic.  Per iteration and lane:
// 4 x 49 FFMA + 16 shuffles + 4 block barriers, against 7 x 49 FFMA + 16 shuffles + 4 barriers before.
//
// Iteration (same update order, exit rule |r^T P^-1 r| < exit_tol and 0-based iteration index as
// src/gato_pcg.cuh:300-437; r^ = M r, p^ = M^-T p, so every scalar equals the reference's in exact arithmetic):
//     v     = p^.S^ p^ = sum_b p_b.(p_b + 2 u_b)                  (p^T O^ p = 2 sum_b p_b . psi_b p_{b-1})
//     r~    = P^^-1 r^ = r^ - O^ r^
//     eta'  = r^.r~    = sum_b r_b.(r_b - 2 u'_b)
// the w parts travel through shared memory and are consumed after the barrier of the reduction they ride on.
// Replaces in the reference: parallelPCG (src/gato_pcg.cuh:270-470) and gato_compute_dz (src/gato_schur.cuh:758-879).
#pragma once
#include "gato_kernel_v2.cuh"

#ifndef GATO_V5_CTAS
#define GATO_V5_CTAS 2
#endif

namespace gato {

template <int S, int C, int NT>
__global__ void __launch_bounds__(NT, GATO_V5_CTAS) gato_pcg_scaled(const SolveParams2 q) {
    using L = Layout2<S, C>;
    using SL = ScaledLayout<S, C>;
    using MC = MapControl<S, C>;
    constexpr int LPK = L::LPK, OWN = L::OWN, OWNC = L::OWNC, H = L::H, R = L::R, VS = L::VS, HALFP = L::HALFP,
                  HALFU = L::HALFU, LDC = L::LDC;
    static_assert(H == 2 && LPK == 4 && R == 7 && HALFU == 7 && OWN == 4 && HALFP == 8 && L::VEC_OWN,
                  "the half-swapped vector loads below are written for S = 14");
    constexpr int NW = NT / 32, NKP = NW * L::KPW, NW4 = (NW + 3) / 4;
    constexpr int XS = (NKP + 2) * VS;
    extern __shared__ __align__(16) float smem[];
    const SolveParams& p = q.p;
    const int K = p.K;
    float* xp = smem;            // p^  with one-knot halos (knot k at slot k + 1)
    float* xr = xp + XS;         // r^  likewise
    float* xwp = xr + XS;        // psi^T p parts, knot k at slot k + 1 (written by the lanes of knot k + 1)
    float* xwr = xwp + XS;       // psi^T r parts
    float* xz = xwr + XS;        // dz phase: q - lambda - A^T lambda_{b+1} per knot   [NKP VS]
    float* xc = xz + NKP * VS;   // dz phase: r - B^T lambda_{b+1} per knot            [NKP LDC]
    float* redA = xc + NKP * LDC;
    float* redB = redA + 32;
    constexpr int NFLOATS = 4 * XS + NKP * (VS + LDC) + 64;

    const int tid = (int)threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int lik = lane % LPK, b = warp * L::KPW + lane / LPK;
    const bool active = b < K;
    const int h = lik % H, g = lik / H;
    const int po0 = g * HALFP + h * OWN;                  // first of this lane's OWN consecutive positions
    bool ovr[OWN];
    int orow[OWN];
#pragma unroll
    for (int j = 0; j < OWN; ++j) {
        ovr[j] = L::own_valid(lik, j);
        orow[j] = ovr[j] ? L::own_row(lik, j) : 0;
    }
    auto ld4 = [](const float* a, float (&v)[4]) {
        const float4 t = *reinterpret_cast<const float4*>(a);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    };
    auto st4 = [](float* a, const float (&v)[4]) { *reinterpret_cast<float4*>(a) = make_float4(v[0], v[1], v[2], v[3]); };
    // row `row` of a [.][VS] matrix in pos layout -> registers indexed by true column
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
    // reduce-scatter of R partial sums over the two lanes of a pair (lane ^ mask): the lane with flag = 0 ends with
    // entries [0, OWN), its partner with [OWN, R) (+ a zero pad); the partner's registers are half-swapped
    auto scatter = [&](const float (&acc)[R], float (&out)[OWN], int flag, int mask) {
#pragma unroll
        for (int j = 0; j < OWN; ++j) {
            if (OWN + j < R) {
                out[j] = acc[j] + __shfl_xor_sync(0xffffffffu, acc[OWN + j], mask);
            } else {
                const float recv = __shfl_xor_sync(0xffffffffu, acc[j], mask);
                out[j] = flag ? 0.f : acc[j] + recv;
            }
        }
    };

#pragma unroll 1
    for (int sys = (int)blockIdx.x; sys < p.nsys; sys += (int)gridDim.x) {
        const float* scr = q.scr + (long long)sys * q.scr_ld;
        const float* kg = scr + (long long)(active ? b : 0) * SL::KSCR;     // this knot's scratch
        {   // halos, pads and the never-written slots must read as zero
            float4* s4 = reinterpret_cast<float4*>(smem);
            for (int i = tid; i < NFLOATS / 4; i += NT) s4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        float psi[R][HALFU];
        {
            float raw[SL::PSIL];
            const float4* src = reinterpret_cast<const float4*>(kg + SL::oPSI + lik * SL::PSIL);
#pragma unroll
            for (int e = 0; e < SL::PSIL / 4; ++e) {
                const float4 t = active ? __ldcg(src + e) : make_float4(0.f, 0.f, 0.f, 0.f);
                raw[4 * e] = t.x; raw[4 * e + 1] = t.y; raw[4 * e + 2] = t.z; raw[4 * e + 3] = t.w;
            }
#pragma unroll
            for (int i = 0; i < R; ++i)
#pragma unroll
                for (int j = 0; j < HALFU; ++j) psi[i][j] = raw[i * HALFU + j];
        }
        float lam[OWN], rr[OWN], pp[OWN];
        {
            const float4 t = active ? __ldcg(reinterpret_cast<const float4*>(kg + SL::oGAM + po0)) : make_float4(0.f, 0.f, 0.f, 0.f);
            rr[0] = t.x; rr[1] = t.y; rr[2] = t.z; rr[3] = t.w;     // pad positions hold zeros
        }
#pragma unroll
        for (int j = 0; j < OWN; ++j) lam[j] = 0.f;
        __syncthreads();

        float* xp_k = xp + (b + 1) * VS + po0;
        float* xr_k = xr + (b + 1) * VS + po0;
        // u = psi_b x_{b-1} (own rows, combined over the column halves) and this lane's slice of w = psi_b^T x_b, which
        // goes to the LEFT knot's slot of xw
        auto apply = [&](const float* x, float* xw, float (&uo)[OWN]) {
            const float* xl = x + b * VS + h * HALFP;           // knot b - 1, column half h
            const float* xo = x + (b + 1) * VS + g * HALFP;     // own knot, row half g
            float P[4], Q[4], P2[4], Q2[4];
            ld4(xl + (g ? 4 : 0), P);
            ld4(xl + (g ? 0 : 4), Q);
            ld4(xo + (h ? 4 : 0), P2);
            ld4(xo + (h ? 0 : 4), Q2);
            const float eL[HALFU] = {P[0], P[1], P[2], g ? Q[3] : P[3], Q[0], Q[1], Q[2]};
            const float eO[R] = {P2[0], P2[1], P2[2], h ? Q2[3] : P2[3], Q2[0], Q2[1], Q2[2]};
            float u[R], w[HALFU];
#pragma unroll
            for (int i = 0; i < R; ++i) u[i] = 0.f;
#pragma unroll
            for (int j = 0; j < HALFU; ++j) w[j] = 0.f;
#pragma unroll
            for (int i = 0; i < R; ++i)
#pragma unroll
                for (int j = 0; j < HALFU; ++j) {
                    u[i] = fmaf(psi[i][j], eL[j], u[i]);
                    w[j] = fmaf(psi[i][j], eO[i], w[j]);
                }
            float wo[OWN];
            scatter(u, uo, h, 1);
            scatter(w, wo, g, 2);
            if (active) st4(xw + b * VS + h * HALFP + g * OWN, wo);
        };

        // r~0 = r - O^ r, eta = r.r~0, p = r~0
        if (active) st4(xr_k, rr);
        __syncthreads();
        float uo[OWN], wv[OWN];
        apply(xr, xwr, uo);
        {
            float part = 0.f;
#pragma unroll
            for (int j = 0; j < OWN; ++j) part = fmaf(rr[j], fmaf(-2.f, uo[j], rr[j]), part);
            part = warp_sum32(part);
            if (lane == 0) redB[warp] = part;
        }
        __syncthreads();
        float eta = block_partials_sum(redB, NW4);
        ld4(xwr + (b + 1) * VS + po0, wv);
#pragma unroll
        for (int j = 0; j < OWN; ++j) pp[j] = (rr[j] - uo[j]) - wv[j];
        if (active) st4(xp_k, pp);
        __syncthreads();

        int iters = p.max_iters;
        float eta_last = eta;
        for (int it = 0; it < p.max_iters; ++it) {
            apply(xp, xwp, uo);
            {
                float part = 0.f;
#pragma unroll
                for (int j = 0; j < OWN; ++j) part = fmaf(pp[j], fmaf(2.f, uo[j], pp[j]), part);
                part = warp_sum32(part);
                if (lane == 0) redA[warp] = part;
            }
            __syncthreads();
            const float v = block_partials_sum(redA, NW4);
            const float alpha = fast_div(eta, v);
            ld4(xwp + (b + 1) * VS + po0, wv);
#pragma unroll
            for (int j = 0; j < OWN; ++j) {
                const float ups = (pp[j] + uo[j]) + wv[j];
                lam[j] = fmaf(alpha, pp[j], lam[j]);
                rr[j] = fmaf(-alpha, ups, rr[j]);
            }
            if (active) st4(xr_k, rr);
            const float eta_seed = rcp_seed(eta);
            __syncthreads();
            apply(xr, xwr, uo);
            {
                float part = 0.f;
#pragma unroll
                for (int j = 0; j < OWN; ++j) part = fmaf(rr[j], fmaf(-2.f, uo[j], rr[j]), part);
                part = warp_sum32(part);
                if (lane == 0) redB[warp] = part;
            }
            __syncthreads();
            const float eta_new = block_partials_sum(redB, NW4);
            eta_last = eta_new;
            if (fabsf(eta_new) < p.exit_tol) { iters = it; break; }
            const float beta = fast_div_seeded(eta_new, eta, eta_seed);
            ld4(xwr + (b + 1) * VS + po0, wv);
#pragma unroll
            for (int j = 0; j < OWN; ++j) pp[j] = fmaf(beta, pp[j], (rr[j] - uo[j]) - wv[j]);
            if (active) st4(xp_k, pp);
            eta = eta_new;
            __syncthreads();
        }

#include "gato_kernel_v5_dz.inc"
        if (tid == 0) {
            p.iters_out[sys] = iters;
            if (p.eta_out != nullptr) p.eta_out[sys] = fabsf(eta_last);
        }
        __syncthreads();   // the next system clears the arrays
    }
}

}  // namespace gato
