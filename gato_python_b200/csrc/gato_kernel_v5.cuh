// gato_kernel_v5.cuh -- SECOND HALF OF KERNEL 5: PCG on the scaled system S^ = I + O^ and the dz recovery, sm_100a.
//
// The first half (gato_fused_solve_v2<.., SCALED = true>, gato_kernel_v5_schur.inc) leaves per knot, in global scratch
// (ScaledLayout): psi_b = M_b phi_b M_{b-1}^T, gamma^_b, M_b and a copy of the knot region.
//
// Mapping: TWO lanes per knot.  Lane (b, g) keeps rows g S/2 .. of psi_b -- 7 x 14 = 98 words for S = 14, against
// 196 matrix words per lane and four lanes per knot in the one-CTA kernel -- and uses every word twice per product,
//     u_b     += psi_b   x_{b-1}      (its rows of this knot: complete in the lane, no combine)
//     w_{b-1} += psi_b^T x_b          (partial over its rows; the LEFT knot's rows: O^ is symmetric, S^[b-1, b] = psi_b^T)
// so an IIWA-50 system is four warps, a CTA needs < 170 registers per thread and ~15 KB of shared memory, and THREE
// systems are resident per SM (one warp of each per sub-partition): one system's reduction chains and barriers hide
// behind the others' arithmetic.  Per iteration and lane: 4 x 98 FFMA against 8 + 4 128-bit shared-memory loads and
// 14 + 10 shuffles -- the first version of this kernel (four lanes per knot, 7 x 7 blocks) issued twice the shared
// memory wavefronts and shuffles per knot and sat at 63 % of the LSU data pipe (profiles/r02_k5_*).
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

#ifndef GATO_V5_PREFETCH
#define GATO_V5_PREFETCH 0
#endif
#ifndef GATO_V5_CTAS
#define GATO_V5_CTAS 3
#endif

// Development builds (-DGATO_V5_TRACE): thread 0 of the CTA that takes system 0 stamps clock64() at eleven points of
// each of the first 8 iterations into the buffer behind the work counter (tests/checks/k5_trace.py reads it).
#ifdef GATO_V5_TRACE
#define V5_STAMP(i) do { if (trace_on && it < 8) trace[it * 12 + (i)] = clock64(); } while (0)
#else
#define V5_STAMP(i) do { } while (0)
#endif

namespace gato {

// ONE_RED: the single-reduction form of the same iteration (gato_plan_set_pcg_form).
template <int S, int C, int NT, bool ONE_RED = false>
__global__ void __launch_bounds__(NT, (GATO_V5_CTAS * 128) / NT) gato_pcg_scaled(const SolveParams2 q) {
    using L = Layout2<S, C>;
    using SL = ScaledLayout<S, C>;
    constexpr int VS = L::VS, HALFP = L::HALFP, HALFU = L::HALFU, LDC = L::LDC;
    static_assert(L::H == 2 && HALFP == 8 && HALFU == 7, "the two 128-bit loads per half-slot below are written for S = 14");
    constexpr int HU = HALFU;                // rows per lane, and columns per half
    constexpr int NW = NT / 32, NKP = NT / 2, NW4 = (NW + 3) / 4;
    // An iterate is stored as TWO arrays, one per half of the knots' entries ([slot][HALFP] each, knot k at slot k + 1),
    // the second 16 bytes further round the banks: the two lanes of a knot read different halves in the same
    // instruction, and with the halves interleaved in one array every 128-bit access was a 2-way bank conflict.
    constexpr int XH = (NKP + 2) * HALFP;
    constexpr int XS = 2 * XH + 4;
    constexpr int CU = (C + 1) / 2;          // control rows per lane in the dz phase
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
    const int b = tid >> 1, g = tid & 1;                  // knot, row half
    const bool active = b < K;
    auto half = [](float* x, int slot, int hh) { return x + hh * (XH + 4) + slot * HALFP; };   // HU entries + a zero pad
    auto ld8 = [](const float* a, float (&v)[8]) {
        const float4 t0 = reinterpret_cast<const float4*>(a)[0], t1 = reinterpret_cast<const float4*>(a)[1];
        v[0] = t0.x; v[1] = t0.y; v[2] = t0.z; v[3] = t0.w; v[4] = t1.x; v[5] = t1.y; v[6] = t1.z; v[7] = t1.w;
    };
    auto st8 = [](float* a, const float (&v)[8]) {
        reinterpret_cast<float4*>(a)[0] = make_float4(v[0], v[1], v[2], v[3]);
        reinterpret_cast<float4*>(a)[1] = make_float4(v[4], v[5], v[6], v[7]);
    };
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

    // Systems are handed out by an atomic counter (reset by the first half, which always runs before this kernel on the
    // same stream): iteration counts differ between systems, and a static split left the slowest CTA ~7 % behind.
    int* next_sys = reinterpret_cast<int*>(smem + NFLOATS);
    int* counter = reinterpret_cast<int*>(q.gbuf);
#pragma unroll 1
    while (true) {
        if (tid == 0) *next_sys = atomicAdd(counter, 1);
        __syncthreads();
        const int sys = *next_sys;
        if (sys >= p.nsys) break;
        const float* scr = q.scr + (long long)sys * q.scr_ld;
        const float* kg = scr + (long long)(active ? b : 0) * SL::KSCR;     // this knot's scratch
        {   // halos, pads and the never-written slots must read as zero
            float4* s4 = reinterpret_cast<float4*>(smem);
            for (int i = tid; i < NFLOATS / 4; i += NT) s4[i] = make_float4(0.f, 0.f, 0.f, 0.f);   // (not next_sys: behind NFLOATS)
        }
        // psi[i][c']: row g HU + i, column order [own half g | other half] (so the lane keeps the first HU partial
        // sums of w and sends the last HU, whatever g)
        float psi[HU][2 * HU];
        {
            const float4* src = reinterpret_cast<const float4*>(kg + SL::oPSI + g * SL::PSIL);
            float raw[SL::PSIL];
#pragma unroll
            for (int e = 0; e < SL::PSIL / 4; ++e) {
                const float4 t = active ? __ldcg(src + e) : make_float4(0.f, 0.f, 0.f, 0.f);
                raw[4 * e] = t.x; raw[4 * e + 1] = t.y; raw[4 * e + 2] = t.z; raw[4 * e + 3] = t.w;
            }
#pragma unroll
            for (int i = 0; i < HU; ++i)
#pragma unroll
                for (int c = 0; c < 2 * HU; ++c)   // column c of the lane's order = block c / HU, column c % HU of it
                    psi[i][c] = raw[(c / HU) * SL::PSIB + i * HU + c % HU];
        }
        if (GATO_V5_PREFETCH && active) {   // what the dz phase will read (A, B, R^-1, q, r, M; Q^-1 unless diagonal): start it towards L2 now
            const char* kb = reinterpret_cast<const char*>(kg);
            auto pf = [&](int lo, int hi) {   // float offsets; the two lanes of the knot take alternate 128-byte lines
                for (int o = lo * 4 + g * 128; o < hi * 4; o += 256) asm volatile("prefetch.global.L2 [%0];" ::"l"(kb + o));
            };
            pf(L::oA, L::oQ);
            if (!q.g_diag) pf(L::oQ, L::oR);
            pf(L::oR, SL::REG);
            pf(SL::oM, SL::oGAM);
        }
        float lam[8], rr[8], pp[8];   // entry 7 is the pad: always zero
        {
            float t[8];
            if (active) {
                // gamma^: from the scratch, or (matrices shared by the call) from the per-system array gato_gamma_scaled filled
                const float* gsrc = (q.gam != nullptr ? q.gam + (long long)sys * q.gam_ld + b * VS : kg + SL::oGAM) + g * HALFP;
                const float4 t0 = __ldcg(reinterpret_cast<const float4*>(gsrc));
                const float4 t1 = __ldcg(reinterpret_cast<const float4*>(gsrc) + 1);
                t[0] = t0.x; t[1] = t0.y; t[2] = t0.z; t[3] = t0.w; t[4] = t1.x; t[5] = t1.y; t[6] = t1.z; t[7] = t1.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) t[j] = 0.f;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) { rr[j] = t[j]; lam[j] = 0.f; pp[j] = 0.f; }
        }
        __syncthreads();

        float* xp_k = half(xp, b + 1, g);
        float* xr_k = half(xr, b + 1, g);
        // u = psi_b x_{b-1} (this lane's rows, complete) and the lane pair's w = psi_b^T x_b, of which this lane keeps
        // the half g and stores it into the LEFT knot's slot of xw.  `own` = this lane's entries of x_b.
        // (Packed FP32 FMAs -- fma.rn.f32x2, pairs along the columns -- were built here twice: 196 FFMA2 instead of
        // 392 FFMA per iteration, but ptxas adds 210-280 MOVs per iteration to keep their operands in aligned register
        // pairs at 168 registers, even with the pairs taken straight from 64-bit halves of the 128-bit loads;
        // 3.45 vs 3.50 ms per 8192 systems the first time, more spills the second.  Not kept.)
        auto apply = [&](float* x, const float (&own)[8], float* xw, float (&u)[8]) {
            float e0[8], e1[8];
            ld8(half(x, b, g), e0);                             // knot b - 1 (slot b), its half g: columns [0, HU) of psi
            ld8(half(x, b, 1 - g), e1);                         // its other half
            float w[2 * HU];
#pragma unroll
            for (int i = 0; i < HU; ++i) u[i] = 0.f;
            u[7] = 0.f;
#pragma unroll
            for (int c = 0; c < 2 * HU; ++c) w[c] = 0.f;
#pragma unroll
            for (int i = 0; i < HU; ++i)
#pragma unroll
                for (int c = 0; c < 2 * HU; ++c) {
                    u[i] = fmaf(psi[i][c], c < HU ? e0[c] : e1[c - HU], u[i]);
                    w[c] = fmaf(psi[i][c], own[i], w[c]);
                }
            float wo[8];
#pragma unroll
            for (int j = 0; j < HU; ++j) wo[j] = w[j] + __shfl_xor_sync(0xffffffffu, w[HU + j], 1);
            wo[7] = 0.f;
            if (active) st8(half(xw, b, g), wo);                // knot b - 1 sits at slot b
        };

        int iters = p.max_iters;
        float eta_last = 0.f;
        if constexpr (ONE_RED) {
            // Single-reduction PCG (Chronopoulos-Gear): with r~ = P^^-1 r and w = S^ r~ formed first, the two dot
            // products gamma = r.r~ and delta = r~.w are reduced TOGETHER, and
            //     beta = gamma / gamma_old,  alpha = gamma / (delta - beta gamma / alpha_old),
            //     p = r~ + beta p,  s = w + beta s (= S^ p),  lambda += alpha p,  r -= alpha s.
            // Algebraically the iteration of src/gato_pcg.cuh:300-437 (same gamma_k = r_k^T P^-1 r_k tested against
            // exit_tol after every update, same 0-based index); one shuffle tree, one partial sum and one division
            // per iteration instead of two of each.  k = number of updates done.
            float ss[8], ut[8], u1[8], wv[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) ss[j] = 0.f;
            float* xu = xp;          // r~ travels in the array the two-reduction loop uses for p
            float gamma_old = 1.f, alpha_old = 1.f;
            for (int k = 0;; ++k) {
                if (active) st8(xr_k, rr);
                __syncthreads();
                apply(xr, rr, xwr, u1);                       // u1 = (O^ r) own part; w parts -> xwr
                __syncthreads();
                ld8(half(xwr, b + 1, g), wv);
#pragma unroll
                for (int j = 0; j < HU; ++j) ut[j] = (rr[j] - u1[j]) - wv[j];
                ut[7] = 0.f;
                if (active) st8(half(xu, b + 1, g), ut);
                __syncthreads();
                apply(xu, ut, xwp, u1);                       // u1 = (O^ r~) own part; w parts -> xwp
                {
                    float pg = 0.f, pd = 0.f;
#pragma unroll
                    for (int j = 0; j < HU; ++j) {
                        pg = fmaf(rr[j], ut[j], pg);
                        pd = fmaf(ut[j], fmaf(2.f, u1[j], ut[j]), pd);
                    }
                    pg = warp_sum32(pg);
                    pd = warp_sum32(pd);
                    if (lane == 0) { redA[warp] = pg; redB[warp] = pd; }
                }
                __syncthreads();
                const float gamma = block_partials_sum(redA, NW4), delta = block_partials_sum(redB, NW4);
                eta_last = gamma;
                if (k >= 1 && fabsf(gamma) < p.exit_tol) { iters = k - 1; break; }
                if (k >= p.max_iters) break;
                const float beta = k == 0 ? 0.f : fast_div(gamma, gamma_old);
                const float alpha = fast_div(gamma, k == 0 ? delta : fmaf(-beta, fast_div(gamma, alpha_old), delta));
                ld8(half(xwp, b + 1, g), wv);
#pragma unroll
                for (int j = 0; j < HU; ++j) {
                    const float w = (ut[j] + u1[j]) + wv[j];  // (S^ r~) own part
                    pp[j] = fmaf(beta, pp[j], ut[j]);
                    ss[j] = fmaf(beta, ss[j], w);
                    lam[j] = fmaf(alpha, pp[j], lam[j]);
                    rr[j] = fmaf(-alpha, ss[j], rr[j]);
                }
                gamma_old = gamma;
                alpha_old = alpha;
            }
        } else {
        // r~0 = r - O^ r, eta = r.r~0, p = r~0
        if (active) st8(xr_k, rr);
        __syncthreads();
        float uo[8], wv[8];
        apply(xr, rr, xwr, uo);
        {
            float part = 0.f;
#pragma unroll
            for (int j = 0; j < HU; ++j) part = fmaf(rr[j], fmaf(-2.f, uo[j], rr[j]), part);
            part = warp_sum32(part);
            if (lane == 0) redB[warp] = part;
        }
        __syncthreads();
        float eta = block_partials_sum(redB, NW4);
        ld8(half(xwr, b + 1, g), wv);
#pragma unroll
        for (int j = 0; j < HU; ++j) pp[j] = (rr[j] - uo[j]) - wv[j];
        if (active) st8(xp_k, pp);
        __syncthreads();

        eta_last = eta;
#ifdef GATO_V5_TRACE
        const bool trace_on = sys == 0 && tid == 0;
        long long* trace = reinterpret_cast<long long*>(q.gbuf) + 8;
#endif
        for (int it = 0; it < p.max_iters; ++it) {
            V5_STAMP(0);
            apply(xp, pp, xwp, uo);
            V5_STAMP(1);
            {
                float part = 0.f;
#pragma unroll
                for (int j = 0; j < HU; ++j) part = fmaf(pp[j], fmaf(2.f, uo[j], pp[j]), part);
                part = warp_sum32(part);
                if (lane == 0) redA[warp] = part;
            }
            V5_STAMP(2);
            __syncthreads();
            V5_STAMP(3);
            const float v = block_partials_sum(redA, NW4);
            const float alpha = fast_div(eta, v);
            ld8(half(xwp, b + 1, g), wv);
#pragma unroll
            for (int j = 0; j < HU; ++j) {
                const float ups = (pp[j] + uo[j]) + wv[j];
                lam[j] = fmaf(alpha, pp[j], lam[j]);
                rr[j] = fmaf(-alpha, ups, rr[j]);
            }
            if (active) st8(xr_k, rr);
            const float eta_seed = rcp_seed(eta);
            V5_STAMP(4);
            __syncthreads();
            V5_STAMP(5);
            apply(xr, rr, xwr, uo);
            V5_STAMP(6);
            {
                float part = 0.f;
#pragma unroll
                for (int j = 0; j < HU; ++j) part = fmaf(rr[j], fmaf(-2.f, uo[j], rr[j]), part);
                part = warp_sum32(part);
                if (lane == 0) redB[warp] = part;
            }
            V5_STAMP(7);
            __syncthreads();
            V5_STAMP(8);
            const float eta_new = block_partials_sum(redB, NW4);
            eta_last = eta_new;
            if (fabsf(eta_new) < p.exit_tol) { iters = it; break; }
            const float beta = fast_div_seeded(eta_new, eta, eta_seed);
            ld8(half(xwr, b + 1, g), wv);
#pragma unroll
            for (int j = 0; j < HU; ++j) pp[j] = fmaf(beta, pp[j], (rr[j] - uo[j]) - wv[j]);
            if (active) st8(xp_k, pp);
            eta = eta_new;
            V5_STAMP(9);
            __syncthreads();
            V5_STAMP(10);
        }

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
