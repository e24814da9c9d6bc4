// gato_kernel_v5_lat.cuh -- SECOND HALF OF KERNEL 5, LATENCY VARIANT (single solves and batches that do not fill the device).
//
// Same iteration and same scratch as gato_pcg_scaled (gato_kernel_v5.cuh), mapped FOUR lanes per knot: lane (g, h) keeps
// one S/2 x S/2 block of psi_b (49 words) and a system is seven warps.  A warp's instruction stream per iteration is
// 366 instructions against 618 with two lanes per knot, and a lone warp issues one instruction every other cycle, so
// one system alone iterates ~1.6 x faster; with the device full the two-lane kernel wins (half the shared-memory
// wavefronts and shuffles per knot: this kernel sat at 63 % of the LSU data pipe, profiles/r02_ncu_hotspots_k5_pcg_4lane.txt).
// Both reduce-scatters are select-free: lanes with h = 1 keep their register ROWS half-swapped, lanes with g = 1 their
// register COLUMNS (applied when psi is loaded), and the vector loads swap their two 128-bit halves accordingly.
#pragma once
#include "gato_kernel_v2.cuh"

namespace gato {

template <int S, int C, int NT>
__global__ void __launch_bounds__(NT, 2) gato_pcg_scaled4(const SolveParams2 q) {
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
    // the half-swap permutation of R rows (or HALFU columns): [OWN .. R) first, so that the rows a lane keeps and the
    // rows it sends in a reduce-scatter sit in the same registers for both lanes of a pair
    auto hswap = [](int i) { return (i + OWN < R) ? i + OWN : (i >= OWN ? i - OWN : i); };
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
            // the scratch holds psi in the order of the throughput kernel (lane g: [block of column half g | other half],
            // each R x HALFU, row-major, natural order); this lane's block is the one of column half h
            float raw[SL::PSIB];
            const float4* src = reinterpret_cast<const float4*>(kg + SL::oPSI + g * SL::PSIL + (h ^ g) * SL::PSIB);
#pragma unroll
            for (int e = 0; e < SL::PSIB / 4; ++e) {
                const float4 t = active ? __ldcg(src + e) : make_float4(0.f, 0.f, 0.f, 0.f);
                raw[4 * e] = t.x; raw[4 * e + 1] = t.y; raw[4 * e + 2] = t.z; raw[4 * e + 3] = t.w;
            }
#pragma unroll
            for (int i = 0; i < R; ++i)
#pragma unroll
                for (int j = 0; j < HALFU; ++j) {   // register (i, j) <- row rp_h(i), column cp_g(j): see `scatter`
                    const int si = hswap(i), sj = hswap(j);
                    const float a0 = h ? raw[si * HALFU + j] : raw[i * HALFU + j];
                    const float a1 = h ? raw[si * HALFU + sj] : raw[i * HALFU + sj];
                    psi[i][j] = g ? a1 : a0;
                }
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

#include "gato_kernel_v5_dz4.inc"
        if (tid == 0) {
            p.iters_out[sys] = iters;
            if (p.eta_out != nullptr) p.eta_out[sys] = fabsf(eta_last);
        }
        __syncthreads();   // the next system clears the arrays
    }
}

}  // namespace gato
