// gato_kernel_v5_gamma.cuh -- kernel 5 when every system of a call shares ONE matrix pair (ldG == ldC == 0: BASELINE.json
// config 4 read as "randomised g / c", the sampling / MPC-rollout workload).
//
// theta, phi, M and psi then belong to the call, not to a system: the first half runs once, and all that differs
// between systems is the right-hand side.  This kernel forms it for every system:
//     gamma_b = Q_b^-1 q_b - c_b + phi_b q_{b-1} + (B R^-1)_{b-1} r_{b-1}     (src/gato_schur.cuh:306-339; block row 0:
//     gamma^  = M gamma  (FP64 accumulate)                                      Q_0^-1 q_0, c[0:S] ignored, :110-147)
// from the caller's g and c and the one scratch the first half left (regions, phi, M in FP64), in the same order of
// operations as the Schur phase of the one-CTA kernel, four lanes per knot, one CTA per system.
#pragma once
#include "gato_kernel_v2.cuh"

namespace gato {

template <int S, int C>
__global__ void __launch_bounds__(256, 2) gato_gamma_scaled(const SolveParams2 q) {
    using L = Layout2<S, C>;
    using SL = ScaledLayout<S, C>;
    using MS = MapState<S, C>;
    constexpr int LPK = L::LPK, OWN = L::OWN, VS = L::VS, LDC = L::LDC;
    const SolveParams& p = q.p;
    const int K = p.K;
    const int tid = (int)threadIdx.x, lane = tid & 31;
    const int lik = lane % LPK, b = tid / LPK;
    const bool active = b < K;
    const int base = lane & ~(LPK - 1);
    int orow[OWN];
    bool ovr[OWN];
#pragma unroll
    for (int j = 0; j < OWN; ++j) {
        ovr[j] = L::own_valid(lik, j);
        orow[j] = ovr[j] ? L::own_row(lik, j) : 0;
    }
    const int po0 = L::own_valid(lik, 0) ? L::pos(L::own_row(lik, 0)) : 0;
    auto load_row = [&](const float* rowbase, float (&v)[S]) {
#pragma unroll
        for (int ci = 0; ci < VS / 4; ++ci) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(rowbase) + ci);
            const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (L::pos_valid(4 * ci + u)) v[L::pos_row(4 * ci + u)] = e[u];
        }
    };
    // row `row` of Q^-1 times q; with a diagonal cost pattern the scratch holds only the diagonal of Q^-1 (same bits)
    auto qdot = [&](const float* rowbase, int row, int dpos, const float (&qv)[S]) {
        float acc = 0.f;
        if (q.g_diag) {
            float qd = 0.f;
#pragma unroll
            for (int k = 0; k < S; ++k) qd = (k == row) ? qv[k] : qd;
            acc = fmaf(__ldg(rowbase + dpos), qd, 0.f);
        } else {
            float qrow[S];
            load_row(rowbase, qrow);
#pragma unroll
            for (int k = 0; k < S; ++k) acc = fmaf(qrow[k], qv[k], acc);
        }
        return acc;
    };
    const float* kg = q.scr + (long long)(active ? b : 0) * SL::KSCR;            // the ONE scratch (system 0 of the call)
    const float* kp = q.scr + (long long)(active && b > 0 ? b - 1 : 0) * SL::KSCR;
    const float* xk = q.shx + (long long)(active ? b : 0) * SL::KSHX;
    // (B R^-1) rows of this lane, as the Schur phase forms them: the same for every system of the call
    float br[OWN][C];
#pragma unroll
    for (int i = 0; i < OWN; ++i)
#pragma unroll
        for (int c = 0; c < C; ++c) br[i][c] = 0.f;
    if (active && b > 0) {
        if (q.g_diag) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const float ri = __ldg(kp + L::oR + c * LDC + c);
                const float4 t = __ldg(reinterpret_cast<const float4*>(kp + L::oB + c * VS + po0));
                const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                for (int i = 0; i < OWN; ++i) br[i][c] = e[i] * ri;
            }
        } else {
#pragma unroll 1
            for (int k = 0; k < C; ++k) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(kp + L::oB + k * VS + po0));
                const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const float y = __ldg(kp + L::oR + k * LDC + c);
#pragma unroll
                    for (int i = 0; i < OWN; ++i) br[i][c] = fmaf(e[i], y, br[i][c]);
                }
            }
        }
    }
    for (int sys = (int)blockIdx.x; sys < p.nsys; sys += (int)gridDim.x) {
        const float* gv = p.g + (long long)sys * p.ldg;
        const float* cv = p.c + (long long)sys * p.ldc;
        float gam[OWN];
#pragma unroll
        for (int i = 0; i < OWN; ++i) gam[i] = 0.f;
        if (active) {
            float qv[S];
#pragma unroll
            for (int k = 0; k < S; ++k) qv[k] = __ldg(gv + b * (S + C) + k);
            if (b == 0) {
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (ovr[i]) {
                        gam[i] = qdot(kg + L::oQ + orow[i] * VS, orow[i], po0 + i, qv);
                    }
            } else {
                float qp[S], rp[C];
#pragma unroll
                for (int k = 0; k < S; ++k) qp[k] = __ldg(gv + (b - 1) * (S + C) + k);
#pragma unroll
                for (int k = 0; k < C; ++k) rp[k] = __ldg(gv + (b - 1) * (S + C) + S + k);
                float a1[OWN];       // phi_b q_{b-1}, own rows: column k of phi_b holds them as one float4
#pragma unroll
                for (int i = 0; i < OWN; ++i) a1[i] = 0.f;
#pragma unroll
                for (int k = 0; k < S; ++k) {
                    const float4 t = __ldg(reinterpret_cast<const float4*>(xk + SL::xPHI + k * VS + po0));
                    a1[0] = fmaf(t.x, qp[k], a1[0]); a1[1] = fmaf(t.y, qp[k], a1[1]);
                    a1[2] = fmaf(t.z, qp[k], a1[2]); a1[OWN - 1] = fmaf(t.w, qp[k], a1[OWN - 1]);
                }
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (ovr[i]) {
                        const float a0 = qdot(kg + L::oQ + orow[i] * VS, orow[i], po0 + i, qv);
                        float a2 = 0.f;
#pragma unroll
                        for (int k = 0; k < C; ++k) a2 = fmaf(br[i][k], rp[k], a2);
                        gam[i] = (a0 - __ldg(cv + b * S + orow[i])) + (a2 + a1[i]);
                    }
            }
        }
        // gamma^ = M gamma: every lane of the knot needs the whole gamma_b (shuffles), then its own rows of M in FP64
        float gfull[S];
#pragma unroll
        for (int r = 0; r < S; ++r) gfull[r] = __shfl_sync(0xffffffffu, gam[MS::reg(r)], base + MS::lik(r));
        float gh[OWN];
#pragma unroll
        for (int i = 0; i < OWN; ++i) {
            gh[i] = 0.f;
            if (active && ovr[i]) {
                const double2* mrow = reinterpret_cast<const double2*>(reinterpret_cast<const double*>(xk + SL::xM64) + orow[i] * S);
                double acc = 0.0;
                static_assert(S % 2 == 0 && (SL::xM64 * 4) % 16 == 0 && (SL::KSHX * 4) % 16 == 0, "rows of M64 as 16-byte loads");
#pragma unroll
                for (int j = 0; j < S; j += 2) {   // zero above the diagonal
                    const double2 m = __ldg(mrow + j / 2);
                    acc = fma(m.x, (double)gfull[j], acc);
                    acc = fma(m.y, (double)gfull[j + 1], acc);
                }
                gh[i] = (float)acc;
            }
        }
        if (active)
            *reinterpret_cast<float4*>(q.gam + (long long)sys * q.gam_ld + b * VS + po0) = make_float4(gh[0], gh[1], gh[2], gh[OWN - 1]);
    }
}

}  // namespace gato
