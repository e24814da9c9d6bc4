// tf32_experiment.cu -- the tensor-core option for the Schur GEMMs, MEASURED (BASELINE.json north_star: "tensor cores only
// for the batched per-knot block GEMMs in Schur formation and only if they hold the tolerance").
//
// theta'_b = phi_b A^T + BR B^T  (src/gato_schur.cuh:342-385: S x S x S and S x C x S products per knot, S = 14, C = 7)
// for a batch of knots, three ways on the same inputs:
//   ffma     FP32 FFMA, four lanes per knot, rows in registers (the mapping of the product kernels)
//   tf32     mma.sync.m16n8k8 TF32 (10-bit mantissa inputs, FP32 accumulate), one warp per knot, tiles padded to 16 x 16 x 16
//   tf32x3   the same with the 3 x TF32 split (a = a_hi + a_lo: a_hi b_hi + a_lo b_hi + a_hi b_lo), the usual way to
//            recover FP32-like accuracy from TF32 tensor cores
// Reports ns per knot at full occupancy and the error of theta' against an FP64 evaluation (max over knots of the
// max-norm relative error), as one JSON line.  tcgen05.mma (M >= 64 tiles) is not applicable to 14 x 14 blocks without
// stacking five knots per tile and block-diagonal masking; mma.sync is the tensor-core path that fits.
//
//     nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tf32_experiment.cu -o build/tf32_experiment
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_runtime.h>

constexpr int S = 14, C = 7;
constexpr int SS = S * S, SC = S * C;
// per knot, row-major: phi [S][S] | A [S][S] | BR [S][C] | B [S][C]
constexpr int KIN = 2 * SS + 2 * SC;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) k_ffma(const float* __restrict__ in, float* __restrict__ out, int nk) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int k = t / 4, l = t % 4;
    if (k >= nk) return;
    const float* phi = in + (size_t)k * KIN; const float* A = phi + SS; const float* BR = A + SS; const float* B = BR + SC;
    const int r0 = (l / 2) * 7 + (l % 2) * 4, nr = (l % 2) ? 3 : 4;     // the own-row split of Layout2
    float acc[4][S];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < S; ++c) acc[i][c] = 0.f;
#pragma unroll 2
    for (int kk = 0; kk < S; ++kk) {
        float y[S];
#pragma unroll
        for (int c = 0; c < S; ++c) y[c] = __ldg(A + c * S + kk);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float x = i < nr ? __ldg(phi + (r0 + i) * S + kk) : 0.f;
#pragma unroll
            for (int c = 0; c < S; ++c) acc[i][c] = fmaf(x, y[c], acc[i][c]);
        }
    }
#pragma unroll
    for (int kk = 0; kk < C; ++kk) {
        float y[S];
#pragma unroll
        for (int c = 0; c < S; ++c) y[c] = __ldg(B + c * C + kk);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float x = i < nr ? __ldg(BR + (r0 + i) * C + kk) : 0.f;
#pragma unroll
            for (int c = 0; c < S; ++c) acc[i][c] = fmaf(x, y[c], acc[i][c]);
        }
    }
    for (int i = 0; i < nr; ++i)
#pragma unroll
        for (int c = 0; c < S; ++c) out[(size_t)k * SS + (r0 + i) * S + c] = acc[i][c];
}

__device__ __forceinline__ unsigned to_tf32(float x) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// one warp per knot; D (16 x 16, two n-tiles) += X (16 x K8) Y^T, X row-major [S][ldx], Y row-major [S][ldy] (so that
// B-fragment element (k, n) = Y[n][k]); SPLIT = 3 x TF32
template <bool SPLIT>
__device__ __forceinline__ void tile_product(float (&d)[2][4], const float* X, int ldx, const float* Y, int ldy, int kdim, int lane) {
    const int ar = lane / 4, ac = lane % 4;
    for (int k0 = 0; k0 < kdim; k0 += 8) {
        float af[4];
        const int rows[4] = {ar, ar + 8, ar, ar + 8}, cols[4] = {k0 + ac, k0 + ac, k0 + ac + 4, k0 + ac + 4};
#pragma unroll
        for (int e = 0; e < 4; ++e) af[e] = (rows[e] < S && cols[e] < kdim) ? __ldg(X + rows[e] * ldx + cols[e]) : 0.f;
        unsigned ah[4], al[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            ah[e] = to_tf32(af[e]);
            al[e] = SPLIT ? to_tf32(af[e] - __uint_as_float(ah[e])) : 0u;
        }
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            const int n = nt * 8 + lane / 4;
            float bf[2];
            const int ks[2] = {k0 + lane % 4, k0 + lane % 4 + 4};
#pragma unroll
            for (int e = 0; e < 2; ++e) bf[e] = (n < S && ks[e] < kdim) ? __ldg(Y + n * ldy + ks[e]) : 0.f;
            unsigned bh[2], bl[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                bh[e] = to_tf32(bf[e]);
                bl[e] = SPLIT ? to_tf32(bf[e] - __uint_as_float(bh[e])) : 0u;
            }
            if (SPLIT) {
                mma_tf32(d[nt], al, bh);
                mma_tf32(d[nt], ah, bl);
            }
            mma_tf32(d[nt], ah, bh);
        }
    }
}

template <bool SPLIT>
__global__ void __launch_bounds__(256) k_mma(const float* __restrict__ in, float* __restrict__ out, int nk) {
    const int k = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x % 32;
    if (k >= nk) return;
    const float* phi = in + (size_t)k * KIN; const float* A = phi + SS; const float* BR = A + SS; const float* B = BR + SC;
    float d[2][4] = {};
    tile_product<SPLIT>(d, phi, S, A, S, S, lane);
    tile_product<SPLIT>(d, BR, C, B, C, C, lane);
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const int r = lane / 4, c = nt * 8 + 2 * (lane % 4);
        const int rr[4] = {r, r, r + 8, r + 8}, cc[4] = {c, c + 1, c, c + 1};
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (rr[e] < S && cc[e] < S) out[(size_t)k * SS + rr[e] * S + cc[e]] = d[nt][e];
    }
}

int main(int argc, char** argv) {
    const int nk = argc > 1 ? atoi(argv[1]) : 8192 * 50;
    std::vector<float> h((size_t)nk * KIN);
    srand(1);
    auto rnd = [] { return (float)rand() / RAND_MAX * 2.f - 1.f; };
    // magnitudes of the IIWA-shaped generator: phi = A Q^-1 with A = I + 0.01 N, Q^-1 in {0.1, 10}; BR = B R^-1 (R = 1e-4), B ~ 0.02
    for (int k = 0; k < nk; ++k) {
        float* phi = h.data() + (size_t)k * KIN; float* A = phi + SS; float* BR = A + SS; float* B = BR + SC;
        for (int r = 0; r < S; ++r)
            for (int c = 0; c < S; ++c) {
                A[r * S + c] = (r == c ? 1.f : 0.f) + (r >= 7 ? 0.01f * rnd() : (c == r + 7 ? 0.01f : 0.f));
                phi[r * S + c] = A[r * S + c] * (c < 7 ? 0.1f : 10.f);
            }
        for (int r = 0; r < S; ++r)
            for (int c = 0; c < C; ++c) {
                B[r * C + c] = r >= 7 ? 0.01f * (2.f * (r - 7 == c) + 0.3f * rnd()) : 0.f;
                BR[r * C + c] = B[r * C + c] * 1e4f;
            }
    }
    float *din, *dout;
    CK(cudaMalloc(&din, h.size() * 4));
    CK(cudaMalloc(&dout, (size_t)nk * SS * 4));
    CK(cudaMemcpy(din, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    // FP64 evaluation of a sample of knots
    const int ns = 2000;
    std::vector<double> ref((size_t)ns * SS);
    for (int s = 0; s < ns; ++s) {
        const int k = (int)((long long)s * nk / ns);
        const float* phi = h.data() + (size_t)k * KIN; const float* A = phi + SS; const float* BR = A + SS; const float* B = BR + SC;
        for (int r = 0; r < S; ++r)
            for (int c = 0; c < S; ++c) {
                double a = 0;
                for (int kk = 0; kk < S; ++kk) a += (double)phi[r * S + kk] * A[c * S + kk];
                for (int kk = 0; kk < C; ++kk) a += (double)BR[r * C + kk] * B[c * C + kk];
                ref[(size_t)s * SS + r * S + c] = a;
            }
    }
    std::vector<float> o((size_t)nk * SS);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char* names[3] = {"ffma", "tf32", "tf32x3"};
    printf("{\"knots\": %d", nk);
    for (int v = 0; v < 3; ++v) {
        auto launch = [&] {
            if (v == 0) k_ffma<<<(nk * 4 + 255) / 256, 256>>>(din, dout, nk);
            else if (v == 1) k_mma<false><<<(nk * 32 + 255) / 256, 256>>>(din, dout, nk);
            else k_mma<true><<<(nk * 32 + 255) / 256, 256>>>(din, dout, nk);
        };
        for (int i = 0; i < 3; ++i) launch();
        CK(cudaEventRecord(e0));
        const int reps = 20;
        for (int i = 0; i < reps; ++i) launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        CK(cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost));
        double worst = 0, mean = 0;
        for (int s = 0; s < ns; ++s) {
            const int k = (int)((long long)s * nk / ns);
            double mx = 0, md = 0;
            for (int i = 0; i < SS; ++i) {
                mx = fmax(mx, fabs(ref[(size_t)s * SS + i]));
                md = fmax(md, fabs((double)o[(size_t)k * SS + i] - ref[(size_t)s * SS + i]));
            }
            worst = fmax(worst, md / mx);
            mean += md / mx / ns;
        }
        printf(", \"%s\": {\"ns_per_knot\": %.3f, \"us_per_iiwa50_system\": %.3f, \"max_rel_err\": %.3e, \"mean_rel_err\": %.3e}",
               names[v], ms / reps * 1e6 / nk, ms / reps * 1e3 / nk * 50, worst, mean);
    }
    printf("}\n");
    return 0;
}
