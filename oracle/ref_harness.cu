// ref_harness.cu -- TEST INFRASTRUCTURE ONLY (checker + baseline; never on the product path).
//
// Include-harness around the UNMODIFIED reference headers: this file
// #includes /root/reference/src/gato_utils.cuh, gato_schur.cuh and gato_pcg.cuh as they lie and drives
// the reference's own kernels and host functions on PRE-ALLOCATED buffers, which the reference's only
// entry point (gpu_library.cu:25-83: 7 cudaMalloc + 2 cudaMemset + 7 cudaFree per solve) never does.
// Two jobs:
//   (a) dump the intermediates of the reference itself -- d_G_dense / d_C_dense right after
//       gato_convert_kkt_format (src/gato_schur.cuh:902), d_S / d_Pinv / d_gamma after form_schur
//       (:885-1009), G^-1 (d_G_dense after gato_form_schur_jacobi), lambda, dz -- for bit-exact / 2e-5
//       parity tests of the new kernels against the reference rather than against the NumPy restatement;
//   (b) time the reference's five kernels (gato_convert_kkt_format, gato_form_schur_jacobi, gato_form_ss,
//       parallelPCG, gato_compute_dz) with CUDA events around launches on pre-allocated buffers, i.e.
//       WITHOUT the per-solve cudaMalloc/cudaFree, the two cudaDeviceSynchronize and the printf of
//       gpu_library.cu:36-81 / src/gato_schur.cuh:899-981: BASELINE.md section 2.1 rows R3 (kernel-only us)
//       and R4 (us per PCG iteration).
// mode 0 = the reference's host API (form_schur / solve_pcg<float> / compute_dz) on the harness's buffers;
// mode 1 = the same five __global__ functions launched directly with the launch shapes those host
//          functions use (src/gato_schur.cuh:893-897,902,947,971,1017; src/gato_pcg.cuh:495-527).
// Built by oracle/build_ref.py into oracle/_ref/S{S}_C{C}_K{K}/libref_harness.so (git-ignored).
// No reference source is copied: compile-time sizes come from -DSTATE_SIZE -DCONTROL_SIZE -DKNOT_POINTS
// exactly as in the reference's CMakeLists.txt:18.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#include "include/gato_defines.h"
#include "include/types.h"
#include "src/gato_utils.cuh"
#include "src/gato_schur.cuh"
#include "src/gato_pcg.cuh"

namespace {

constexpr int S_ = STATE_SIZE, C_ = CONTROL_SIZE, K_ = KNOT_POINTS;
constexpr size_t kGDense = KKT_G_DENSE_SIZE_BYTES / sizeof(float);
constexpr size_t kCDense = KKT_C_DENSE_SIZE_BYTES / sizeof(float);
constexpr size_t kS3 = (size_t)3 * STATES_SQ * KNOT_POINTS;
constexpr size_t kNl = (size_t)STATE_SIZE * KNOT_POINTS;
constexpr size_t kNz = (size_t)STATES_S_CONTROLS * KNOT_POINTS - CONTROL_SIZE;

// Zero padding (floats) behind every buffer: the reference's dz kernel reads one block past the end of
// d_C_dense and S floats past the end of d_lambda for the last state rows (SURVEY 9.6,
// src/gato_schur.cuh:827-838); with zeros there it produces the intended value.  refh_set_padding(0)
// restores the reference's exact allocation sizes (for the compute-sanitizer run that documents 9.6).
size_t g_pad = 2 * (STATES_SQ + STATES_P_CONTROLS);

struct Buffers {
    int *G_row = nullptr, *G_col = nullptr, *C_row = nullptr, *C_col = nullptr;
    float *G_val = nullptr, *C_val = nullptr, *g = nullptr, *c = nullptr;
    float *G_dense = nullptr, *C_dense = nullptr, *S = nullptr, *Pinv = nullptr, *gamma = nullptr;
    float *lambda = nullptr, *dz = nullptr;
    // solve_pcg's scratch (src/gato_pcg.cuh:486-492), allocated once here
    float *r = nullptr, *p = nullptr, *v = nullptr, *eta_new = nullptr;
    int* iters = nullptr;
    int nnzG = 0, nnzC = 0;
    bool ok = false;
} B;

const char* g_err = "";
#define RH(expr)                                                     \
    do {                                                             \
        cudaError_t e_ = (expr);                                     \
        if (e_ != cudaSuccess) { g_err = cudaGetErrorString(e_); fprintf(stderr, "ref_harness: %s failed: %s\n", #expr, g_err); return (int)e_; } \
    } while (0)

template <class T>
int alloc0(T** p, size_t n) {
    RH(cudaMalloc((void**)p, (n + g_pad) * sizeof(T)));
    RH(cudaMemset(*p, 0, (n + g_pad) * sizeof(T)));
    return 0;
}

void release() {
    void* ps[] = {B.G_row, B.G_col, B.C_row, B.C_col, B.G_val, B.C_val, B.g, B.c, B.G_dense, B.C_dense, B.S,
                  B.Pinv, B.gamma, B.lambda, B.dz, B.r, B.p, B.v, B.eta_new, B.iters};
    for (void* q : ps)
        if (q) cudaFree(q);
    B = Buffers();
}

int upload(const int* G_row, const int* G_col, const float* G_val, int nnzG, const int* C_row, const int* C_col,
           const float* C_val, int nnzC, const float* g, const float* c) {
    release();
    int rc;
    B.nnzG = nnzG; B.nnzC = nnzC;
    if ((rc = alloc0(&B.G_row, kNz + 1)) || (rc = alloc0(&B.G_col, (size_t)nnzG)) || (rc = alloc0(&B.G_val, (size_t)nnzG)) ||
        (rc = alloc0(&B.C_row, kNl + 1)) || (rc = alloc0(&B.C_col, (size_t)nnzC)) || (rc = alloc0(&B.C_val, (size_t)nnzC)) ||
        (rc = alloc0(&B.g, kNz)) || (rc = alloc0(&B.c, kNl)))
        return rc;
    // same order as gato_linsys (gpu_library.cu:36-45)
    if ((rc = alloc0(&B.G_dense, kGDense)) || (rc = alloc0(&B.C_dense, kCDense)) || (rc = alloc0(&B.S, kS3)) ||
        (rc = alloc0(&B.Pinv, kS3)) || (rc = alloc0(&B.gamma, kNl)) || (rc = alloc0(&B.dz, kNz)) ||
        (rc = alloc0(&B.lambda, kNl)))
        return rc;
    if ((rc = alloc0(&B.r, kNl)) || (rc = alloc0(&B.p, kNl)) || (rc = alloc0(&B.v, 1)) || (rc = alloc0(&B.eta_new, 1)) ||
        (rc = alloc0(&B.iters, 1)))
        return rc;
    RH(cudaMemcpy(B.G_row, G_row, (kNz + 1) * sizeof(int), cudaMemcpyHostToDevice));
    RH(cudaMemcpy(B.G_col, G_col, (size_t)nnzG * sizeof(int), cudaMemcpyHostToDevice));
    RH(cudaMemcpy(B.G_val, G_val, (size_t)nnzG * sizeof(float), cudaMemcpyHostToDevice));
    RH(cudaMemcpy(B.C_row, C_row, (kNl + 1) * sizeof(int), cudaMemcpyHostToDevice));
    RH(cudaMemcpy(B.C_col, C_col, (size_t)nnzC * sizeof(int), cudaMemcpyHostToDevice));
    RH(cudaMemcpy(B.C_val, C_val, (size_t)nnzC * sizeof(float), cudaMemcpyHostToDevice));
    RH(cudaMemcpy(B.g, g, kNz * sizeof(float), cudaMemcpyHostToDevice));
    RH(cudaMemcpy(B.c, c, kNl * sizeof(float), cudaMemcpyHostToDevice));
    B.ok = true;
    return 0;
}

// what cuda_calloc does for the two dense buffers and d_lambda at the top of every gato_linsys call
int zero_dense() {
    RH(cudaMemsetAsync(B.G_dense, 0, KKT_G_DENSE_SIZE_BYTES, 0));
    RH(cudaMemsetAsync(B.C_dense, 0, KKT_C_DENSE_SIZE_BYTES, 0));
    RH(cudaMemsetAsync(B.lambda, 0, kNl * sizeof(float), 0));
    return 0;
}

// launch shapes of form_schur / compute_dz (src/gato_schur.cuh:893-897, 1014-1015)
dim3 small_block() { return dim3(((std::max(STATES_SQ, CONTROLS_SQ) / 32) + 1) * 32); }
int g_schur_block = 0;
int schur_block() {
    if (g_schur_block == 0) {
        int gridsize = 0, blocksize = 0;
        cudaOccupancyMaxPotentialBlockSize<void (*)(float*, float*, float*, float*, float*, float*, float*)>(
            &gridsize, &blocksize, gato_form_schur_jacobi, 0, 0);
        g_schur_block = blocksize;
    }
    return g_schur_block;
}

int k_convert(float rho) {
    gato_convert_kkt_format<<<dim3(K_), small_block()>>>(B.G_row, B.G_col, B.G_val, B.C_row, B.C_col, B.C_val,
                                                        B.G_dense, B.C_dense, rho);
    RH(cudaPeekAtLastError());
    return 0;
}
int k_schur() {
    gato_form_schur_jacobi<<<K_, schur_block()>>>(B.G_dense, B.C_dense, B.g, B.c, B.S, B.Pinv, B.gamma);
    RH(cudaPeekAtLastError());
    return 0;
}
int k_ss() {
#if SS_PRECON
    gato_form_ss<<<dim3(K_), small_block()>>>(B.S, B.Pinv, B.gamma);
    RH(cudaPeekAtLastError());
#endif
    return 0;
}
// the cooperative launch of solve_pcg (src/gato_pcg.cuh:495-527) on pre-allocated scratch
int g_pcg_blocks = -1;
int k_pcg(float eps, int max_iter) {
    void* kern = (void*)parallelPCG<float, false>;
    dim3 grid(K_, 1, 1), block(S_, 1, 1);
    if (g_pcg_blocks < 0) g_pcg_blocks = check_sms<float>(kern, block);
    if (g_pcg_blocks != K_) { g_err = "KNOT_POINTS exceeds the co-resident blocks: the reference would take its _fixed path"; return -2; }
    unsigned sharedMemSize = (2 * 3 * STATES_SQ + 12 * STATE_SIZE) * sizeof(float);
    void* args[] = {(void*)&B.S, (void*)&B.Pinv, (void*)&B.gamma, (void*)&B.lambda, (void*)&B.r, (void*)&B.p,
                    (void*)&B.v, (void*)&B.eta_new, (void*)&B.iters, (void*)&max_iter, (void*)&eps};
    RH(cudaLaunchCooperativeKernel(kern, grid, block, args, sharedMemSize));
    return 0;
}
int k_dz() {
    gato_compute_dz<<<dim3(K_), small_block()>>>(B.G_dense, B.C_dense, B.g, B.lambda, B.dz);
    RH(cudaPeekAtLastError());
    return 0;
}

}  // namespace

extern "C" {

const char* refh_last_error(void) { return g_err; }
void refh_sizes(int* S, int* C, int* K) { *S = S_; *C = C_; *K = K_; }
void refh_set_padding(int floats) { g_pad = floats > 0 ? (size_t)floats : 0; }
int refh_schur_block(void) { return schur_block(); }
void refh_release(void) { release(); }

int refh_upload(const int* G_row, const int* G_col, const float* G_val, int nnzG, const int* C_row,
                const int* C_col, const float* C_val, int nnzC, const float* g, const float* c) {
    return upload(G_row, G_col, G_val, nnzG, C_row, C_col, C_val, nnzC, g, c);
}

// One solve with dumps of every intermediate (any output pointer may be NULL).
int refh_run(int mode, float rho, float eps, int max_iter, float* G_dense_pre, float* C_dense, float* S,
             float* Pinv, float* gamma, float* G_inv, float* lambda, float* dz, int* iters) {
    if (!B.ok) { g_err = "refh_upload first"; return -1; }
    int rc;
    if ((rc = zero_dense())) return rc;
    // the unpack dump always comes from the reference's conversion kernel launched on its own
    if ((rc = k_convert(rho))) return rc;
    RH(cudaDeviceSynchronize());
    if (G_dense_pre) RH(cudaMemcpy(G_dense_pre, B.G_dense, KKT_G_DENSE_SIZE_BYTES, cudaMemcpyDeviceToHost));
    if (C_dense) RH(cudaMemcpy(C_dense, B.C_dense, KKT_C_DENSE_SIZE_BYTES, cudaMemcpyDeviceToHost));
    int it = -1;
    if (mode == 0) {
        if ((rc = zero_dense())) return rc;
        form_schur(B.G_row, B.G_col, B.G_val, B.G_dense, B.C_row, B.C_col, B.C_val, B.C_dense, B.g, B.c, B.S, B.Pinv,
                   B.gamma, rho);
        RH(cudaDeviceSynchronize());
    } else {
        if ((rc = k_schur()) || (rc = k_ss())) return rc;
        RH(cudaDeviceSynchronize());
    }
    if (S) RH(cudaMemcpy(S, B.S, kS3 * sizeof(float), cudaMemcpyDeviceToHost));
    if (Pinv) RH(cudaMemcpy(Pinv, B.Pinv, kS3 * sizeof(float), cudaMemcpyDeviceToHost));
    if (gamma) RH(cudaMemcpy(gamma, B.gamma, kNl * sizeof(float), cudaMemcpyDeviceToHost));
    if (G_inv) RH(cudaMemcpy(G_inv, B.G_dense, KKT_G_DENSE_SIZE_BYTES, cudaMemcpyDeviceToHost));
    if (mode == 0) {
        it = solve_pcg<float>(B.S, B.Pinv, B.gamma, B.lambda, false, eps, max_iter);
        compute_dz(B.G_dense, B.C_dense, B.g, B.lambda, B.dz);
    } else {
        if ((rc = k_pcg(eps, max_iter)) || (rc = k_dz())) return rc;
        RH(cudaMemcpy(&it, B.iters, sizeof(int), cudaMemcpyDeviceToHost));
    }
    RH(cudaDeviceSynchronize());
    if (lambda) RH(cudaMemcpy(lambda, B.lambda, kNl * sizeof(float), cudaMemcpyDeviceToHost));
    if (dz) RH(cudaMemcpy(dz, B.dz, kNz * sizeof(float), cudaMemcpyDeviceToHost));
    if (iters) *iters = it;
    return 0;
}

// Timing.  total_us[reps]: one event pair around the whole solve (mode 1: the five launches back to back, nothing
// else; mode 0: form_schur + solve_pcg + compute_dz, i.e. with the reference's own syncs, printf, cudaMallocAsync
// and blocking D2H of the iteration count).  phase_us[5][reps] (mode 1 only, second pass): events between the
// kernels -- convert, schur_jacobi, form_ss, parallelPCG, compute_dz.  The zeroing of the dense buffers (the
// reference's cuda_calloc) is outside every bracket.
int refh_time(int mode, float rho, float eps, int max_iter, int reps, int warmup, float* total_us,
              float* phase_us, int* iters_out) {
    if (!B.ok) { g_err = "refh_upload first"; return -1; }
    cudaEvent_t ev[6];
    for (auto& e : ev) RH(cudaEventCreate(&e));
    int rc, it = -1;
    for (int i = -warmup; i < reps; ++i) {
        if ((rc = zero_dense())) return rc;
        RH(cudaEventRecord(ev[0], 0));
        if (mode == 0) {
            form_schur(B.G_row, B.G_col, B.G_val, B.G_dense, B.C_row, B.C_col, B.C_val, B.C_dense, B.g, B.c, B.S, B.Pinv,
                       B.gamma, rho);
            it = solve_pcg<float>(B.S, B.Pinv, B.gamma, B.lambda, false, eps, max_iter);
            compute_dz(B.G_dense, B.C_dense, B.g, B.lambda, B.dz);
        } else {
            if ((rc = k_convert(rho)) || (rc = k_schur()) || (rc = k_ss()) || (rc = k_pcg(eps, max_iter)) || (rc = k_dz()))
                return rc;
        }
        RH(cudaEventRecord(ev[1], 0));
        RH(cudaEventSynchronize(ev[1]));
        float ms = 0.f;
        RH(cudaEventElapsedTime(&ms, ev[0], ev[1]));
        if (i >= 0 && total_us) total_us[i] = ms * 1e3f;
    }
    if (mode == 1) RH(cudaMemcpy(&it, B.iters, sizeof(int), cudaMemcpyDeviceToHost));
    if (iters_out) *iters_out = it;
    if (mode == 1 && phase_us) {
        for (int i = -warmup; i < reps; ++i) {
            if ((rc = zero_dense())) return rc;
            RH(cudaEventRecord(ev[0], 0));
            if ((rc = k_convert(rho))) return rc;
            RH(cudaEventRecord(ev[1], 0));
            if ((rc = k_schur())) return rc;
            RH(cudaEventRecord(ev[2], 0));
            if ((rc = k_ss())) return rc;
            RH(cudaEventRecord(ev[3], 0));
            if ((rc = k_pcg(eps, max_iter))) return rc;
            RH(cudaEventRecord(ev[4], 0));
            if ((rc = k_dz())) return rc;
            RH(cudaEventRecord(ev[5], 0));
            RH(cudaEventSynchronize(ev[5]));
            if (i >= 0)
                for (int ph = 0; ph < 5; ++ph) {
                    float ms = 0.f;
                    RH(cudaEventElapsedTime(&ms, ev[ph], ev[ph + 1]));
                    phase_us[(size_t)ph * reps + i] = ms * 1e3f;
                }
        }
    }
    for (auto& e : ev) cudaEventDestroy(e);
    return 0;
}

}  // extern "C"
