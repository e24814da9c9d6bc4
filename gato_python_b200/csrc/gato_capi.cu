// gato_capi.cu -- host side of the C ABI declared in include/gato_b200.h.
//
// Replaces the reference's host driver (main_call / gato_linsys, gpu_library.cu:25-231): the
// sparsity pattern is analysed once into scatter maps (a "plan"), device buffers are persistent,
// and a solve is H2D -> ONE fused kernel -> D2H with no allocation and a single host sync.
// There is no CPU fallback: without a CUDA device every entry point fails with GATO_ENODEVICE.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gato_b200.h"
#include "gato_kernels.cuh"
#include "gato_kernel_v2.cuh"
#include "gato_kernel_v5.cuh"
#include "gato_kernel_v5_lat.cuh"
#include "gato_kernel_v5_gamma.cuh"

#ifndef GATO_EXTRA_S
#define GATO_EXTRA_S 0
#define GATO_EXTRA_C 0
#endif

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define GATO_CUDA(expr)                                                                   \
    do {                                                                                  \
        cudaError_t e_ = (expr);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver      \
                            ? GATO_ENODEVICE : (int)e_,                                   \
                        "%s: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t ensure(size_t n) {
        if (n <= bytes) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};

struct PinBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t ensure(size_t n) {
        if (n <= bytes) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        bytes = 0;
        cudaError_t e = cudaMallocHost(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
};

// every entry point that touches the device runs on the plan's device, whatever the caller's current one is
struct DeviceGuard {
    int prev = 0;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); cudaSetDevice(dev); }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

}  // namespace

struct gato_plan {
    int S, C, K, nnzG, nnzC, device, max_batch;
    int n_z, n_l, g_dense, c_dense;
    int cl_override = 0;
    int kernel = 0;         // 0 = automatic, 1 = cluster kernel (v1), 2 = one-CTA-per-system kernel (v2), 3 / 4 = v2 over a
                            // cluster / a cooperative grid, 5 = scaled system in two kernels
    int g_diag = 0;         // G's pattern is purely diagonal
    int sm_count = 0;
    int smem_optin = 0;
    int staging = 1;        // v2: TMA staging of the value arrays when it fits (gato_plan_set_staging)
    int precond = 0;        // GATO_PRECOND_STAIR / GATO_PRECOND_BLOCK_JACOBI (gato_plan_set_preconditioner)
    int zero_copy = 1;      // gato_solve: staged kernels read the pinned input buffer in place (gato_plan_set_zero_copy)
    int lam_zc = 0;         // where gato_solve left its last lambda: 0 = device buffer, 1 = pinned result buffer
    int lam_slot = 0;       // gato_solve ping-pongs its output between two slots of dlo ...
    bool have_lambda = false;   // ... so the previous lambda stays readable for GATO_WARM_RESIDENT
    // device: pattern maps
    int *d_gmap = nullptr, *d_cmap = nullptr, *d_gstart = nullptr, *d_cstart = nullptr;
    int *d_gdst2 = nullptr, *d_cdst2 = nullptr;   // v2: absolute shared-memory destinations
    int* d_cperm = nullptr;                       // v2, one CTA: (destination << kPermShift) | C_val index, in bank-sorted order
    int ncperm = 0;
    // device: staged inputs / outputs for the host-buffer entry points
    DevBuf dG, dC, dg, dc, dl, dlo, ddz, dit, dbg;
    DevBuf dgrid;           // kernel 4: global scratch of the cooperative-grid variant
    // kernel 5: per-system scratch between its two kernels (ScaledLayout), one chunk of systems -- one buffer per
    // stream the plan has been used on (launches on different streams may overlap; the chunk pipeline uses four)
    static constexpr int kScr = 8;
    DevBuf dscr[kScr];
    cudaStream_t scr_stream[kScr] = {};
    int nscr = 0;
    int v5_chunk = 0;       // kernel 5: systems per chunk (0 = automatic)
    int v5_shared = 0;      // kernel 5, calls with one matrix pair for every system (ldG == ldC == 0): 0 = first half once, 1 = per system
    int v5_variant = 0;     // kernel 5, second half: 0 = by call size, 1 = latency kernel (4 lanes per knot), 2 = throughput kernel
    int profile = 0;        // kernel 5: record CUDA events around each of its launches (gato_plan_set_profiling)
    std::vector<cudaEvent_t> prof_ev;   // 3 per chunk of the last profiled call: before first half | between | after second
    int prof_chunks = 0, prof_launches = 0;
    int batch_hint = 0;     // > 0 inside gato_solve_batched: the size of the whole call, so that every chunk of its
                            // pipeline runs the kernel the whole batch selects
    PinBuf hin, hout;   // pinned staging for the single-solve path
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // chunk pipeline of the host-buffer batched path: H2D of chunk i+1 overlaps the kernel of chunk i
    static constexpr int kPipe = 4;
    int pipe_chunk = 0;     // systems per chunk; 0 = automatic, <0 = pipeline off
    cudaStream_t pstream[kPipe] = {};
    cudaEvent_t pev[kPipe] = {};
    cudaEvent_t ev_shared = nullptr;
    // one-CTA kernel: resident CTAs per SM for the shared-memory size this plan launches with (computed once per
    // size; per plan, not per process: plans of different horizons / devices do not disturb each other)
    mutable int occ_smem = -1, occ_val = 0;
    mutable int occ5_smem = -1, occ5_val = 0;   // the same for the first half of kernel 5
    mutable int occp_val = 0;                   // second half of kernel 5: resident CTAs per SM
    mutable int occl_val = 0;                   // its latency variant
    mutable int occp_form = 0;
    int pcg_form = 0;       // kernel 5, throughput second half: 0 = the reference's two reductions per iteration, 1 = single reduction
};

// ------------------------------------------------------------------------------------------------
// kernel dispatch
// ------------------------------------------------------------------------------------------------
namespace {

struct LaunchCfg { int cl, kpc, maxw; size_t smem; };

template <int S, int C>
int pick_cfg(const gato_plan* pl, LaunchCfg* out) {
    using L = gato::Layout<S, C>;
    const int K = pl->K;
    auto fits = [&](int cl, int kpc) {
        return kpc <= 32 && (size_t)L::cta_floats(kpc, cl) * 4 <= (size_t)pl->smem_optin;
    };
    int cl = pl->cl_override;
    if (cl <= 0) {
        // default: the fewest CTAs whose warps (one per knot) still fit one CTA each -- fewer
        // CTAs means fewer DSMEM hops per cluster barrier
        for (cl = 1; cl <= 16; cl *= 2)
            if (fits(cl, (K + cl - 1) / cl)) break;
    }
    if (cl < 1 || cl > 16) return fail(GATO_EUNSUPPORTED, "K=%d does not fit one 16-CTA cluster", K);
    const int kpc = (K + cl - 1) / cl;
    if (!fits(cl, kpc))
        return fail(GATO_EUNSUPPORTED, "cluster of %d CTAs x %d knots does not fit (K=%d, smem %d B)",
                    cl, kpc, K, pl->smem_optin);
    out->cl = cl;
    out->kpc = kpc;
    out->maxw = kpc <= 16 ? 16 : 32;
    out->smem = (size_t)L::cta_floats(kpc, cl) * 4;
    return GATO_OK;
}

template <int S, int C, int MAXW>
int launch_t(const LaunchCfg& cfg, const gato::SolveParams& p, cudaStream_t st) {
    auto kern = gato::gato_fused_solve<S, C, MAXW>;
    GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    if (cfg.cl > 8)
        GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)(p.nsys * cfg.cl), 1, 1);
    lc.blockDim = dim3((unsigned)(cfg.kpc * 32), 1, 1);
    lc.dynamicSmemBytes = cfg.smem;
    lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)cfg.cl;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    lc.attrs = at;
    lc.numAttrs = 1;
    GATO_CUDA(cudaLaunchKernelEx(&lc, kern, p));
    return GATO_OK;
}

template <int S, int C>
int launch_sc(const gato_plan* pl, const gato::SolveParams& p, cudaStream_t st) {
    LaunchCfg cfg;
    int rc = pick_cfg<S, C>(pl, &cfg);
    if (rc) return rc;
    return cfg.maxw == 16 ? launch_t<S, C, 16>(cfg, p, st) : launch_t<S, C, 32>(cfg, p, st);
}

template <int S, int C>
int unpack_sc(const gato_plan* pl, const float* dG, const float* dC, float* Gd, float* Cd, float rho,
              cudaStream_t st) {
    const int warps_per_block = 4;
    const int blocks = (pl->K + warps_per_block - 1) / warps_per_block;
    gato::gato_unpack_kernel<S, C><<<blocks, warps_per_block * 32, 0, st>>>(
        pl->d_gmap, pl->d_cmap, pl->d_gstart, pl->d_cstart, dG, dC, Gd, Cd, pl->K, rho);
    GATO_CUDA(cudaGetLastError());
    return GATO_OK;
}


// ---- v2: one CTA per system ---------------------------------------------------------------------
// thread-count menu per state size: ptxas gets 65536 / NT registers per thread, so NT bounds the
// number of knots (8 lanes each) one CTA can hold
template <int S> struct V2Menu { static constexpr int n = 3; static constexpr int nt[3] = {64, 256, 1024}; };
template <> struct V2Menu<14> { static constexpr int n = 3; static constexpr int nt[3] = {64, 128, 224}; };

template <int S, int C>
int v2_pick_nt(const gato_plan* pl) {   // smallest menu entry that holds K knots and fits shared memory; 0 = none
    using L = gato::Layout2<S, C>;
    for (int i = 0; i < V2Menu<S>::n; ++i) {
        const int nt = V2Menu<S>::nt[i];
        if (nt >= L::LPK * pl->K && (size_t)(L::smem_floats(pl->K, nt) + 8) * 4 <= (size_t)pl->smem_optin) return nt;
    }
    return 0;
}

#ifndef GATO_V5_COPY_WARP
#define GATO_V5_COPY_WARP 0   // measured: the copy phase shrinks by 2 000 cycles per system and the other phases grow by as much (off)
#endif
template <int S, int C, int NT, bool SCALED = false>
int launch_v2_t(const gato_plan* pl, gato::SolveParams2 q, cudaStream_t st) {
    using L = gato::Layout2<S, C>;
    // first half of kernel 5 at the largest CTA shape: one warp more than the layout's NT threads, for the region copy
    // (256 threads x 255 registers fill the register file exactly; the smaller shapes run several CTAs per SM instead)
    constexpr bool XW = SCALED && NT == 224 && GATO_V5_COPY_WARP;
    constexpr int NTT = XW ? NT + 32 : NT;
    auto kern = gato::gato_fused_solve_v2<S, C, NT, false, false, SCALED, XW>;
    int& occ_smem = SCALED ? pl->occ5_smem : pl->occ_smem;
    int& occ_val = SCALED ? pl->occ5_val : pl->occ_val;
    const int core = L::smem_floats(q.p.K, NT);
    // TMA staging of the C / G / g values (+ 16 bytes for the mbarrier) when it fits beside the system
    const int stage = gato::stage_len(q.nnzC) + gato::stage_len(q.nnzG) + gato::stage_len(pl->n_z);
    const int tail = 8;   // mbarrier + loop state of the persistent CTA
    const bool staged = pl->staging && (size_t)(core + stage + tail) * 4 <= (size_t)pl->smem_optin;
    q.stage_off = staged ? core : 0;
    q.tail_off = core + (staged ? stage : 0);
    q.kpc = q.p.K;
    const size_t smem = (size_t)(q.tail_off + tail) * 4;
    // persistent grid: as many CTAs as are resident at once, each walking systems i, i + grid, ...
    // MaxDynamicSharedMemorySize is state of (function, device): it is raised ONCE per device to the opt-in maximum
    // (under a lock), so host threads driving plans of different shared-memory sizes never lower it under each other.
    {
        constexpr int kMaxDev = 64;
        static std::mutex mu;
        static bool raised[kMaxDev];
        std::lock_guard<std::mutex> lk(mu);
        const int di = pl->device % kMaxDev;
        if (!raised[di]) {
#ifdef GATO_S_TRACE
            GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl->smem_optin - 256));   // trace sums are static shared memory
#else
            GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl->smem_optin));
#endif
            raised[di] = true;
        }
    }
    if (occ_smem != (int)smem) {
        int occ = 0;
        GATO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NTT, smem));
        if (occ < 1) return fail(GATO_EUNSUPPORTED, "one-CTA kernel does not fit an SM (%zu B shared memory)", smem);
        occ_smem = (int)smem;
        occ_val = occ;
    }
    const int resident = occ_val * (pl->sm_count > 0 ? pl->sm_count : 148);
    const unsigned grid = (unsigned)(q.p.nsys < resident ? q.p.nsys : resident);
    kern<<<grid, NTT, smem, st>>>(q);
    GATO_CUDA(cudaGetLastError());
    return GATO_OK;
}

template <int S, int C>
int launch_v2_sc(const gato_plan* pl, const gato::SolveParams2& q, cudaStream_t st) {
    const int nt = v2_pick_nt<S, C>(pl);
    if (nt == V2Menu<S>::nt[0]) return launch_v2_t<S, C, V2Menu<S>::nt[0]>(pl, q, st);
    if (nt == V2Menu<S>::nt[1]) return launch_v2_t<S, C, V2Menu<S>::nt[1]>(pl, q, st);
    if constexpr (V2Menu<S>::n > 2) {
        if (nt == V2Menu<S>::nt[2]) return launch_v2_t<S, C, V2Menu<S>::nt[2]>(pl, q, st);
    }
    return fail(GATO_EUNSUPPORTED, "K=%d does not fit the one-CTA kernel", pl->K);
}

// ---- kernel 5: the scaled system in two kernels (gato_kernel_v5.cuh) -------------------------------------------
// first half = the one-CTA kernel's unpack + Schur phases + the FP64 congruence (1 CTA per SM, 255 registers), second
// half = PCG on S^ = I + O^ + dz (2 CTAs per SM); they meet in a per-plan scratch buffer that holds one chunk of
// systems (ScaledLayout::KSCR floats per knot) and is reused chunk after chunk, so it stays in the 126 MB L2.
template <int S> struct HasV5 { static constexpr bool value = (S == 14); };
constexpr int kV5Threads = 128;     // second half: two lanes per knot, up to 64 knots

template <int S, int C>
int v5_ok_sc(const gato_plan* pl) {
    if constexpr (HasV5<S>::value) {
        using L = gato::Layout2<S, C>;
        return (2 * pl->K <= kV5Threads && v2_pick_nt<S, C>(pl) > 0) ? 1 : 0;
    } else {
        return 0;
    }
}

template <int S, int C>
long long v5_scratch_per_system(const gato_plan* pl) {
    if constexpr (HasV5<S>::value) return (long long)pl->K * gato::ScaledLayout<S, C>::KSCR;
    else return 0;
}

gato::SolveParams slice_params(const gato::SolveParams& base, int r0, int nr) {
    gato::SolveParams p = base;
    p.nsys = nr;
    p.G_val = base.G_val + (long long)r0 * base.ldG;
    p.C_val = base.C_val + (long long)r0 * base.ldC;
    p.g = base.g + (long long)r0 * base.ldg;
    p.c = base.c + (long long)r0 * base.ldc;
    if (base.lambda_in) p.lambda_in = base.lambda_in + (long long)r0 * base.ldl;
    p.lambda_out = base.lambda_out + (long long)r0 * base.ldlo;
    p.dz_out = base.dz_out + (long long)r0 * base.lddz;
    p.iters_out = base.iters_out + r0;
    if (base.eta_out) p.eta_out = base.eta_out + r0;
    return p;
}

template <int S, int C>
int launch_v5_first(gato_plan* pl, const gato::SolveParams2& q, cudaStream_t st) {
    if constexpr (HasV5<S>::value) {
        const int nt = v2_pick_nt<S, C>(pl);
        if (nt == V2Menu<S>::nt[0]) return launch_v2_t<S, C, V2Menu<S>::nt[0], true>(pl, q, st);
        if (nt == V2Menu<S>::nt[1]) return launch_v2_t<S, C, V2Menu<S>::nt[1], true>(pl, q, st);
        if constexpr (V2Menu<S>::n > 2) {
            if (nt == V2Menu<S>::nt[2]) return launch_v2_t<S, C, V2Menu<S>::nt[2], true>(pl, q, st);
        }
    }
    return fail(GATO_EUNSUPPORTED, "kernel 5 does not hold S=%d K=%d", pl->S, pl->K);
}

// latency variant of the second half (four lanes per knot, seven warps per system): single solves and batches below
// one system per SM
template <int S, int C>
int launch_v5_second_lat(gato_plan* pl, const gato::SolveParams2& q, cudaStream_t st) {
    if constexpr (HasV5<S>::value) {
        using L = gato::Layout2<S, C>;
        constexpr int NT = 224, NKP = (NT / 32) * L::KPW;
        constexpr size_t smem = (size_t)(4 * (NKP + 2) * L::VS + NKP * (L::VS + L::LDC) + 64) * sizeof(float);
        auto kern = gato::gato_pcg_scaled4<S, C, NT>;
        if (pl->occl_val == 0) {
            int occ = 0;
            GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            GATO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, smem));
            if (occ < 1) return fail(GATO_EUNSUPPORTED, "second half of kernel 5 does not fit an SM");
            pl->occl_val = occ;
        }
        const int resident = pl->occl_val * (pl->sm_count > 0 ? pl->sm_count : 148);
        const unsigned grid = (unsigned)(q.p.nsys < resident ? q.p.nsys : resident);
        kern<<<grid, NT, smem, st>>>(q);
        GATO_CUDA(cudaGetLastError());
        return GATO_OK;
    }
    return fail(GATO_EUNSUPPORTED, "kernel 5 does not hold S=%d K=%d", pl->S, pl->K);
}

// throughput second half with NT threads per CTA (two lanes per knot: horizons up to NT / 2 knots).  The kernel needs
// 168 registers whatever NT is, so a shorter horizon gets more CTAs per SM instead of idle warps: 128 threads x 3 CTAs,
// 96 x 4, 64 x 6, 32 x 12 (measured at K = 48: 1.52 ms per 8192 systems with 96 x 4 against 1.86 ms with 128 x 3).
template <int S, int C, int NT>
int launch_v5_second_nt(gato_plan* pl, const gato::SolveParams2& q, cudaStream_t st) {
    using L = gato::Layout2<S, C>;
    constexpr int NKP = NT / 2;
    constexpr size_t smem = (size_t)(4 * (2 * (NKP + 2) * L::HALFP + 4) + NKP * (L::VS + L::LDC) + 64 + 4) * sizeof(float);
    auto kern = pl->pcg_form == 1 ? gato::gato_pcg_scaled<S, C, NT, true> : gato::gato_pcg_scaled<S, C, NT, false>;
    if (pl->occp_val == 0 || pl->occp_form != pl->pcg_form) {
        int occ = 0;
        GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        GATO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, smem));
        if (occ < 1) return fail(GATO_EUNSUPPORTED, "second half of kernel 5 does not fit an SM");
        pl->occp_val = occ;
        pl->occp_form = pl->pcg_form;
    }
    const int resident = pl->occp_val * (pl->sm_count > 0 ? pl->sm_count : 148);
    const unsigned grid = (unsigned)(q.p.nsys < resident ? q.p.nsys : resident);
    kern<<<grid, NT, smem, st>>>(q);
    GATO_CUDA(cudaGetLastError());
    return GATO_OK;
}

// Horizons a few knots beyond a multiple of 16 leave the last warp nearly empty (K = 50: 4 of 32 lanes), and that warp
// costs a full share of issue slots.  Built and measured against it: the knots beyond 48 as "tail" knots on 16 lanes
// each (psi rows and columns from shared memory, no shuffles) as a short second pass of one warp, 96 threads and four
// systems per SM -- 2.20 ms per 8192 systems against 1.87 ms: the tail's 16 128-bit shared-memory loads per product land
// on the LSU pipe that is already 58 % busy, and the warp that runs them lengthens every barrier interval.  Not kept.
// Second attempt, 128 threads: the fourth warp keeps its place but runs an instruction stream of its own (16 lanes per
// knot, row and column of psi in the registers the other warps use for their rows: 28 FFMA and no shuffles per product)
// -- correct, no spills, and 1.99 ms against 1.87 ms.  The nearly empty warp is not what the other warps wait for.
template <int S, int C>
int launch_v5_second(gato_plan* pl, const gato::SolveParams2& q, cudaStream_t st, bool lat) {
    if constexpr (HasV5<S>::value) {
        if (lat) return launch_v5_second_lat<S, C>(pl, q, st);
        if (2 * pl->K <= 32) return launch_v5_second_nt<S, C, 32>(pl, q, st);
        if (2 * pl->K <= 64) return launch_v5_second_nt<S, C, 64>(pl, q, st);
        if (2 * pl->K <= 96) return launch_v5_second_nt<S, C, 96>(pl, q, st);
        return launch_v5_second_nt<S, C, kV5Threads>(pl, q, st);
    }
    return fail(GATO_EUNSUPPORTED, "kernel 5 does not hold S=%d K=%d", pl->S, pl->K);
}

constexpr int kV5SharedChunk = 1 << 16;   // systems per gamma^ array of the shared-matrix path (3.2 KB each at IIWA-50)

// first_only: run just the first half on the first system and leave its scratch in pl->dscr (parity hook)
template <int S, int C>
int launch_v5_sc(gato_plan* pl, gato::SolveParams2 q, cudaStream_t st, bool first_only) {
    if (!v5_ok_sc<S, C>(pl)) return fail(GATO_EUNSUPPORTED, "kernel 5 does not hold S=%d K=%d", pl->S, pl->K);
    const long long per = v5_scratch_per_system<S, C>(pl);
    const int sms = pl->sm_count > 0 ? pl->sm_count : 148;
    // measured (profiles/r02_k5_chunk_sweep.json): larger chunks win although the scratch then leaves L2 -- the tail of
    // each second-half launch (iteration counts differ between systems) costs more than the DRAM round trip
    int chunk = pl->v5_chunk > 0 ? pl->v5_chunk : 64 * sms;
    if (chunk > q.p.nsys) chunk = q.p.nsys;
    int si = 0;
    while (si < pl->nscr && pl->scr_stream[si] != st) ++si;
    if (si == pl->nscr) {
        if (pl->nscr == gato_plan::kScr)
            return fail(GATO_EUNSUPPORTED, "kernel 5 keeps one scratch buffer per stream: more than %d streams on one plan", gato_plan::kScr);
        pl->scr_stream[pl->nscr++] = st;
    }
    GATO_CUDA(pl->dscr[si].ensure((size_t)chunk * (size_t)per * sizeof(float) + 4096));
    q.scr = (float*)pl->dscr[si].p;
    q.gbuf = q.scr + (size_t)chunk * (size_t)per;     // behind the scratch: the second half's work counter
    q.scr_ld = per;
    const gato::SolveParams base = q.p;
    // which second half: decided by the size of the whole call (every chunk of it runs the same kernel, so the result
    // does not depend on the chunking); gato_plan_set_scaled_variant overrides
    const int whole = pl->batch_hint > 0 ? pl->batch_hint : base.nsys;
    const bool lat = 4 * pl->K <= 224 && (pl->v5_variant == 1 || (pl->v5_variant == 0 && whole < sms));
    const bool prof = pl->profile && !first_only;
    // One matrix pair for the whole call (ldG == ldC == 0): theta, phi, M and psi are computed ONCE; a small kernel
    // forms gamma^ of every system and the second half reads the one scratch.  (Throughput variant only.)
    // (not with a true warm start: gamma^ of that path knows nothing of lambda0)
    const bool shared = !first_only && !lat && pl->v5_shared != 1 && base.nsys > 1 && base.ldG == 0 && base.ldC == 0 && base.warm != 2;
    const int nchunk = shared ? (base.nsys + kV5SharedChunk - 1) / kV5SharedChunk : (base.nsys + chunk - 1) / chunk;
    if (prof) {
        while ((int)pl->prof_ev.size() < 3 * nchunk) {
            cudaEvent_t e;
            GATO_CUDA(cudaEventCreate(&e));
            pl->prof_ev.push_back(e);
        }
        pl->prof_chunks = 0;
    }
    if (shared) {
        if constexpr (HasV5<S>::value) {
            using SL = gato::ScaledLayout<S, C>;
            using L = gato::Layout2<S, C>;
            const int cmax = base.nsys < kV5SharedChunk ? base.nsys : kV5SharedChunk;
            const size_t o_shx = (size_t)per * sizeof(float) + 4096;          // scratch of ONE system | counter | phi, M64 | gamma^
            const size_t o_gam = o_shx + (((size_t)pl->K * SL::KSHX * sizeof(float) + 255) & ~(size_t)255);
            const long long gam_ld = (long long)pl->K * L::VS;
            GATO_CUDA(pl->dscr[si].ensure(o_gam + (size_t)cmax * (size_t)gam_ld * sizeof(float)));
            char* b0 = (char*)pl->dscr[si].p;
            q.scr = (float*)b0;
            q.gbuf = q.scr + (size_t)per;
            q.scr_ld = 0;
            q.shx = (float*)(b0 + o_shx);
            q.p = slice_params(base, 0, 1);
            if (prof) GATO_CUDA(cudaEventRecord(pl->prof_ev[0], st));
            int rc = launch_v5_first<S, C>(pl, q, st);
            if (rc) return rc;
            q.gam = (float*)(b0 + o_gam);
            q.gam_ld = gam_ld;
            const int gnt = 32 * L::warps(pl->K);
            for (int r0 = 0, ci = 0; r0 < base.nsys; r0 += kV5SharedChunk, ++ci) {
                const int nr = base.nsys - r0 < kV5SharedChunk ? base.nsys - r0 : kV5SharedChunk;
                q.p = slice_params(base, r0, nr);
                if (prof && ci > 0) GATO_CUDA(cudaEventRecord(pl->prof_ev[3 * ci], st));
                if (ci > 0) GATO_CUDA(cudaMemsetAsync(q.gbuf, 0, sizeof(int), st));   // the work counter (the first half reset it for chunk 0)
                const unsigned ggrid = (unsigned)(nr < 8 * sms ? nr : 8 * sms);
                gato::gato_gamma_scaled<S, C><<<ggrid, gnt, 0, st>>>(q);
                GATO_CUDA(cudaGetLastError());
                if (prof) GATO_CUDA(cudaEventRecord(pl->prof_ev[3 * ci + 1], st));
                if ((rc = launch_v5_second<S, C>(pl, q, st, false))) return rc;
                if (prof) { GATO_CUDA(cudaEventRecord(pl->prof_ev[3 * ci + 2], st)); pl->prof_chunks = ci + 1; }
            }
            pl->prof_launches = 1 + 2 * nchunk;
            return GATO_OK;
        }
    }
    for (int r0 = 0, ci = 0; r0 < base.nsys; r0 += chunk, ++ci) {
        const int nr = base.nsys - r0 < chunk ? base.nsys - r0 : chunk;
        q.p = slice_params(base, r0, nr);
        if (prof) GATO_CUDA(cudaEventRecord(pl->prof_ev[3 * ci], st));
        int rc = launch_v5_first<S, C>(pl, q, st);
        if (rc || first_only) return rc;
        if (prof) GATO_CUDA(cudaEventRecord(pl->prof_ev[3 * ci + 1], st));
        if ((rc = launch_v5_second<S, C>(pl, q, st, lat))) return rc;
        if (prof) { GATO_CUDA(cudaEventRecord(pl->prof_ev[3 * ci + 2], st)); pl->prof_chunks = ci + 1; }
    }
    pl->prof_launches = 2 * nchunk;
    return GATO_OK;
}

// ---- v2 over a thread-block cluster (kernel 3): horizons beyond one SM's registers ---------------------
struct ClusterCfg { int ncta, kpc, nt; size_t smem; };

template <int S, int C>
int v2cl_cfg(const gato_plan* pl, ClusterCfg* out) {
    using L = gato::Layout2<S, C>;
    if (pl->K < 2) return fail(GATO_EUNSUPPORTED, "the cluster variant of the one-CTA kernel needs K >= 2");
    const int nt_max = V2Menu<S>::nt[V2Menu<S>::n - 1];
    const int maxk = (nt_max / 32) * L::KPW - 2;        // two of a CTA's knot slots are halo knots
    if (maxk < 1) return fail(GATO_EUNSUPPORTED, "no cluster variant for S=%d", S);
    int ncta;
    if (pl->kernel == 3 && pl->cl_override > 0) {
        ncta = pl->cl_override;                         // explicit cluster size (gato_plan_set_cluster)
    } else {
        // default: CTAs of about one warp per SM sub-partition (4 warps; two of the knot slots are halos) -- the
        // per-iteration compute of a CTA halves against the 7-warp shape while the exchange cost stays (measured)
        const int pref = 4 * L::KPW - 2;
        ncta = (pl->K + pref - 1) / pref;
        if (ncta > 16) ncta = (pl->K + maxk - 1) / maxk;
    }
    if (ncta < 2) ncta = 2;
    if (ncta > 16) return fail(GATO_EUNSUPPORTED, "K=%d needs %d CTAs: more than one 16-CTA cluster", pl->K, ncta);
    const int kpc = (pl->K + ncta - 1) / ncta;
    ncta = (pl->K + kpc - 1) / kpc;                      // no empty CTA at the end
    if (ncta < 2) return fail(GATO_EUNSUPPORTED, "K=%d does not split over %d CTAs", pl->K, ncta);
    int nt = 0;
    for (int i = 0; i < V2Menu<S>::n; ++i)
        if ((V2Menu<S>::nt[i] / 32) * L::KPW >= kpc + 2) { nt = V2Menu<S>::nt[i]; break; }
    if (nt == 0) return fail(GATO_EUNSUPPORTED, "%d knots per CTA do not fit the cluster variant (max %d)", kpc, maxk);
    const size_t smem = (size_t)(L::smem_floats_cl(kpc, nt) + 12) * 4;   // + control tail (3 mbarriers, loop state, parities)
    if (smem > (size_t)pl->smem_optin)
        return fail(GATO_EUNSUPPORTED, "cluster variant: %zu B shared memory per CTA > %d", smem, pl->smem_optin);
    out->ncta = ncta; out->kpc = kpc; out->nt = nt; out->smem = smem;
    return GATO_OK;
}

template <int S, int C, int NT>
int launch_v2cl_t(const ClusterCfg& cfg, gato::SolveParams2 q, cudaStream_t st) {
    using L = gato::Layout2<S, C>;
    auto kern = gato::gato_fused_solve_v2<S, C, NT, true>;
    q.stage_off = 0;
    q.tail_off = L::smem_floats_cl(cfg.kpc, NT);
    q.kpc = cfg.kpc;
    GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    if (cfg.ncta > 8) GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)(q.p.nsys * cfg.ncta), 1, 1);
    lc.blockDim = dim3((unsigned)NT, 1, 1);
    lc.dynamicSmemBytes = cfg.smem;
    lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)cfg.ncta;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    lc.attrs = at;
    lc.numAttrs = 1;
    GATO_CUDA(cudaLaunchKernelEx(&lc, kern, q));
    return GATO_OK;
}

template <int S, int C>
int launch_v2cl_sc(const gato_plan* pl, gato::SolveParams2 q, cudaStream_t st) {
    ClusterCfg cfg;
    int rc = v2cl_cfg<S, C>(pl, &cfg);
    if (rc) return rc;
    if (cfg.nt == V2Menu<S>::nt[0]) return launch_v2cl_t<S, C, V2Menu<S>::nt[0]>(cfg, q, st);
    if (cfg.nt == V2Menu<S>::nt[1]) return launch_v2cl_t<S, C, V2Menu<S>::nt[1]>(cfg, q, st);
    if constexpr (V2Menu<S>::n > 2) {
        if (cfg.nt == V2Menu<S>::nt[2]) return launch_v2cl_t<S, C, V2Menu<S>::nt[2]>(cfg, q, st);
    }
    return fail(GATO_EUNSUPPORTED, "no cluster variant with %d threads", cfg.nt);
}

// ---- v2 over a cooperative grid (kernel 4): horizons beyond one 16-CTA cluster --------------------------
// CTAs of V2Menu::nt[1] threads (four warps at S = 14), up to 30 own knots each, all co-resident; slices and
// partials through a global scratch buffer, grid.sync() at the synchronisation points.  One system per launch.
template <int S, int C>
int v2grid_cfg(const gato_plan* pl, ClusterCfg* out) {
    using L = gato::Layout2<S, C>;
    constexpr int NT = V2Menu<S>::nt[1];
    if (pl->K < 2) return fail(GATO_EUNSUPPORTED, "the grid variant needs K >= 2");
    const int maxk = (NT / 32) * L::KPW - 2;
    if (maxk < 1) return fail(GATO_EUNSUPPORTED, "no grid variant for S=%d", S);
    int ncta = (pl->K + maxk - 1) / maxk;
    if (ncta < 2) ncta = 2;
    const int kpc = (pl->K + ncta - 1) / ncta;
    ncta = (pl->K + kpc - 1) / kpc;
    if (ncta < 2) return fail(GATO_EUNSUPPORTED, "K=%d does not split over CTAs", pl->K);
    const size_t smem = (size_t)(L::smem_floats_cl(kpc, NT) + 12) * 4;
    if (smem > (size_t)pl->smem_optin)
        return fail(GATO_EUNSUPPORTED, "grid variant: %zu B shared memory per CTA > %d", smem, pl->smem_optin);
    auto kern = gato::gato_fused_solve_v2<S, C, NT, true, true>;
    int occ = 0;
    GATO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    GATO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, smem));
    const int cap = occ * (pl->sm_count > 0 ? pl->sm_count : 148);
    if (ncta > cap || ncta > 256)
        return fail(GATO_EUNSUPPORTED, "K=%d needs %d co-resident CTAs, the device holds %d", pl->K, ncta, cap < 256 ? cap : 256);
    out->ncta = ncta; out->kpc = kpc; out->nt = NT; out->smem = smem;
    return GATO_OK;
}

template <int S, int C>
int launch_v2grid_sc(gato_plan* pl, gato::SolveParams2 q, cudaStream_t st) {
    using L = gato::Layout2<S, C>;
    constexpr int NT = V2Menu<S>::nt[1];
    ClusterCfg cfg;
    int rc = v2grid_cfg<S, C>(pl, &cfg);
    if (rc) return rc;
    GATO_CUDA(pl->dgrid.ensure((size_t)(512 + cfg.ncta * 6 * L::VS) * sizeof(float)));
    auto kern = gato::gato_fused_solve_v2<S, C, NT, true, true>;
    q.stage_off = 0;
    q.tail_off = L::smem_floats_cl(cfg.kpc, NT);
    q.kpc = cfg.kpc;
    q.ncta = cfg.ncta;
    q.gbuf = (float*)pl->dgrid.p;
    const gato::SolveParams base = q.p;
    for (int sys = 0; sys < base.nsys; ++sys) {   // a cooperative grid holds one system: batches run back to back
        gato::SolveParams& p = q.p;
        p = base;
        p.nsys = 1;
        p.G_val = base.G_val + (long long)sys * base.ldG;
        p.C_val = base.C_val + (long long)sys * base.ldC;
        p.g = base.g + (long long)sys * base.ldg;
        p.c = base.c + (long long)sys * base.ldc;
        if (base.lambda_in) p.lambda_in = base.lambda_in + (long long)sys * base.ldl;
        p.lambda_out = base.lambda_out + (long long)sys * base.ldlo;
        p.dz_out = base.dz_out + (long long)sys * base.lddz;
        p.iters_out = base.iters_out + sys;
        if (base.eta_out) p.eta_out = base.eta_out + sys;
        void* args[] = {(void*)&q};
        GATO_CUDA(cudaLaunchCooperativeKernel((void*)kern, dim3((unsigned)cfg.ncta), dim3((unsigned)NT), args, cfg.smem, st));
    }
    return GATO_OK;
}

template <int S, int C>
void v2_fill_maps(const gato_plan* pl, const std::vector<int>& gmap, const std::vector<int>& gstart,
                  const std::vector<int>& cmap, const std::vector<int>& cstart, std::vector<int>* gdst,
                  std::vector<int>* cdst) {
    // gmap / cmap hold the reference's column-major in-block offsets; re-target them to
    // the v2 shared-memory layout (Layout2: pos-padded rows, Q and R row-major)
    using L = gato::Layout2<S, C>;
    for (int k = 0; k < pl->K; ++k) {
        for (int t = gstart[k]; t < gstart[k + 1]; ++t) {
            const int off = gmap[t] & gato::kMapMask;
            int dst;
            if (off < S * S) {
                const int c = off / S, r = off % S;
                dst = L::oQ + r * L::VS + L::pos(c);
            } else {
                const int c = (off - S * S) / C, r = (off - S * S) % C;
                dst = L::oR + r * L::LDC + c;
            }
            (*gdst)[t] = (k * L::KN + dst) | (gmap[t] & gato::kMapDiagBit);
        }
        for (int t = cstart[k]; t < cstart[k + 1]; ++t) {
            if (cmap[t] < 0) { (*cdst)[t] = -1; continue; }
            const int col = cmap[t] / S, r = cmap[t] % S;
            const int dst = col < S ? L::oA + col * L::VS + L::pos(r) : L::oB + (col - S) * L::VS + L::pos(r);
            (*cdst)[t] = k * L::KN + dst;
        }
    }
}

// The scatter of C's non-zeros into the knot regions, reordered so that every 32 consecutive entries (one warp store)
// fall into 32 distinct shared-memory banks -- and come from 32 distinct banks of the staged copy of C_val.  CSR order
// puts the entries of one row of [A | B] next to each other, and those sit VS = 16 words apart: 16 of 32 lanes on one
// bank (measured 21.1 M shared-memory wavefronts per launch where 2.2 M are needed).  Greedy edge colouring of the
// bipartite (source bank, destination bank) multigraph: each entry goes to the first warp store that has both banks
// free; the unused lanes of a store carry kPermPad.  -> one word per entry, (destination << kPermShift) | source.
std::vector<int> v2_bank_sorted_scatter(const std::vector<int>& cdst) {
    struct Grp { unsigned src = 0, dst = 0; int n = 0; int e[32]; };
    std::vector<Grp> grp;
    size_t first_open = 0;
    for (size_t t = 0; t < cdst.size(); ++t) {
        if (cdst[t] < 0) continue;
        const unsigned bs = 1u << (t & 31), bd = 1u << (cdst[t] & 31);
        size_t g = first_open;
        while (g < grp.size() && ((grp[g].src & bs) || (grp[g].dst & bd))) ++g;
        if (g == grp.size()) grp.emplace_back();
        grp[g].src |= bs; grp[g].dst |= bd;
        grp[g].e[grp[g].n++] = (int)t;
        while (first_open < grp.size() && grp[first_open].n == 32) ++first_open;
    }
    std::vector<int> out;
    out.reserve(grp.size() * 32);
    for (const Grp& g : grp)
        for (int i = 0; i < 32; ++i) out.push_back(i < g.n ? (cdst[g.e[i]] << gato::kPermShift) | g.e[i] : gato::kPermPad);
    return out;
}

// -DGATO_EXTRA_S/-DGATO_EXTRA_C add one more (STATE_SIZE, CONTROL_SIZE) pair to every kernel family
#if defined(GATO_ONLY_IIWA)   // development builds: the (14, 7) kernels only (a fifth of the compile time)
#define GATO_FOR_SIZES(X) X(14, 7)
#elif GATO_EXTRA_S > 0
#define GATO_FOR_SIZES(X) X(2, 1) X(4, 2) X(7, 3) X(14, 7) X(GATO_EXTRA_S, GATO_EXTRA_C)
#else
#define GATO_FOR_SIZES(X) X(2, 1) X(4, 2) X(7, 3) X(14, 7)
#endif

bool size_supported(int S, int C) {
#define X(s, c) if (S == s && C == c) return true;
    GATO_FOR_SIZES(X)
#undef X
#if GATO_EXTRA_S > 0
    if (S == GATO_EXTRA_S && C == GATO_EXTRA_C) return true;
#endif
    return false;
}

int launch_solve(const gato_plan* pl, const gato::SolveParams& p, cudaStream_t st) {
#define X(s, c) if (pl->S == s && pl->C == c) return launch_sc<s, c>(pl, p, st);
    GATO_FOR_SIZES(X)
#undef X
#if GATO_EXTRA_S > 0
    if (pl->S == GATO_EXTRA_S && pl->C == GATO_EXTRA_C)
        return launch_sc<GATO_EXTRA_S, GATO_EXTRA_C>(pl, p, st);
#endif
    return fail(GATO_EUNSUPPORTED, "no kernels for S=%d C=%d", pl->S, pl->C);
}

int launch_unpack(const gato_plan* pl, const float* dG, const float* dC, float* Gd, float* Cd,
                  float rho, cudaStream_t st) {
#define X(s, c) if (pl->S == s && pl->C == c) return unpack_sc<s, c>(pl, dG, dC, Gd, Cd, rho, st);
    GATO_FOR_SIZES(X)
#undef X
#if GATO_EXTRA_S > 0
    if (pl->S == GATO_EXTRA_S && pl->C == GATO_EXTRA_C)
        return unpack_sc<GATO_EXTRA_S, GATO_EXTRA_C>(pl, dG, dC, Gd, Cd, rho, st);
#endif
    return fail(GATO_EUNSUPPORTED, "no kernels for S=%d C=%d", pl->S, pl->C);
}

int cfg_for(const gato_plan* pl, LaunchCfg* cfg) {
#define X(s, c) if (pl->S == s && pl->C == c) return pick_cfg<s, c>(pl, cfg);
    GATO_FOR_SIZES(X)
#undef X
#if GATO_EXTRA_S > 0
    if (pl->S == GATO_EXTRA_S && pl->C == GATO_EXTRA_C)
        return pick_cfg<GATO_EXTRA_S, GATO_EXTRA_C>(pl, cfg);
#endif
    return fail(GATO_EUNSUPPORTED, "no kernels for S=%d C=%d", pl->S, pl->C);
}


int v2_nt_for(const gato_plan* pl) {
#define X(s, c) if (pl->S == s && pl->C == c) return v2_pick_nt<s, c>(pl);
    GATO_FOR_SIZES(X)
#undef X
    return 0;
}

bool use_v2(const gato_plan* pl) {
    if (pl->kernel == 1 || pl->kernel == 3 || pl->kernel == 4 || pl->kernel == 5 || (pl->kernel == 0 && pl->cl_override > 0))
        return false;
    return v2_nt_for(pl) > 0;
}

bool v5_ok(const gato_plan* pl) {
#define X(s, c) if (pl->S == s && pl->C == c) return v5_ok_sc<s, c>(pl) != 0;
    GATO_FOR_SIZES(X)
#undef X
    return false;
}

bool use_v5(const gato_plan* pl) { return pl->kernel == 5 && pl->precond == 0 && v5_ok(pl); }

// automatic choice for a call of `nsys` systems: kernel 5 for batches that fill the device (>= two systems per SM) (its two kernels keep two to
// three systems per SM in flight: measured 2.3 M vs 2.0 M solves/s at 8192 IIWA-50 systems), the one-CTA kernel for
// single solves and small batches (one launch; the zero-copy single solve needs its TMA staging)
bool use_v5_for(const gato_plan* pl, int nsys, int warm) {
    if (pl->precond != 0) return false;      // kernel 5 is the stair preconditioner in scaled coordinates: block-Jacobi runs on kernels 1-4
    if (pl->kernel == 5) return v5_ok(pl);
    (void)warm;   // (true warm start included since the first half forms r0 = gamma - S lambda0)
    if (pl->kernel != 0 || pl->cl_override > 0) return false;
    if (pl->batch_hint > 0) nsys = pl->batch_hint;
    return nsys >= 2 * (pl->sm_count > 0 ? pl->sm_count : 148) && v5_ok(pl);   // below two waves of the one-CTA kernel it is not ahead
}

long long v5_scratch_floats(const gato_plan* pl) {
#define X(s, c) if (pl->S == s && pl->C == c) return v5_scratch_per_system<s, c>(pl);
    GATO_FOR_SIZES(X)
#undef X
    return 0;
}

int v2cl_cfg_for(const gato_plan* pl, ClusterCfg* cfg) {
#define X(s, c) if (pl->S == s && pl->C == c) return v2cl_cfg<s, c>(pl, cfg);
    GATO_FOR_SIZES(X)
#undef X
    return fail(GATO_EUNSUPPORTED, "no kernels for S=%d C=%d", pl->S, pl->C);
}

int v2grid_cfg_for(const gato_plan* pl, ClusterCfg* cfg) {
#define X(s, c) if (pl->S == s && pl->C == c) return v2grid_cfg<s, c>(pl, cfg);
    GATO_FOR_SIZES(X)
#undef X
    return fail(GATO_EUNSUPPORTED, "no kernels for S=%d C=%d", pl->S, pl->C);
}

// kernel 3 (the one-CTA kernel spread over a cluster): explicit choice, or automatic for horizons the
// one-CTA kernel cannot hold
bool use_v2cl(const gato_plan* pl) {
    if (pl->kernel == 3) return true;
    if (pl->kernel != 0 || pl->cl_override > 0 || v2_nt_for(pl) > 0) return false;
    if (pl->S <= 8) return false;   // small states: 1024-thread CTAs make its barriers dearer than kernel 1's (measured)
    ClusterCfg cfg;
    const std::string keep = g_err;
    const bool ok = v2cl_cfg_for(pl, &cfg) == GATO_OK;
    g_err = keep;
    return ok;
}

// kernel 4 (the same over a cooperative grid): explicit choice, or automatic when nothing else holds the horizon
bool use_v2grid(const gato_plan* pl) {
    if (pl->kernel == 4) return true;
    if (pl->kernel != 0 || pl->cl_override > 0 || use_v2(pl) || use_v2cl(pl)) return false;
    const std::string keep = g_err;
    LaunchCfg c1;
    ClusterCfg cg_;
    const bool v1_ok = cfg_for(pl, &c1) == GATO_OK;
    const bool ok = !v1_ok && v2grid_cfg_for(pl, &cg_) == GATO_OK;
    g_err = keep;
    return ok;
}

// does the one-CTA kernel run with TMA staging for this plan?
template <int S, int C>
bool v2_stages_sc(const gato_plan* pl) {
    using L = gato::Layout2<S, C>;
    const int nt = v2_pick_nt<S, C>(pl);
    if (nt == 0 || !pl->staging) return false;
    const int stage = gato::stage_len(pl->nnzC) + gato::stage_len(pl->nnzG) + gato::stage_len(pl->n_z);
    return (size_t)(L::smem_floats(pl->K, nt) + stage + 8) * 4 <= (size_t)pl->smem_optin;
}

bool v2_stages(const gato_plan* pl) {
    if (!use_v2(pl)) return false;
#define X(s, c) if (pl->S == s && pl->C == c) return v2_stages_sc<s, c>(pl);
    GATO_FOR_SIZES(X)
#undef X
    return false;
}

int launch_solve_v2(const gato_plan* pl, const gato::SolveParams& p, cudaStream_t st) {
    gato::SolveParams2 q = {};
    q.p = p;
    q.gdst = pl->d_gdst2; q.cdst = pl->d_cdst2;
    q.cperm = pl->d_cperm; q.ncperm = pl->ncperm;
    q.nnzG = pl->nnzG; q.nnzC = pl->nnzC; q.g_diag = pl->g_diag;
#define X(s, c) if (pl->S == s && pl->C == c) return launch_v2_sc<s, c>(pl, q, st);
    GATO_FOR_SIZES(X)
#undef X
    return fail(GATO_EUNSUPPORTED, "no v2 kernels for S=%d C=%d", pl->S, pl->C);
}

int launch_solve_v5(gato_plan* pl, const gato::SolveParams& p, cudaStream_t st, bool first_only = false) {
    gato::SolveParams2 q = {};
    q.p = p;
    q.gdst = pl->d_gdst2; q.cdst = pl->d_cdst2;
    q.cperm = pl->d_cperm; q.ncperm = pl->ncperm;
    q.nnzG = pl->nnzG; q.nnzC = pl->nnzC; q.g_diag = pl->g_diag;
#define X(s, c) if (pl->S == s && pl->C == c) return launch_v5_sc<s, c>(pl, q, st, first_only);
    GATO_FOR_SIZES(X)
#undef X
    return fail(GATO_EUNSUPPORTED, "no kernel 5 for S=%d C=%d", pl->S, pl->C);
}

int launch_solve_v2cl(const gato_plan* pl, const gato::SolveParams& p, cudaStream_t st) {
    gato::SolveParams2 q = {};
    q.p = p;
    q.gdst = pl->d_gdst2; q.cdst = pl->d_cdst2;
    q.cperm = pl->d_cperm; q.ncperm = pl->ncperm;
    q.nnzG = pl->nnzG; q.nnzC = pl->nnzC; q.g_diag = pl->g_diag;
#define X(s, c) if (pl->S == s && pl->C == c) return launch_v2cl_sc<s, c>(pl, q, st);
    GATO_FOR_SIZES(X)
#undef X
    return fail(GATO_EUNSUPPORTED, "no v2 kernels for S=%d C=%d", pl->S, pl->C);
}

void v2_maps_for(const gato_plan* pl, const std::vector<int>& gmap, const std::vector<int>& gstart,
                 const std::vector<int>& cmap, const std::vector<int>& cstart, std::vector<int>* gdst,
                 std::vector<int>* cdst) {
#define X(s, c) if (pl->S == s && pl->C == c) return v2_fill_maps<s, c>(pl, gmap, gstart, cmap, cstart, gdst, cdst);
    GATO_FOR_SIZES(X)
#undef X
}

int launch_solve_v2grid(gato_plan* pl, const gato::SolveParams& p, cudaStream_t st) {
    gato::SolveParams2 q = {};
    q.p = p;
    q.gdst = pl->d_gdst2; q.cdst = pl->d_cdst2;
    q.cperm = pl->d_cperm; q.ncperm = pl->ncperm;
    q.nnzG = pl->nnzG; q.nnzC = pl->nnzC; q.g_diag = pl->g_diag;
#define X(s, c) if (pl->S == s && pl->C == c) return launch_v2grid_sc<s, c>(pl, q, st);
    GATO_FOR_SIZES(X)
#undef X
    return fail(GATO_EUNSUPPORTED, "no v2 kernels for S=%d C=%d", pl->S, pl->C);
}

// the kernel the plan selects; dumps of the Schur intermediates exist only in the cluster kernel
int launch_auto(gato_plan* pl, const gato::SolveParams& p, cudaStream_t st) {
    if (p.dbg_S == nullptr && use_v5_for(pl, p.nsys, p.warm)) return launch_solve_v5(pl, p, st);
    if (p.dbg_S == nullptr && use_v2(pl)) return launch_solve_v2(pl, p, st);
    if (p.dbg_S == nullptr && use_v2cl(pl)) return launch_solve_v2cl(pl, p, st);
    if (p.dbg_S == nullptr && use_v2grid(pl)) return launch_solve_v2grid(pl, p, st);
    return launch_solve(pl, p, st);
}

}  // namespace

#include "gato_capi_entry.inc"

// ---- kernel 5 without any allocation: its scratch in a file-scope device array ----------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops).  Everything else in this library sizes its scratch with cudaMalloc on first
// use (pl->dscr); this entry point uses ONE fixed array that every GPU which loads the library holds until exit.  All
// calls share it and it keeps the last call's contents: calls that overlap on the GPU (two streams, two plans, two host
// threads) race on it.
namespace {
constexpr long long kNoAllocFloats = 2ll << 20;          // 8 MiB
constexpr long long kNoAllocTail = 1024;                 // floats behind the systems' scratch: the second half's work counter
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];

// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address: the address a kernel can
// use comes from cudaGetSymbolAddress, once per device, with that device current.
int noalloc_scratch_ptr(int device, float** out) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static float* ptr[kMaxDev];
    if (device < 0 || device >= kMaxDev) return fail(GATO_EINVAL, "device %d: the no-allocation path serves devices 0..%d", device, kMaxDev - 1);
    std::lock_guard<std::mutex> lk(mu);
    if (ptr[device] == nullptr) {
        int cur = -1;
        GATO_CUDA(cudaGetDevice(&cur));
        if (cur != device) return fail(GATO_EINVAL, "device %d is not current (current: %d)", device, cur);
        float* p = nullptr;
        GATO_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (p == nullptr) return fail((int)cudaErrorInvalidSymbol, "cudaGetSymbolAddress returned a null address");
        ptr[device] = p;
    }
    *out = ptr[device];
    return GATO_OK;
}

template <int S, int C>
int launch_v5_noalloc_sc(gato_plan* pl, gato::SolveParams2 q, cudaStream_t st) {
    if (!v5_ok_sc<S, C>(pl)) return fail(GATO_EUNSUPPORTED, "kernel 5 does not hold S=%d K=%d", pl->S, pl->K);
    const long long per = v5_scratch_per_system<S, C>(pl);
    const long long cap = (kNoAllocFloats - kNoAllocTail) / per;
    if (q.p.nsys > cap)
        return fail(GATO_EINVAL, "nsys=%d: the fixed scratch (%lld bytes) holds at most %lld systems of K=%d (%lld bytes each)",
                    q.p.nsys, kNoAllocFloats * 4, cap, pl->K, per * 4);
    float* scr = nullptr;
    int rc = noalloc_scratch_ptr(pl->device, &scr);
    if (rc) return rc;
    q.scr = scr;
    q.gbuf = scr + (kNoAllocFloats - kNoAllocTail);
    q.scr_ld = per;
    // what must start at zero: the second half's work counter (the first half also resets it; the array keeps the last
    // call's contents).  Every scratch word the second half reads is written by the first half of the same call.
    GATO_CUDA(cudaMemsetAsync(q.gbuf, 0, 256, st));
    const int sms = pl->sm_count > 0 ? pl->sm_count : 148;
    const bool lat = 4 * pl->K <= 224 && (pl->v5_variant == 1 || (pl->v5_variant == 0 && q.p.nsys < sms));   // as launch_v5_sc decides
    if ((rc = launch_v5_first<S, C>(pl, q, st))) return rc;
    return launch_v5_second<S, C>(pl, q, st, lat);
}
}  // namespace

extern "C" {

int gato_noalloc_max_systems(const gato_plan* pl) {
    if (!pl) return fail(GATO_EINVAL, "plan is NULL");
    if (!v5_ok(pl)) return 0;
    const long long per = v5_scratch_floats(pl);
    return per > 0 ? (int)((kNoAllocFloats - kNoAllocTail) / per) : 0;
}

int gato_solve_batched_device_noalloc(gato_plan* pl, int nsys, const float* dG_val, long long ldG, const float* dC_val,
                                      long long ldC, const float* dg, long long ldg, const float* dc, long long ldc,
                                      float* dlambda_out, float* ddz_out, int* diters_out, float exit_tol, int max_iters,
                                      float rho, void* stream) {
    int rc = check_solve_args(pl, nsys, max_iters, 0, nullptr);
    if (rc) return rc;
    if (!dG_val || !dC_val || !dg || !dc || !dlambda_out || !ddz_out || !diters_out)
        return fail(GATO_EINVAL, "NULL device array");
    if (pl->precond != 0) return fail(GATO_EUNSUPPORTED, "the no-allocation path is kernel 5: stair preconditioner only");
    DeviceGuard guard(pl->device);
    gato::SolveParams2 q = {};
    gato::SolveParams& p = q.p;
    p.gmap = pl->d_gmap; p.cmap = pl->d_cmap; p.gstart = pl->d_gstart; p.cstart = pl->d_cstart;
    p.G_val = dG_val; p.ldG = ldG; p.C_val = dC_val; p.ldC = ldC;
    p.g = dg; p.ldg = ldg; p.c = dc; p.ldc = ldc;
    p.lambda_out = dlambda_out; p.ldlo = pl->n_l;
    p.dz_out = ddz_out; p.lddz = pl->n_z;
    p.iters_out = diters_out;
    p.K = pl->K; p.nsys = nsys; p.max_iters = max_iters; p.warm = 0;
    p.rho = rho; p.exit_tol = exit_tol;
    q.gdst = pl->d_gdst2; q.cdst = pl->d_cdst2;
    q.cperm = pl->d_cperm; q.ncperm = pl->ncperm;
    q.nnzG = pl->nnzG; q.nnzC = pl->nnzC; q.g_diag = pl->g_diag;
#define X(s, c) if (pl->S == s && pl->C == c) return launch_v5_noalloc_sc<s, c>(pl, q, (cudaStream_t)stream);
    GATO_FOR_SIZES(X)
#undef X
    return fail(GATO_EUNSUPPORTED, "no kernel 5 for S=%d C=%d", pl->S, pl->C);
}

}  // extern "C"
