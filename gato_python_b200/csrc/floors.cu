// floors.cu -- micro-benchmarks of the latency floors a single KKT solve runs against (SURVEY.md 8d:
// "per-PCG-iteration latency against the barrier/launch floor").  Stand-alone binary, prints one JSON object.
//
//   launch_us           null kernel, launch -> completion, CUDA events (median)
//   syncthreads_ns      one __syncthreads() of a 224-thread CTA (7 warps, the IIWA-50 shape)
//   cluster_sync_ns[n]  one barrier.cluster arrive.release + wait.acquire over n CTAs of 224 threads
//   dsmem_push_sync_ns  one 16-byte DSMEM store to the neighbour CTA followed by that cluster barrier
//   reduce_chain_ns     the dot-product chain of one PCG reduction: 5 dependent warp shuffles + partial to
//                       shared memory + block barrier + fixed-order sum of the 7 partials + reciprocal
//   grid_sync_ns        cooperative-groups grid.sync() over 50 blocks of 32 threads -- what the reference's
//                       PCG pays four times per iteration (src/gato_pcg.cuh:315-437)
//
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -rdc=true floors.cu -o floors
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <vector>

namespace cg = cooperative_groups;

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            std::fprintf(stderr, "%s: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return 1;                                                                          \
        }                                                                                      \
    } while (0)

__device__ __forceinline__ long long gtime() {
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__global__ void k_null() {}

__global__ void k_syncthreads(int reps, long long* out) {
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < reps; ++i) __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[0] = t1 - t0;
}

__global__ void k_cluster_sync(int reps, int push, long long* out) {
    extern __shared__ float sm[];
    cg::cluster_group cl = cg::this_cluster();
    const int r = (int)cl.block_rank(), n = (int)cl.dim_blocks().x;
    float* rem = cl.map_shared_rank(sm, (r + 1) % n);
    cl.sync();
    const long long t0 = clock64();
    for (int i = 0; i < reps; ++i) {
        if (push && threadIdx.x < 4) reinterpret_cast<float4*>(rem)[threadIdx.x] = make_float4(i, i, i, i);
        cl.sync();
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
    if (sm[0] == -1.f) out[1] = 0;   // keep the stores alive
}

__global__ void k_reduce_chain(int reps, long long* out, float* sink) {
    __shared__ __align__(16) float red[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float x = 1.0f + threadIdx.x * 1e-3f, eta = 1.f;
    if (threadIdx.x < 8) red[threadIdx.x] = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < reps; ++i) {
        float v = x * eta;
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        if (lane == 0) red[warp] = v;
        __syncthreads();
        const float4 a = reinterpret_cast<const float4*>(red)[0], b = reinterpret_cast<const float4*>(red)[1];
        const float s = ((a.x + b.x) + (a.y + b.y)) + ((a.z + b.z) + (a.w + b.w));
        float rc;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(s));
        eta = fmaf(fmaf(-s, eta * rc, eta), rc, eta * rc);
        __syncthreads();   // partials are overwritten next round (the PCG loop has its exchange barrier here)
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) { out[0] = t1 - t0; sink[0] = eta; }
}

// FP32 FMA issue: 8 independent accumulator chains per thread, scalar FFMA vs packed FFMA2 (fma.rn.f32x2)
template <bool PACKED>
__global__ void k_fma(int reps, long long* out, float* sink) {
    float a[16];
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
    const float b = 1.0001f, c = 1e-3f;
    unsigned long long p[8], bb, cc;
    asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b), "f"(b));
    asm("mov.b64 %0, {%1, %2};" : "=l"(cc) : "f"(c), "f"(c));
    for (int i = 0; i < 8; ++i) asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(a[2 * i]), "f"(a[2 * i + 1]));
    __syncthreads();
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
        if (PACKED) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(p[i]) : "l"(p[i]), "l"(bb), "l"(cc));
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(a[i]) : "f"(a[i]), "f"(b), "f"(c));
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
    for (int i = 0; i < 16; ++i) s += a[i];
    for (int i = 0; i < 8; ++i) s += __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32));
    if (threadIdx.x == 0) out[0] = t1 - t0;
    sink[threadIdx.x] = s;
}

__global__ void k_grid_sync(int reps, long long* out) {
    cg::grid_group g = cg::this_grid();
    g.sync();
    const long long t0 = gtime();
    for (int i = 0; i < reps; ++i) g.sync();
    const long long t1 = gtime();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
}

int main() {
    int dev = 0, khz = 0;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    long long* d_out;
    float* d_sink;
    CK(cudaMalloc(&d_out, 64));
    CK(cudaMalloc(&d_sink, 4096));
    long long h[2];
    const int reps = 2000;

    // launch latency
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 50; ++i) k_null<<<1, 32>>>();
    CK(cudaDeviceSynchronize());
    std::vector<float> ts;
    for (int i = 0; i < 300; ++i) {
        CK(cudaEventRecord(e0));
        k_null<<<1, 32>>>();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        ts.push_back(ms * 1e3f);
    }
    std::sort(ts.begin(), ts.end());
    const float launch_us = ts[ts.size() / 2];

    // clock64 ticks at the SM clock; measure the actual rate against %globaltimer is overkill here: report
    // cycles and convert with the attribute clock (the boxes run at max clock under these tiny loads)
    auto ns = [&](long long cyc) { return (double)cyc / reps * 1e6 / (double)khz; };

    k_syncthreads<<<1, 224>>>(reps, d_out);
    CK(cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost));
    const double sync_ns = ns(h[0]), sync_cyc = (double)h[0] / reps;

    k_reduce_chain<<<1, 224>>>(reps, d_out, d_sink);
    CK(cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost));
    const double chain_ns = ns(h[0]);

    // 16 FMAs per thread per round either way (16 FFMA or 8 FFMA2), 8 warps per SM = 2 per scheduler
    double fma_cyc[2];
    for (int pk = 0; pk < 2; ++pk) {
        if (pk) k_fma<true><<<1, 256>>>(reps, d_out, d_sink); else k_fma<false><<<1, 256>>>(reps, d_out, d_sink);
        CK(cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost));
        fma_cyc[pk] = (double)h[0] / reps;
    }

    std::printf("{\"sm_clock_khz\": %d, \"launch_us\": %.2f, \"syncthreads_ns\": %.1f, \"syncthreads_cycles\": %.1f, "
                "\"reduce_chain_ns\": %.1f, \"cycles_per_16_fma_per_thread_2_warps_per_scheduler\": {\"ffma\": %.1f, \"ffma2\": %.1f}",
                khz, launch_us, sync_ns, sync_cyc, chain_ns, fma_cyc[0], fma_cyc[1]);

    CK(cudaFuncSetAttribute(k_cluster_sync, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    for (int push = 0; push < 2; ++push) {
        std::printf(", \"%s\": {", push ? "dsmem_push_sync_ns" : "cluster_sync_ns");
        const int sizes[] = {1, 2, 4, 5, 8, 16};
        bool first = true;
        for (int n : sizes) {
            cudaLaunchConfig_t lc = {};
            lc.gridDim = dim3(n);
            lc.blockDim = dim3(224);
            lc.dynamicSmemBytes = 1024;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = n; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            lc.attrs = at;
            lc.numAttrs = 1;
            if (cudaLaunchKernelEx(&lc, k_cluster_sync, reps, push, d_out) != cudaSuccess) { cudaGetLastError(); continue; }
            if (cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); continue; }
            std::printf("%s\"%d\": %.1f", first ? "" : ", ", n, ns(h[0]));
            first = false;
        }
        std::printf("}");
    }

    {   // the reference's synchronisation primitive: cooperative grid.sync over 50 blocks of 32 threads
        int r = reps;
        void* args[] = {&r, &d_out};
        if (cudaLaunchCooperativeKernel((void*)k_grid_sync, dim3(50), dim3(32), args, 0, 0) == cudaSuccess &&
            cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost) == cudaSuccess)
            std::printf(", \"grid_sync_ns\": %.1f", (double)h[0] / reps);
        else
            cudaGetLastError();
    }
    std::printf("}\n");
    return 0;
}
