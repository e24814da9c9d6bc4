/* gato_b200.h -- C ABI of the B200-native GATO KKT linear-system hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  Every entry point
 * cites the reference interface (relative to /root/reference) that it replaces.  All entry points
 * return 0 on success or a negative GATO_E* / positive cudaError_t code; they never call exit()
 * (the reference's gpuErrchk does, include/gato_defines.h:42-51).  gato_last_error() returns a
 * thread-local description of the last failure.
 *
 * Data conventions: FP32 values, int32 indices, CSR in SciPy order
 * (row = indptr, col = indices, val = data); g / dz have (S+C)K-C entries laid out per knot as
 * [x(S) | u(C)] (last knot x only); c / lambda have S*K entries.
 */
#ifndef GATO_B200_H_
#define GATO_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

#define GATO_OK 0
#define GATO_EINVAL (-1)      /* bad argument / size mismatch                                  */
#define GATO_EPATTERN (-2)    /* CSR pattern has entries the block layout cannot hold          */
#define GATO_EUNSUPPORTED (-3)/* (S, C) not compiled in, or K too large for one cluster        */
#define GATO_ENODEVICE (-4)   /* no CUDA device / driver: there is NO CPU fallback              */

/* warm_start modes for the solve entry points */
#define GATO_WARM_OFF 0        /* lambda_in ignored                                             */
#define GATO_WARM_REFERENCE 1  /* what the reference does with warm_start=True: lambda_in is
                                  accepted and has NO numerical effect (src/gato_pcg.cuh:300-304) */
#define GATO_WARM_TRUE 2       /* extension: r0 = gamma - S*lambda_in                            */

#define GATO_WARM_RESIDENT 3   /* extension, gato_solve only: GATO_WARM_TRUE with lambda_in = the lambda
                                  this plan's previous gato_solve left on the device (nothing is
                                  uploaded; the first solve of a plan starts cold)                 */

typedef struct gato_plan gato_plan; /* opaque: sparsity pattern maps + persistent device workspace */

/* Concurrency.  A plan owns staging buffers, scratch and streams: use it from ONE host thread at a time (the pybind11
 * module serialises its callers with a mutex; gato_solve_batched_multi gives every device its own plan and thread).
 * gato_solve_batched_device is asynchronous and may be issued on several streams of one plan from that thread:
 * kernels 1-3 keep no per-plan device state, kernel 5 keeps one scratch buffer per stream it has seen (at most 8),
 * kernel 4 (cooperative grid) keeps ONE scratch buffer per plan, so its launches must be on one stream (the chunk
 * pipeline of gato_solve_batched switches itself off for it). */

/* Which (STATE_SIZE, CONTROL_SIZE) pairs this build of the library carries kernels for.
 * Replaces the compile-time knobs of CMakeLists.txt:18 / install.bash:4-19 at library level; the
 * pybind11 module `gpu_library` still fixes S, C, K at compile time.  Returns the number of pairs;
 * writes up to `cap` pairs into sizes[2*i], sizes[2*i+1]. */
int gato_supported_sizes(int* sizes, int cap);

const char* gato_last_error(void);
int gato_device_count(void);

/* Build a plan for one sparsity pattern on CUDA device `device`.
 * Replaces the per-call index arithmetic of csr_to_custom_G / csr_to_custom_C
 * (src/gato_schur.cuh:674-743) by a one-off host pass that turns each CSR non-zero into its
 * destination inside the per-knot dense block, and the per-call cudaMalloc/cudaFree of
 * gato_linsys (gpu_library.cu:36-45, 74-81) and main_call (gpu_library.cu:140-157, 211-218) by a
 * persistent workspace sized for `max_batch` systems.  Host pointers. */
int gato_plan_create(int S, int C, int K,
                     const int* G_row, const int* G_col, int nnzG,
                     const int* C_row, const int* C_col, int nnzC,
                     int device, int max_batch, gato_plan** plan_out);

/* Block-input plan: no CSR at all.  The `G_val` / `C_val` arguments of every solve entry point are
 * then the reference's own dense device layouts (include/gato_defines.h:36-37):
 *   G_val = G dense, (S^2+C^2)K - C^2 floats: knot k at k(S^2+C^2): Q_k col-major, then R_k col-major
 *   C_val = C dense, (S^2+SC)(K-1) floats:    stage k at k(S^2+SC): A_k col-major, then B_k col-major,
 *           A_k / B_k as stored in C's block row k+1 (sign included, i.e. -A, -B for x+ = Ax + Bu)
 * i.e. what gato_convert_kkt_format (src/gato_schur.cuh:745-756) would have produced (rho is still
 * added to the Q/R diagonals by the solve).  Replaces the caller-side dense -> CSR conversion of
 * test_IIWA50.py:11-18 plus the CSR -> dense kernel.  Results equal those of a CSR plan over the
 * fully dense pattern. */
int gato_plan_create_blocks(int S, int C, int K, int device, int max_batch, gato_plan** plan_out);
void gato_plan_destroy(gato_plan* plan);

/* Cluster shape override for tuning: `cluster_ctas` CTAs per system (0 = automatic). */
int gato_plan_set_cluster(gato_plan* plan, int cluster_ctas);
int gato_plan_get_cluster(const gato_plan* plan, int* cluster_ctas, int* knots_per_cta);

/* Kernel selection for tuning and tests: 0 = automatic, 1 = warp-per-knot cluster kernel (one thread-block
 * cluster per system, DSMEM exchange; any K up to 16 CTAs x 32 knots), 2 = one-CTA-per-system kernel (whole
 * system resident in one SM's registers, block barriers only; the batched path and the default whenever K
 * fits), 3 = the one-CTA kernel spread over a thread-block cluster (each CTA owns a run of knots plus two
 * halo knots; neighbour slices and dot-product partials pushed through DSMEM, cluster barriers; the default
 * for IIWA-sized states once K exceeds one SM, e.g. K = 256 -> 9 CTAs x 29 knots), 4 = the same decomposition
 * over a cooperative grid (slices and partials through a global scratch buffer, grid.sync; the default when no
 * cluster holds the horizon, e.g. K = 1000 -> 34 CTAs x 30 knots, up to one CTA per SM; this is the job of the
 * reference's grid-stride parallelPCG_fixed, src/gato_pcg.cuh:17-268, 529-553).  All replace the same
 * reference functions (parallelPCG, src/gato_pcg.cuh:441-470, and the three Schur/dz kernels,
 * src/gato_schur.cuh:462-494, 652-670, 869-879). */
int gato_plan_set_kernel(gato_plan* plan, int kernel);
int gato_plan_get_kernel(const gato_plan* plan);

/* One-CTA kernel only: stage each system's value arrays into shared memory with TMA bulk copies
 * (cp.async.bulk; the next system's copy overlaps the current solve in the persistent CTA) whenever the
 * staging area fits beside the system (default on); 0 = always read them with per-thread loads.  Same
 * results bit for bit.  Replaces the per-row scalar loads of csr_to_custom_G / csr_to_custom_C
 * (src/gato_schur.cuh:674-743). */
int gato_plan_set_staging(gato_plan* plan, int on);

/* Preconditioner of the PCG loop.  The reference selects it at compile time (include/gato_defines.h:9-11:
 * BLOCK_J_PRECON, SS_PRECON, PRECONDITIONER_BANDWIDTH "must be set for compile"; consulted at
 * src/gato_schur.cuh:407-429 and :496, :965) and ships 1 / 1 / 3 = the symmetric stair preconditioner, the only
 * setting for which it initialises every block of d_Pinv it reads.  Here it is a plan switch:
 *   GATO_PRECOND_STAIR         P^-1 = D^-1 - D^-1 O D^-1 (default; the parity mode -- iteration counts match the
 *                              reference build)
 *   GATO_PRECOND_BLOCK_JACOBI  P^-1 = D^-1 = diag(theta_b^-1): what SS_PRECON = 0 intends (gato_form_ss not run,
 *                              off-diagonal blocks of Pinv taken as zero).  More iterations, each with one
 *                              exchange + barrier and 6 S^2 multiply-adds per knot less.  Same solution within the
 *                              exit tolerance; iteration counts differ from the reference build by design. */
#define GATO_PRECOND_STAIR 0
#define GATO_PRECOND_BLOCK_JACOBI 1
int gato_plan_set_preconditioner(gato_plan* plan, int preconditioner);
int gato_plan_get_preconditioner(const gato_plan* plan);

/* gato_solve only: when the kernel stages its inputs with TMA, let those bulk copies read the plan's pinned
 * host staging buffer in place over PCIe instead of copying it to the device first (default on).  Removes
 * one of the two DMA operations of the reference's bracket (the H2D of gpu_library.cu:150-157).  Same results. */
int gato_plan_set_zero_copy(gato_plan* plan, int on);

/* Chunk pipeline of gato_solve_batched: the batch is cut into chunks of `systems_per_chunk` systems
 * whose H2D copy, kernel and D2H copy run on rotating streams, so the copy of chunk i+1 overlaps the
 * solve of chunk i (the reference copies, solves and copies back strictly in sequence on stream 0,
 * gpu_library.cu:150-157, 174, 71-72).  0 = automatic (default), < 0 = off (one chunk). */
int gato_plan_set_pipeline(gato_plan* plan, int systems_per_chunk);

/* One KKT solve, HOST buffers in and out.  Replaces main_call + gato_linsys
 * (gpu_library.cu:25-83, 85-231): H2D of values, one fused kernel, D2H of lambda and dz.
 * lambda_in may be NULL unless warm_start == GATO_WARM_TRUE.  iters_out (may be NULL) receives the
 * count the reference only prints (gpu_library.cu:189-191): the 0-based index of the iteration
 * that met |r^T P^-1 r| < exit_tol, or max_iters.  device_ms_out (may be NULL) receives the CUDA
 * event time of H2D + kernel + D2H (the reference's timed bracket, gpu_library.cu:172-187). */
int gato_solve(gato_plan* plan,
               const float* G_val, const float* C_val, const float* g, const float* c,
               const float* lambda_in, float* lambda_out, float* dz_out,
               float exit_tol, int max_iters, int warm_start, float rho,
               int* iters_out, float* device_ms_out);

/* Batch of `nsys` independent systems over the plan's pattern, HOST buffers.  ld* are element
 * strides between consecutive systems; ld == 0 means the array is shared by all systems.
 * lambda_out is [nsys][S*K], dz_out is [nsys][(S+C)K-C], iters_out is [nsys] (may be NULL).
 * The reference has no batched entry point (gpu_library.cu:236-239 exports one function); this is
 * what a loop over linsys_solve would compute. */
int gato_solve_batched(gato_plan* plan, int nsys,
                       const float* G_val, long long ldG, const float* C_val, long long ldC,
                       const float* g, long long ldg, const float* c, long long ldc,
                       const float* lambda_in, long long ldl,
                       float* lambda_out, float* dz_out, int* iters_out,
                       float exit_tol, int max_iters, int warm_start, float rho,
                       float* device_ms_out);

/* MPC sequence (BASELINE.json config 5): `nsteps` CONSECUTIVE solves over the plan's pattern, HOST
 * buffers; step t uses row t of each input array (ld == 0: shared by all steps) and rho[t] (rho ==
 * NULL: rho0 for every step).  With warm_start == GATO_WARM_TRUE step t starts from the lambda of
 * step t-1, which never leaves the device (step 0 from lambda0, or cold if lambda0 == NULL); with
 * GATO_WARM_OFF / GATO_WARM_REFERENCE every step starts cold like the reference
 * (src/gato_pcg.cuh:300-304).  One H2D of all inputs, nsteps back-to-back launches on one stream,
 * one D2H of lambda_out [nsteps][S*K], dz_out [nsteps][(S+C)K-C], iters_out [nsteps].  Replaces a
 * host loop over linsys_solve (gpu_library.cu:85-231) that re-uploads and re-allocates every step. */
int gato_solve_sequence(gato_plan* plan, int nsteps,
                        const float* G_val, long long ldG, const float* C_val, long long ldC,
                        const float* g, long long ldg, const float* c, long long ldc,
                        const float* rho, float rho0, const float* lambda0,
                        float* lambda_out, float* dz_out, int* iters_out,
                        float exit_tol, int max_iters, int warm_start, float* device_ms_out);

/* Same, DEVICE buffers on the plan's device, asynchronous on `stream` (a cudaStream_t passed as
 * void*; NULL = legacy default stream, which is what the reference uses, include/gato_defines.h:22).
 * No allocation, no host synchronisation.  Replaces the internal seam gato_linsys
 * (gpu_library.cu:25-30), whose CSR/g/c arguments are device pointers. */
int gato_solve_batched_device(gato_plan* plan, int nsys,
                              const float* dG_val, long long ldG, const float* dC_val, long long ldC,
                              const float* dg, long long ldg, const float* dc, long long ldc,
                              const float* dlambda_in, long long ldl,
                              float* dlambda_out, float* ddz_out, int* diters_out,
                              float exit_tol, int max_iters, int warm_start, float rho,
                              void* stream);

/* CSR -> per-knot dense blocks only, HOST buffers (parity hook for gato_convert_kkt_format,
 * src/gato_schur.cuh:745-756).  G_dense_out has (S^2+C^2)K-C^2 floats, C_dense_out has
 * (S^2+SC)(K-1) floats, both in the reference's layout (include/gato_defines.h:36-37). */
int gato_unpack(gato_plan* plan, const float* G_val, const float* C_val, float rho,
                float* G_dense_out, float* C_dense_out);

/* One process, several GPUs (batches shard, no collective).  plans[i] must have been created for the
 * same problem on ndev distinct devices.  The batch is cut into contiguous slices, one host thread per device runs
 * gato_solve_batched on its slice of the caller's host arrays (a shared array, ld == 0, is read by every device), and
 * the results land in disjoint slices of lambda_out / dz_out / iters_out: the only "gather" is each device's own D2H.
 * device_ms_out receives the longest device time.  Replaces the reference's single-device assumption
 * (src/gato_utils.cuh:831, include/gato_defines.h:22). */
int gato_solve_batched_multi(gato_plan* const* plans, int ndev, int nsys, const float* G_val, long long ldG,
                             const float* C_val, long long ldC, const float* g, long long ldg, const float* c,
                             long long ldc, const float* lambda_in, long long ldl, float* lambda_out, float* dz_out,
                             int* iters_out, float exit_tol, int max_iters, int warm_start, float rho,
                             float* device_ms_out);

/* Kernel 5 ("scaled system", gato_plan_set_kernel(plan, 5); STATE_SIZE = 14, K <= 56): the hot path as TWO kernels.
 * The first forms the reference's Schur blocks theta_b, phi_b, gamma_b (src/gato_schur.cuh:13-494) and, in FP64, the
 * congruence S^ = M S M^T (theta_b = L_b L_b^T, M_b = L_b^-1) whose diagonal blocks are the identity; the second runs
 * the reference's PCG (src/gato_pcg.cuh:300-437, same update order, exit rule and iteration numbering) on
 * S^ = I + O^ with P^^-1 = I - O^ -- the stair preconditioner (src/gato_schur.cuh:498-649) in those coordinates -- and
 * recovers lambda = M^T lambda^ and dz (src/gato_schur.cuh:758-879).  Identical iterates in exact arithmetic; one
 * 7 x 7 block of psi_b = M_b phi_b M_{b-1}^T per lane instead of 196 matrix words, two systems resident per SM.
 * gato_scaled_layout: {floats per knot, region copy, psi offset, psi floats per lane, M offset, gamma^ offset, VS,
 * HALFP} of the per-knot scratch; gato_scaled_debug: run the first kernel on one system and return its scratch
 * (parity hook for psi, M, gamma^); gato_plan_set_scaled_chunk: systems per pass through the scratch buffer
 * (0 = automatic: two per SM, which keeps the scratch in L2). */
int gato_scaled_layout(const gato_plan* plan, int* layout8);
/* the kernel a call with `nsys` systems and this warm_start runs (automatic selection: kernel 5 for batches of at
 * least two systems per SM when it holds the problem, else what gato_plan_get_kernel reports) */
int gato_plan_get_kernel_for(const gato_plan* plan, int nsys, int warm_start);
int gato_scaled_debug(gato_plan* plan, const float* G_val, const float* C_val, const float* g, const float* c,
                      float rho, float* scratch_out, long long nfloats);
int gato_plan_set_scaled_chunk(gato_plan* plan, int systems_per_chunk);
/* Kernel 5 has two second halves over the same scratch: a THROUGHPUT kernel (two lanes per knot, four warps per system,
 * three systems per SM) and a LATENCY kernel (four lanes per knot, seven warps per system: a lone system iterates
 * ~1.6 x faster).  0 = chosen by the size of the whole call (latency below one system per SM), 1 / 2 = forced.  The two
 * agree within the rounding of their summation orders, not bit for bit. */
int gato_plan_set_scaled_variant(gato_plan* plan, int variant);
/* Kernel 5, batched calls whose systems share ONE matrix pair (ldG == 0 and ldC == 0: the reference's G / C with many
 * right-hand sides g, c -- sampling, rollouts): theta, phi, the factor M and psi then belong to the call, so the first
 * half (src/gato_schur.cuh:26-496 plus the congruence) runs once, gato_gamma_scaled forms gamma^ = M gamma of every
 * system, and the second half reads the one scratch.  on = 1 (default) takes that path with the throughput second
 * half; on = 0 recomputes everything per system (the path the parity tests compare it with). */
int gato_plan_set_shared_matrix_path(gato_plan* plan, int on);
/* NO COUNTERPART IN THE REFERENCE.  For callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it): kernel 5 on device-resident inputs and
 * outputs, as gato_solve_batched_device, with its scratch (unpack + Schur + congruence results handed from its first
 * kernel to its second) in ONE file-scope `static __device__` array of GATO_NOALLOC_SCRATCH_BYTES instead of a buffer
 * the plan grows with cudaMalloc.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine);
 * the plan itself was allocated when it was created.  Limits: stair preconditioner, no true warm start (lambda_in is not
 * taken), S = 14 and K <= 56 (kernel 5), and nsys <= gato_noalloc_max_systems(plan) = (8 MiB - 4 KiB) / (K * 4352 B):
 * 38 systems at K = 50; beyond that GATO_EINVAL with the limit in gato_last_error().  Every GPU that loads the library
 * holds the 8 MiB until exit.  ALL calls on a device share the array and it keeps the last call's contents (the call
 * clears the one word that must start at zero, the work counter): calls that overlap on the GPU -- two streams, two
 * plans, two host threads -- race on it; order them on one stream.  Results: the same bits as gato_solve_batched_device
 * with kernel 5 selected (same kernels, same second-half variant rule).  Under stream capture: make one call outside the
 * capture first (it resolves the array's device address and sets the kernels' attributes, once per device and plan);
 * later calls enqueue a memset node and two kernel nodes and nothing else (tests/checks/noalloc_graph.py). */
#define GATO_NOALLOC_SCRATCH_BYTES (8ll << 20)
int gato_noalloc_max_systems(const gato_plan* plan);   /* 0: kernel 5 does not hold this plan's sizes */
int gato_solve_batched_device_noalloc(gato_plan* plan, int nsys, const float* dG_val, long long ldG, const float* dC_val,
                                      long long ldC, const float* dg, long long ldg, const float* dc, long long ldc,
                                      float* dlambda_out, float* ddz_out, int* diters_out, float exit_tol, int max_iters,
                                      float rho, void* stream);
/* Kernel 5, throughput second half: 0 = the update order of src/gato_pcg.cuh:348-421 (two reductions per iteration),
 * 1 = the single-reduction form of the same iteration (Chronopoulos-Gear: r~ = P^-1 r and w = S r~ first, then
 * gamma = r.r~ and delta = r~.w reduced together; s = S p by recurrence).  Same exit rule and iteration numbering;
 * identical in exact arithmetic. */
int gato_plan_set_pcg_form(gato_plan* plan, int form);
/* Kernel 5 launches two kernels per chunk; with profiling on, CUDA events are recorded on the launching stream
 * around each of them, and gato_plan_last_profile returns the device time of each half summed over the chunks of the
 * last call (it waits for that call's last event) and the number of kernel launches the call made. */
int gato_plan_set_profiling(gato_plan* plan, int on);
int gato_plan_last_profile(gato_plan* plan, float* first_half_ms, float* second_half_ms, int* launches);

/* Solve one system and also return the Schur-phase intermediates in the reference's device layout
 * and (negated) storage signs: S_out, Pinv_out 3*S^2*K floats (gpu_library.cu:40-41), gamma_out
 * S*K floats, Ginv_out (S^2+C^2)K-C^2 floats (what form_schur leaves in d_G_dense,
 * src/gato_schur.cuh:238-259).  clk_ns_out (16 entries) receives %globaltimer stamps of the
 * kernel's phase boundaries: 0 start, 1 smem cleared, 2 unpacked, 3 Q/R inverted, 4 Schur rows
 * formed, 5 stair blocks formed, 6 rows in registers, 7 PCG loop start, 8 PCG loop end, 9 dz
 * written.  Any of the five may be NULL.  Parity / profiling hook only. */
int gato_solve_debug(gato_plan* plan,
                     const float* G_val, const float* C_val, const float* g, const float* c,
                     float* lambda_out, float* dz_out,
                     float exit_tol, int max_iters, float rho, int* iters_out,
                     float* S_out, float* Pinv_out, float* gamma_out, float* Ginv_out,
                     long long* clk_ns_out);

#ifdef __cplusplus
}
#endif
#endif /* GATO_B200_H_ */
