/*
 * dgmm_b200.h -- C ABI of the B200-native deconvolution-GMM hot path.
 *
 * This is the drop-in boundary.  The reference (bayesiains/density-deconvolution) has no
 * FFI layer of its own: its hot path is ~70 lines of batched torch ops inside the Python
 * classes of deconv/gmm/.  Each entry point below names the reference lines it replaces
 * (paths relative to the reference checkout).  The Python shim that keeps the reference's
 * class API on top of these calls lives in density_deconvolution_b200/ and the ctypes
 * binding a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain C: pointers, sizes, a stream handle; no torch types.
 *   - all float32, row-major, exactly the reference's tensor layouts:
 *       X [N,D], noise S [N,D,D], means [K,D], covars [K,D,D].
 *   - "device" entry points take caller-owned DEVICE pointers, never allocate (one exception: for K > 256
 *     the component-chunked passes take (n_chunks + 1) N floats from the stream-ordered allocator), and are
 *     asynchronous on `stream` (a cudaStream_t passed as void*; NULL = default stream).
 *     Workspace is caller-provided; query its size with dgmm_workspace_bytes().
 *   - "host" entry points (suffix _host) take HOST pointers, stage through device memory
 *     they own inside an opaque dgmm_ctx, and return after the result is on the host.
 *   - return value: 0 = ok, <0 = argument / capability error (see DGMM_ERR_*), >0 = CUDA
 *     error code from the runtime (cudaError_t).  Numerical failure (some Sigma_j + S_i not
 *     positive definite) is NOT a return code: it is counted in `status[0]` (device int) or
 *     in the `nonpd` slot of the packed statistics, and the affected ll_i are -inf, which
 *     is how the reference reports it (deconv/gmm/deconv_gmm.py:44-47 returns -inf).
 *   - there is no CPU fallback anywhere in this library.
 */
#ifndef DGMM_B200_H
#define DGMM_B200_H

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define DGMM_API __attribute__((visibility("default")))
#else
#define DGMM_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define DGMM_ABI_VERSION 2
#define DGMM_MAX_D 32
#define DGMM_MAX_K 1024 /* one launch serves 256 components (512 for D <= 8); larger K runs in component chunks (see dgmm_estep_stats_f32) */

/* error codes (negative) */
#define DGMM_OK 0
#define DGMM_ERR_BAD_ARG (-1)        /* null pointer, N<0, K<1 ...                      */
#define DGMM_ERR_UNSUPPORTED_D (-2)  /* D outside 1..DGMM_MAX_D                         */
#define DGMM_ERR_UNSUPPORTED_K (-3)  /* K outside 1..DGMM_MAX_K                         */
#define DGMM_ERR_WORKSPACE (-4)      /* workspace too small / null                      */
#define DGMM_ERR_LAYOUT (-5)         /* unknown noise layout                            */
#define DGMM_ERR_NO_DEVICE (-6)      /* no CUDA device / not an sm_100 part             */
#define DGMM_ERR_STATIC_SCRATCH (-8) /* shape needs more scratch than the static array of a *_static_* entry point holds */
#define DGMM_ERR_LL_REQUIRED (-7)    /* dgmm_estep_stats_f32 needs out_ll for this shape (dgmm_stats_needs_ll) */

/* noise-covariance layouts (the reference contract is DGMM_NOISE_FULL:
 * deconv/gmm/data.py:7-17, deconv/gmm/sgd_deconv_gmm.py:31-41) */
#define DGMM_NOISE_FULL 0   /* S [N,D,D]; only the lower triangle is read (potrf 'L') */
#define DGMM_NOISE_DIAG 1   /* S [N,D]   = diagonal of each S_i                          */
#define DGMM_NOISE_NONE 2   /* S ignored (S_i = 0): SGDGMMModule.forward, sgd_gmm.py:49-62 */
#define DGMM_NOISE_SHARED 3 /* S [D,D]   one matrix for all points (data_gen.py:50-53)   */
#define DGMM_NOISE_PACKED 4 /* S [N,D(D+1)/2] lower triangle of each S_i, row-major: (0,0),(1,0),(1,1),(2,0)...
                             * the same information as DGMM_NOISE_FULL in 1.7-1.9x fewer bytes (HBM, PCIe and the
                             * kernels' shared-memory ring); dgmm_pack_noise_f32 / dgmm_pack_noise_host convert */

/* Packed statistics buffer written by dgmm_estep_stats_f32.
 *   [ s0 (K) | s1 (K*D) | S2 (K*D*D, full symmetric) | ll_hi | ll_lo | sum_g | nonpd ]
 * with omega_ij = g_i (r_ij + resp_floor), u_ij = T_ij^{-1}(x_i - mu_j):
 *   s0_j = sum_i omega_ij, s1_j = sum_i omega_ij u_ij, S2_j = sum_i omega_ij (u u^T - T^{-1}),
 *   ll_hi + ll_lo = sum_i ll_i (a float pair carrying a double through a float32
 *   all-reduce), sum_g = sum_i g_i, nonpd = number of non-PD (point,component) pairs.
 * Every slot is additive across shards: one all-reduce(SUM) of this buffer is the whole
 * cross-GPU exchange of an EM iteration or SGD step. */
static inline size_t dgmm_stats_len(int D, int K) {
    return (size_t)K * (size_t)(1 + D + D * D) + 4;
}

/* ---- capability / sizing ------------------------------------------------------------ */

DGMM_API int dgmm_abi_version(void);

/* Fills list[0..*n) with the supported dimensions (1..DGMM_MAX_D); list may be NULL. */
DGMM_API int dgmm_supported_dims(int* list, int* n);

/* Bytes of device workspace needed by dgmm_loglik_f32 / dgmm_estep_stats_f32 /
 * dgmm_estep_full_f32 for this shape on the current device (0 on error). */
DGMM_API size_t dgmm_workspace_bytes(int D, int K);

/* 1 when dgmm_estep_stats_f32 needs a non-NULL out_ll [N] for this shape: above D = 16 the statistics run as
 * two passes (cooperative 8-lane teams per pair, csrc/dgmm_team.cuh) and ll_i travels from the first to the
 * second through out_ll.  0 otherwise (out_ll optional). */
DGMM_API int dgmm_stats_needs_ll(int D, int K, int noise_layout);

/* ---- device entry points ------------------------------------------------------------ */

/* Per-point log-likelihood  ll_i = log sum_j alpha_j N(x_i | mu_j, Sigma_j + S_i).
 * Replaces DeconvGMM._e_step's density part (deconv/gmm/deconv_gmm.py:43-62,77),
 * SGDDeconvGMMModule.forward (deconv/gmm/sgd_deconv_gmm.py:14-28) and the per-batch body of
 * score_batch (deconv/gmm/online_deconv_gmm.py:190-205, deconv/gmm/sgd_gmm.py:228-242).
 *   log_alpha [K] = log mixture weights; out_ll [N]; out_sum (optional, 2 doubles on device:
 *   sum_i ll_i and the non-PD pair count); status (optional, 1 device int, zeroed here). */
DGMM_API int dgmm_loglik_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                    const float* log_alpha, const float* mu, const float* Sigma,
                    float* out_ll, double* out_sum, int* status,
                    void* workspace, size_t workspace_bytes, void* stream);

/* Fused E-step + M-step reductions: ll_i and the packed statistics above, in one pass.
 * Replaces deconv/gmm/deconv_gmm.py:39-107 (E-step + the reductions of the M-step),
 * deconv/gmm/batch_deconv_gmm.py:95-120, deconv/gmm/online_deconv_gmm.py:148-166 and the
 * autograd backward of deconv/gmm/sgd_deconv_gmm.py:14-28 (deconv/gmm/sgd_gmm.py:181).
 * K > 256 (the reference's Gaia runs use D = 7, K = 512, experiments/gmm/gaia/variable_k/fit_em.sh:18): D <= 8 serves
 * K <= 512 in one launch of 512-thread groups; otherwise the components are served in chunks of <= 256 (<= 512) - a log-likelihood pass per chunk, ll_i = logsumexp over the chunks, then a
 * statistics pass per chunk weighted with g_i exp(ll_chunk,i - ll_i), which makes omega_ij exact; a non-zero
 * resp_floor is scaled by the chunk's share of the point.
 *   g [N] optional per-point weights (NULL = 1; = grad_output for the SGD backward);
 *   resp_floor: 0 for EM/SGD, 10*eps for OnlineDeconvGMM (online_deconv_gmm.py:150);
 *   out_ll [N] optional (required where dgmm_stats_needs_ll says so: DGMM_ERR_LL_REQUIRED otherwise);
 *   out_stats [dgmm_stats_len(D,K)]. */
DGMM_API int dgmm_estep_stats_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                         const float* log_alpha, const float* mu, const float* Sigma,
                         const float* g, float resp_floor,
                         float* out_ll, float* out_stats,
                         void* workspace, size_t workspace_bytes, void* stream);

/* Materialising E-step for small N: log r [N,K], b [N,K,D], B [N,K,D,D], ll [N].
 * Replaces the return value of DeconvGMM._e_step (deconv/gmm/deconv_gmm.py:64-79) and
 * SGDDeconvGMM.posterior_params (deconv/gmm/sgd_deconv_gmm.py:71-94).
 * Any of out_log_r / out_b / out_B may be NULL to skip that output. */
DGMM_API int dgmm_estep_full_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                        const float* log_alpha, const float* mu, const float* Sigma,
                        float* out_log_r, float* out_b, float* out_B, float* out_ll, int* status,
                        void* workspace, size_t workspace_bytes, void* stream);

/* Closed-form EM parameter update from the (all-reduced) packed statistics, in place:
 *   delta = Sigma s1/s0; mu += delta; Sigma += Sigma (S2/s0) Sigma - delta delta^T + 2 w I/s0;
 *   alpha = s0/n_total.  Equals deconv/gmm/deconv_gmm.py:81-109 and
 * deconv/gmm/batch_deconv_gmm.py:65-75 (identity: SURVEY.md Appendix C).
 * alpha [K] and log_alpha [K] are both written. */
DGMM_API int dgmm_mstep_update_f32(const float* stats, double n_total, float w, int D, int K,
                          float* mu, float* Sigma, float* alpha, float* log_alpha, void* stream);

/* Online-EM M-step of one minibatch from the (all-reduced) packed statistics, in place
 * (OnlineDeconvGMM._m_step + _adjust, deconv/gmm/online_deconv_gmm.py:134-188; the statistics come from
 * dgmm_estep_stats_f32 with resp_floor = 10 eps, :150,154):
 *   running sums sum_resps [K] and sum_cond_means [K,D] are moved by `step_size` towards the batch's,
 *   mu = sum_cond_means / sum_resps, Sigma by the moment-shift rule of _adjust, alpha = sum_resps /
 *   global_batch; log_alpha is written for the next launch of the statistics kernel.
 * crash_flag (optional, 1 device int, caller-zeroed): a minibatch whose statistics carry a non-PD pair
 * leaves all state untouched and sets the flag, and so does every later call while it is set (the
 * reference breaks out of the epoch there, :79-88) -- no host synchronisation per minibatch. */
DGMM_API int dgmm_online_mstep_f32(const float* stats, float resp_floor, float step_size, float w,
                                   double global_batch, int D, int K, float* mu, float* Sigma, float* sum_resps,
                                   float* sum_cond_means, float* alpha, float* log_alpha, int* crash_flag,
                                   void* stream);

/* Register-only FFMA chain used to measure the sustained FP32 roofline denominator on the
 * box (MEASURED_PEAKS.json has no FP32 entry).  Launches one kernel on `stream`;
 * returns the number of FLOPs it executes through *flops.  sink = 1 device float. */
DGMM_API int dgmm_fp32_probe(float* sink, int iters, double* flops, void* stream);

/* ---- host entry points (HOST buffers; copies inside) -------------------------------- */

typedef struct dgmm_ctx dgmm_ctx;

/* Creates a context on `device` that owns pinned staging buffers, device chunk buffers and
 * two copy/compute streams sized for chunks of `chunk_points` points. */
DGMM_API int dgmm_ctx_create(dgmm_ctx** out, int device, int D, int K, int64_t chunk_points);
DGMM_API void dgmm_ctx_destroy(dgmm_ctx* ctx);

/* One EM iteration over HOST data (X [N,D], S in `noise_layout`), streamed to the device in
 * chunks with copy/compute overlap: E-step statistics over all chunks, then the M-step
 * update; mu/Sigma/alpha (HOST, in/out) are updated in place and *ll_sum receives
 * sum_i ll_i under the parameters the iteration started from (what BaseGMM.fit compares
 * against tol, deconv/gmm/base.py:46-55).  *nonpd receives the non-PD pair count.
 * Replaces one trip of the loop at deconv/gmm/base.py:46-55. */
DGMM_API int dgmm_em_iteration_host(dgmm_ctx* ctx, const float* X, const float* S, int noise_layout, int64_t N,
                           float w, float* alpha, float* mu, float* Sigma,
                           double* ll_sum, int64_t* nonpd);

/* E-step statistics over HOST data without the parameter update: out_stats (HOST,
 * dgmm_stats_len(D,K) floats) receives the packed buffer, ready for a cross-GPU all-reduce
 * followed by dgmm_mstep_update_f32.  alpha/mu/Sigma are HOST inputs.  This is the per-rank
 * half of dgmm_em_iteration_host for the sharded (one process per GPU) case and the body of
 * OnlineDeconvGMM's per-minibatch step (resp_floor = 10 eps, online_deconv_gmm.py:150). */
DGMM_API int dgmm_estep_stats_host(dgmm_ctx* ctx, const float* X, const float* S, int noise_layout, int64_t N,
                          const float* alpha, const float* mu, const float* Sigma, float resp_floor,
                          float* out_stats);

/* sum_i ll_i (and optionally ll_i [N] on the host) for HOST data.
 * Replaces score_batch (deconv/gmm/online_deconv_gmm.py:190-205, deconv/gmm/sgd_gmm.py:228-242). */
DGMM_API int dgmm_score_host(dgmm_ctx* ctx, const float* X, const float* S, int noise_layout, int64_t N,
                    const float* alpha, const float* mu, const float* Sigma,
                    float* out_ll, double* ll_sum, int64_t* nonpd);

/* ---- resident table: upload once, iterate many times --------------------------------------
 * The per-call entries above re-send the whole table every EM iteration (the reference's contract: fit() is
 * handed host tensors).  A fit of E iterations should pay the PCIe transfer once: dgmm_ctx_upload copies the
 * table into device memory owned by the context (chunked, from pageable or pinned host memory) and the
 * _resident entries iterate on that copy.  With pack_full != 0 a DGMM_NOISE_FULL table is packed on the
 * device while it streams in (kept as DGMM_NOISE_PACKED: 1.7-1.9x less HBM).  One table per context; a second
 * upload replaces it.  Replaces the loop of deconv/gmm/base.py:46-55 around data that stays on the GPU. */
DGMM_API int dgmm_ctx_upload(dgmm_ctx* ctx, const float* X, const float* S, int noise_layout, int64_t N,
                             int pack_full);
/* One EM iteration on the resident table; alpha/mu/Sigma are HOST in/out exactly as in dgmm_em_iteration_host. */
DGMM_API int dgmm_em_iteration_resident(dgmm_ctx* ctx, float w, float* alpha, float* mu, float* Sigma,
                                        double* ll_sum, int64_t* nonpd);
/* E iterations of base.py:46-55 on the resident table without a host round trip of the parameters in between:
 * stops when |ll - ll_prev| < tol or a non-PD pair appears (base.py:48-53); *iters_done and ll_trace[0..E)
 * (optional) report what ran.  Parameters are uploaded once and read back once. */
DGMM_API int dgmm_em_fit_resident(dgmm_ctx* ctx, int max_iters, float w, double tol, float* alpha, float* mu,
                                  float* Sigma, int* iters_done, double* ll_trace, int64_t* nonpd);
/* sum_i ll_i (and optionally ll_i [N] on the host) on the resident table. */
DGMM_API int dgmm_score_resident(dgmm_ctx* ctx, const float* alpha, const float* mu, const float* Sigma,
                                 float* out_ll, double* ll_sum, int64_t* nonpd);

/* ---- noise layout converters ---------------------------------------------------------------- */
/* S_full [N,D,D] -> S_packed [N,D(D+1)/2] (lower triangles, row-major).  Device pointers / host pointers. */
DGMM_API int dgmm_pack_noise_f32(const float* S_full, int64_t N, int D, float* S_packed, void* stream);
DGMM_API int dgmm_pack_noise_host(const float* S_full, int64_t N, int D, float* S_packed);

/* ---- initialisation: fused k-means assignment (device pointers) ----------------------- */

/* Packed buffer written by dgmm_kmeans_assign_f32:
 *   [ counts (K) | sums (K*D) | dist_hi | dist_lo ]
 * counts_j = #{i : label_i = j}, sums_j = sum of the x_i assigned to j, dist_hi + dist_lo =
 * sum_i min_j |x_i - c_j| (Euclidean norm; a float pair carrying a double).  Every slot is additive
 * across shards: one all-reduce(SUM) of K*(1+D)+2 floats per k-means pass. */
static inline size_t dgmm_kmeans_stats_len(int D, int K) {
    return (size_t)K * (size_t)(1 + D) + 2;
}

/* Bytes of device workspace needed by dgmm_kmeans_assign_f32 (0 on error). */
DGMM_API size_t dgmm_kmeans_workspace_bytes(int D, int K);

/* Nearest-centroid assignment of X [N,D] against centroids [K,D] and, in the same pass, the
 * per-centroid counts and coordinate sums (ties -> lowest index, as torch.min).  Replaces the
 * [B,K,D] difference tensor, argmin, bincount and mask reduction of
 * deconv/gmm/util.py:19-24 (k_means) and deconv/gmm/util.py:51-64 (minibatch_k_means).
 *   labels (optional) [N] int32; out_stats [dgmm_kmeans_stats_len(D,K)].
 * Deterministic (no atomics; fixed summation order for a given N and device). */
DGMM_API int dgmm_kmeans_assign_f32(const float* X, int64_t N, int D, int K, const float* centroids,
                                    int32_t* labels, float* out_stats, void* workspace, size_t workspace_bytes,
                                    void* stream);

/* dgmm_kmeans_assign_f32 for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc): same inputs, outputs and results, but the per-block partial sums live in a file-scope
 * `static __device__` array of the library (dgmm_kmeans_static_scratch_bytes() = 8 MiB, held by every GPU that loads the
 * library until the process exits) instead of a caller-provided workspace; the call makes no cudaMalloc / cudaFree /
 * allocating library call.  Limit: shapes with dgmm_kmeans_workspace_bytes(D, K) > 8 MiB (on a 148-SM B200:
 * K * ceil((D + 1) / 4) > 442, e.g. K > 147 at D = 10, K > 88 at D = 16) return DGMM_ERR_STATIC_SCRATCH.  All calls on a
 * device share the array: calls that overlap on the GPU (different streams) race; calls on one stream are safe.
 * No counterpart in the reference. */
DGMM_API size_t dgmm_kmeans_static_scratch_bytes(void);
DGMM_API int dgmm_kmeans_assign_static_f32(const float* X, int64_t N, int D, int K, const float* centroids,
                                           int32_t* labels, float* out_stats, void* stream);

/* Centroid update from the (all-reduced) packed buffer, in place.
 *   mode 0, Lloyd (deconv/gmm/util.py:23-24): c_j = sums_j / counts_j; an empty cluster keeps its
 *           centroid (the reference writes NaN there: documented fix).
 *   mode 1, minibatch (deconv/gmm/util.py:56-64): counts_run_j += counts_j;
 *           c_j += (sums_j - counts_j c_j) / counts_run_j   (per-centre rate 1/count).
 * counts_run [K] is read and written for mode 1 only. */
DGMM_API int dgmm_kmeans_update_f32(int mode, const float* stats, int D, int K, float* centroids,
                                    float* counts_run, void* stream);

/* ---- device-resident minibatches --------------------------------------------------------- */

/* Gathers the rows idx[0..B) (int64, device) of X [N,D] and of the noise table (DGMM_NOISE_FULL: [N,D,D],
 * DGMM_NOISE_PACKED: [N,D(D+1)/2], DGMM_NOISE_DIAG: [N,D], DGMM_NOISE_NONE: nothing) into contiguous outX [B,D] / outS [B,...] in one launch.
 * Replaces the DataLoader's per-sample __getitem__ + collate (deconv/gmm/data.py:7-17,
 * deconv/gmm/sgd_gmm.py:139-145, deconv/gmm/online_deconv_gmm.py:43-56) for tables resident on the GPU. */
DGMM_API int dgmm_gather_points_f32(const float* X, const float* S, int noise_layout, const int64_t* idx, int64_t B,
                                    int D, float* outX, float* outS, void* stream);

/* ---- posterior sampling (device pointers) ------------------------------------------------ */

/* M draws per point from the posterior mixture p(v | w_i, S_i) = sum_j r_ij N(b_ij, B_ij) without
 * materialising b [N,K,D] or B [N,K,D,D].  Replaces SGDDeconvGMM._sample_posterior
 * (deconv/gmm/sgd_deconv_gmm.py:107-119; the exact-baseline sampler of the flow experiments,
 * deconv/flow/svi_gmm.py:21-32).  The caller supplies the randomness, so the call is a deterministic
 * map: u [N,M] uniforms in [0,1) pick the component by inverse CDF of r_i (first j with
 * cumsum_j r_ij > u), z [N,M,D] standard normals give the draw b_ij + chol(B_ij) z, which is how
 * torch.distributions.MultivariateNormal(loc, covariance_matrix).sample maps its normals.
 *   out_samples [N,M,D]; out_comp (optional) [N,M] int32 picked components; scratch_log_r [N,K]
 *   (receives log r_ij); out_ll [N] (receives ll_i); status (optional, 1 device int): non-PD
 *   (point,component) pairs plus draws whose B_ij had a non-positive pivot (those samples are NaN). */
DGMM_API int dgmm_posterior_sample_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                                       const float* log_alpha, const float* mu, const float* Sigma, int M,
                                       const float* u, const float* z, float* out_samples, int32_t* out_comp,
                                       float* scratch_log_r, float* out_ll, int* status, void* workspace,
                                       size_t workspace_bytes, void* stream);

/* Log-density of query points V [N,M,D] under each point's posterior mixture,
 *   out_logprob[i,m] = logsumexp_j ( log r_ij + log N(V[i,m] | b_ij, B_ij) ),
 * without materialising b or B.  Replaces SGDDeconvGMM.posterior_log_prob
 * (deconv/gmm/sgd_deconv_gmm.py:96-105).  scratch_log_r [N,K], out_ll [N], status as for
 * dgmm_posterior_sample_f32. */
DGMM_API int dgmm_posterior_logprob_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                                        const float* log_alpha, const float* mu, const float* Sigma, int M,
                                        const float* V, float* out_logprob, float* scratch_log_r, float* out_ll,
                                        int* status, void* workspace, size_t workspace_bytes, void* stream);

/* ---- minibatch SGD: parameter side of one step (device pointers) ------------------------ */

/* Parameter transform of SGDGMMModule (deconv/gmm/sgd_gmm.py:38-47, sgd_deconv_gmm.py:17):
 *   log_alpha = log_softmax(soft_weights) [K],
 *   Sigma_j = L_j L_j^T with L_j = strict_lower(l_lower_j) + diag(exp(l_diag_j))   [K,D,D].
 * l_lower [K, D(D-1)/2] in the order of torch.tril_indices(D, D, -1) (may be NULL for D = 1). */
DGMM_API int dgmm_sgd_params_f32(const float* soft_weights, const float* l_diag, const float* l_lower, int D, int K,
                                 float* log_alpha, float* Sigma, void* stream);

/* One optimiser step of BaseSGDGMM.fit (deconv/gmm/sgd_gmm.py:167-182) from the (all-reduced) packed
 * statistics of the minibatch (dgmm_estep_stats_f32 with g = NULL, resp_floor = 0):
 *   loss = -(1/n_batch) sum_i ll_i + (n_batch/n_total) w sum_{j,d} 1/Sigma_j[d,d]   (reg_loss, :110-112)
 * chain rule to (soft_weights, means, l_diag, l_lower), torch.optim.Adam update (no weight decay, no
 * amsgrad; `step` = 1-based step count for the bias corrections), then the transform of
 * dgmm_sgd_params_f32 on the updated parameters.  All four parameter tensors, their Adam moments
 * exp_avg[4] / exp_avg_sq[4] (HOST arrays of device pointers in the order soft_weights, means, l_diag,
 * l_lower) are updated in place; log_alpha [K] and Sigma [K,D,D] are written for the next launch of
 * the statistics kernel (mu is `means` itself).  epoch_acc (optional, 2 device doubles) accumulates
 * sum_i ll_i and the non-PD pair count of the minibatches (the per-epoch train loss of :176 without
 * a host sync per batch). */
DGMM_API int dgmm_sgd_adam_step_f32(const float* stats, double n_batch, double n_total, float w, float lr,
                                    float beta1, float beta2, float eps, int step, int D, int K,
                                    float* soft_weights, float* means, float* l_diag, float* l_lower,
                                    float* const* exp_avg, float* const* exp_avg_sq, float* log_alpha,
                                    float* Sigma, double* epoch_acc, void* stream);

/* One minibatch of BaseSGDGMM.fit (deconv/gmm/sgd_gmm.py:167-182) in ONE call: gather of the shuffled batch from
 * the resident table (the DataLoader's __getitem__ + collate, :139-145,167-169), fused statistics pass (forward +
 * backward, :173-181) and the Adam step of dgmm_sgd_adam_step_f32 (:182).
 *   X / S: the resident table (layout as noise_layout); perm: device int64 permutation of its rows, the batch is
 *   perm[batch_lo .. batch_lo + B) (perm == NULL: the contiguous rows batch_lo .. batch_lo + B);
 *   bufX [B,D], bufS [B,...]: gather outputs (unused without perm); ll_scratch [B] or NULL (needed when
 *   dgmm_stats_needs_ll); phase: 1 = gather + statistics into `stats` (then the caller all-reduces it),
 *   2 = Adam step from `stats`, 3 = both (single process).  Remaining arguments as dgmm_sgd_adam_step_f32. */
DGMM_API int dgmm_sgd_fused_step_f32(const float* X, const float* S, int noise_layout, int D, int K,
                                     const int64_t* perm, int64_t batch_lo, int64_t B, float* bufX, float* bufS,
                                     float* ll_scratch, int phase, float* stats, double n_batch, double n_total,
                                     float w, float lr, float beta1, float beta2, float eps, int step,
                                     float* soft_weights, float* means, float* l_diag, float* l_lower,
                                     float* const* exp_avg, float* const* exp_avg_sq, float* log_alpha,
                                     float* Sigma, double* epoch_acc, void* workspace, size_t workspace_bytes,
                                     void* stream);

/* The Adam step of dgmm_sgd_adam_step_f32 with the cross-GPU SUM of the packed statistics FUSED into its kernels
 * (no collective call between the statistics pass and the update; replaces the all-reduce of SURVEY.md 8e for the
 * minibatch-SGD step, where a separate 140 KB NCCL all-reduce costs ~10 % of a 1.3 ms step).  One process per GPU
 * of one NVLink box (world <= 8).  Every rank owns a SYMMETRIC buffer mapped by all peers; peer_bufs[r] (HOST array
 * of device pointers) is rank r's buffer as mapped in this process.  Layout in floats: this step's packed statistics
 * at stats_off (callers alternate two regions with the step parity), `world` uint32 hand-shake flags at flags_off.
 * The local statistics must already be in peer_bufs[rank] + stats_off (same stream).  The call stamps flag[rank] = stamp
 * in every peer's buffer (release, system scope); the Adam kernels wait (bounded, ~20 s) for all `world` local flags to
 * reach `stamp` (acquire), then read each statistic from all peers and add them in rank order - bit-identical on
 * every rank.  stamp must grow by one per step.  timeout_flag (device int, zeroed by the caller) is set when a peer
 * never answered; epoch_acc[1] is then poisoned like a non-PD pair. */
DGMM_API int dgmm_sgd_adam_step_peers_f32(void* const* peer_bufs, int world, int rank, int64_t stats_off,
                                          int64_t flags_off, uint32_t stamp, int* timeout_flag, double n_batch,
                                          double n_total, float w, float lr, float beta1, float beta2, float eps,
                                          int step, int D, int K, float* soft_weights, float* means, float* l_diag,
                                          float* l_lower, float* const* exp_avg, float* const* exp_avg_sq,
                                          float* log_alpha, float* Sigma, double* epoch_acc, void* stream);

/* ---- diagnostics (not part of the drop-in surface) ----------------------------------- */

/* Overrides the automatic launch plan of the pair kernel; -1 = automatic for that field.
 * sigp/accp: 0 registers, 1 shared memory, 2 global; rc: register-cap class. */
DGMM_API int dgmm_debug_set_tuning(int sigp, int accp, int nslot, int pt, int nst, int blocks_per_sm, int rc,
                                   int no_bulk);
/* out[0..10] = sigp, accp, rc, GS, NSLOT, TPB, PT, NST, grid, smem bytes, use_bulk. */
DGMM_API int dgmm_debug_get_plan(int mode, int D, int K, int64_t N, int noise_layout, int has_g, int* out);

/* Per-launch device timing of the pair kernel (CUDA events on the launching stream, recorded
 * directly around the launch).  dgmm_debug_pair_timing(1) starts a fresh record;
 * dgmm_debug_pair_times returns the last min(*n, 512) launch times in ms (*n: in = capacity,
 * out = count) after synchronising on them.  Used by bench.py's roofline line. */
DGMM_API int dgmm_debug_pair_timing(int enable);
DGMM_API int dgmm_debug_pair_times(float* out_ms, int* n);

#ifdef __cplusplus
}
#endif
#endif /* DGMM_B200_H */
