// dgmm_sgd.cu -- the parameter side of one minibatch-SGD step as two small kernels, so that a
// step of BaseSGDGMM.fit (deconv/gmm/sgd_gmm.py:167-182) is five launches of this library and no
// autograd graph:
//     prepare_kernel, pair_kernel<MODE_STATS>, finalize_stats_kernel      (forward + backward)
//     [all-reduce of the packed statistics]
//     sgd_component_kernel, sgd_weights_kernel                             (chain rule + Adam + re-parametrise)
// What the two kernels replace, per step, in the reference:
//   * autograd through SGDGMMModule.L / covars (sgd_gmm.py:38-47), softmax (sgd_deconv_gmm.py:17) and
//     reg_loss (sgd_gmm.py:110-112): with the fused statistics s0, s1, S2 of sum_i ll_i
//       d loss/d mu_j      = -s1_j / n
//       d loss/d Sigma_j   = -S2_j / (2n)  -  (n/n_total) w diag(1 / Sigma_j[d,d]^2)      =: G_j (symmetric)
//       d loss/d L_j       = 2 G_j L_j;   l_lower: its strict lower part, l_diag: diag * exp(l_diag)
//       d loss/d soft_w_k  = -(s0_k - alpha_k sum_j s0_j) / n
//   * torch.optim.Adam.step (defaults: betas (0.9, 0.999), eps 1e-8, no weight decay / amsgrad), state in
//     caller-owned exp_avg / exp_avg_sq buffers (the estimator passes the torch optimiser's own state
//     tensors, so `optimiser.state` stays truthful);
//   * the next forward's parameter transform: log_alpha = log_softmax(soft_weights), Sigma_j = L_j L_j^T.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/dgmm_b200.h"

namespace {

#define CUDA_TRY(x)                                  \
    do {                                             \
        cudaError_t e__ = (x);                       \
        if (e__ != cudaSuccess) return (int)e__;     \
    } while (0)

struct AdamCfg {
    float lr, beta1, beta2, eps;
    float bc1, bc2_sqrt;  // 1 - beta1^t, sqrt(1 - beta2^t)
};

// Cross-GPU reduction fused into the consumer (round 2).  Under one process per GPU every rank's packed statistics
// sit in a SYMMETRIC buffer that all peers map over NVLink (torch.distributed._symmetric_memory is only the
// allocator / handle exchange).  Instead of an NCCL all-reduce between the statistics pass and the Adam kernels
// (a separate collective kernel on another stream, ~100 us per step at 140 KB), the two Adam kernels read the
// statistic they need from ALL peers' buffers and add the `world` values themselves, in rank order (bit-identical
// on every rank).  Hand-shake: after its finalise kernel a rank stamps flag[rank] in every peer's buffer
// (st.release.sys after a system fence); a consumer block spins (bounded) until all `world` local flags carry the
// step's stamp (ld.acquire.sys, bounded at ~20 s) before its first peer read.  Buffers alternate with the step parity, so a fast rank
// can never overwrite statistics a slow one is still reading (it cannot get two steps ahead: its next hand-shake
// needs the slow rank's stamp).
constexpr int kMaxPeers = 8;
struct PeerView {
    const float* stats[kMaxPeers];   // this step's packed statistics of every rank (world == 1: the local buffer)
    const unsigned* flags;           // local flags [world] (nullptr: no hand-shake)
    int world;
    unsigned stamp;
    int* timeout;                    // set to 1 when the bounded wait expired (peer died): results are then invalid
};
struct PeerFlags {
    unsigned* flags[kMaxPeers];      // flag arrays of every rank (peer-mapped)
};

// All `world` remote loads are issued before the first add (a rank-by-rank loop would pay the NVLink round trip
// eight times in a row); the sum itself is taken in rank order.
__device__ __forceinline__ float peer_sum(const PeerView& pv, size_t e) {
    float v[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r) v[r] = (r < pv.world) ? __ldcg(pv.stats[r] + e) : 0.f;
    float s = v[0];
#pragma unroll
    for (int r = 1; r < kMaxPeers; ++r) s += v[r];
    return s;
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Block-wide: wait until every rank has published this step's statistics.
__device__ __forceinline__ void peer_wait(const PeerView& pv) {
    if (pv.flags == nullptr) return;
    if ((int)threadIdx.x < pv.world) {
        const long long t0 = clock64();
        while ((int)(ld_acquire_sys(pv.flags + threadIdx.x) - pv.stamp) < 0) {
            if (clock64() - t0 > 40000000000LL) {   // ~20 s at 1.9 GHz: a peer is gone; do not hang the GPU
                if (pv.timeout) *pv.timeout = 1;
                break;
            }
        }
    }
    __syncthreads();
}
__global__ void peer_signal_kernel(PeerFlags pf, int world, int rank, unsigned stamp) {
    if ((int)threadIdx.x < world) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pf.flags[threadIdx.x] + rank), "r"(stamp) : "memory");
    }
}

// torch/optim/adam.py (_single_tensor_adam): exp_avg.lerp_(grad, 1 - beta1);
// exp_avg_sq = beta2 exp_avg_sq + (1 - beta2) grad^2; denom = sqrt(exp_avg_sq) / sqrt(bc2) + eps;
// param -= (lr / bc1) * exp_avg / denom
__device__ __forceinline__ float adam_update(float p, float g, float* m, float* v, const AdamCfg& c) {
    const float mm = *m + (g - *m) * (1.f - c.beta1);
    const float vv = c.beta2 * *v + (1.f - c.beta2) * g * g;
    *m = mm;
    *v = vv;
    const float denom = sqrtf(vv) / c.bc2_sqrt + c.eps;
    return p - (c.lr / c.bc1) * (mm / denom);
}

__device__ __forceinline__ int low_index(int a, int b) { return a * (a - 1) / 2 + b; }  // torch.tril_indices(D, D, -1) order

// One block per component, D*D threads.  update == 0: only the transform Sigma_j = L_j L_j^T.
__global__ void sgd_component_kernel(int update, const PeerView pv, float inv_n, float reg, AdamCfg cfg,
                                     int D, int K, float* __restrict__ means, float* __restrict__ l_diag,
                                     float* __restrict__ l_lower, float* m_mu, float* v_mu, float* m_ld, float* v_ld,
                                     float* m_ll, float* v_ll, float* __restrict__ Sigma) {
    extern __shared__ float sh[];
    float* L = sh;            // [D*D]
    float* G = L + D * D;     // [D*D]
    const int j = blockIdx.x, t = threadIdx.x;
    const int a = t / D, b = t - a * D;
    const int NL = D * (D - 1) / 2;
    float* ld = l_diag + (size_t)j * D;
    float* ll = l_lower + (size_t)j * NL;
    L[t] = (a == b) ? expf(ld[a]) : (a > b ? ll[low_index(a, b)] : 0.f);
    __syncthreads();
    if (update) {
        peer_wait(pv);
        const size_t s1 = (size_t)K + (size_t)j * D;                              // offsets in the packed buffer
        const size_t S2 = (size_t)K + (size_t)K * D + (size_t)j * D * D;
        const float s1_a = (b == 0) ? peer_sum(pv, s1 + a) : 0.f;                 // (both remote reads up front)
        float g = -0.5f * inv_n * peer_sum(pv, S2 + t);
        if (a == b) {
            float sd = 0.f;
            for (int c = 0; c <= a; ++c) sd = fmaf(L[a * D + c], L[a * D + c], sd);   // Sigma_j[a,a]
            g -= reg / (sd * sd);
        }
        G[t] = g;
        __syncthreads();
        float newL = L[t];
        if (a >= b) {
            float dl = 0.f;
            for (int c = b; c < D; ++c) dl = fmaf(G[a * D + c], L[c * D + b], dl);    // L is lower: L[c][b] = 0 for c < b
            dl *= 2.f;
            if (a == b) {
                const int i = j * D + a;
                const float nd = adam_update(ld[a], dl * L[t], m_ld + i, v_ld + i, cfg);
                ld[a] = nd;
                newL = expf(nd);
            } else {
                const int i = j * NL + low_index(a, b);
                newL = adam_update(ll[low_index(a, b)], dl, m_ll + i, v_ll + i, cfg);
                ll[low_index(a, b)] = newL;
            }
        }
        if (b == 0) {
            const int i = j * D + a;
            means[i] = adam_update(means[i], -inv_n * s1_a, m_mu + i, v_mu + i, cfg);
        }
        __syncthreads();
        L[t] = newL;
        __syncthreads();
    }
    const int lo = a < b ? a : b;
    float s = 0.f;
    for (int c = 0; c <= lo; ++c) s = fmaf(L[a * D + c], L[b * D + c], s);
    Sigma[(size_t)j * D * D + t] = s;
}

// One block of roundup(K) threads: 256, 512 or 1024 (K <= 1024).  update == 0: only log_alpha = log_softmax(soft_weights).
// epoch_acc (optional, 2 doubles): += (sum_i ll_i of the minibatch, non-PD pair count) from the stats tail.
__global__ void sgd_weights_kernel(int update, const PeerView pv, float inv_n, AdamCfg cfg, int D, int K,
                                   float* __restrict__ soft_weights, float* m_sw, float* v_sw,
                                   float* __restrict__ log_alpha, double* epoch_acc) {
    __shared__ float red[1024];
    __shared__ float red2[1024];
    const int t = threadIdx.x;
    const int half = blockDim.x >> 1;
    auto block_max = [&](float v) {
        red[t] = v;
        __syncthreads();
        for (int o = half; o > 0; o >>= 1) {
            if (t < o) red[t] = fmaxf(red[t], red[t + o]);
            __syncthreads();
        }
        const float r = red[0];
        __syncthreads();
        return r;
    };
    auto block_sum2 = [&](float v, float w, float& sv, float& sw) {
        red[t] = v;
        red2[t] = w;
        __syncthreads();
        for (int o = half; o > 0; o >>= 1) {
            if (t < o) { red[t] += red[t + o]; red2[t] += red2[t + o]; }
            __syncthreads();
        }
        sv = red[0];
        sw = red2[0];
        __syncthreads();
    };
    float w = t < K ? soft_weights[t] : -INFINITY;
    if (update) {
        peer_wait(pv);
        const float s0 = t < K ? peer_sum(pv, t) : 0.f;
        // tails of the ranks' buffers: lane r reads rank r's (sum ll, non-PD count) in parallel
        __shared__ double tail_ll[kMaxPeers], tail_bad[kMaxPeers];
        if (epoch_acc && t < pv.world) {
            const float* tl = pv.stats[t] + (size_t)K * (1 + D + D * D);
            tail_ll[t] = (double)__ldcg(tl) + (double)__ldcg(tl + 1);
            tail_bad[t] = (double)__ldcg(tl + 3);
        }
        const float mx = block_max(w);
        const float e = t < K ? expf(w - mx) : 0.f;
        float se, ss;
        block_sum2(e, s0, se, ss);
        if (t < K) {
            const float alpha = e / se;
            const float g = -inv_n * (s0 - alpha * ss);
            w = adam_update(w, g, m_sw + t, v_sw + t, cfg);
            soft_weights[t] = w;
        }
        if (t == 0 && epoch_acc) {
            double ll = 0.0, bad = 0.0;
            for (int r = 0; r < pv.world; ++r) { ll += tail_ll[r]; bad += tail_bad[r]; }
            if (pv.timeout && *pv.timeout) bad += 1.0;     // a peer never answered: poison the epoch like a non-PD pair
            epoch_acc[0] += ll;
            epoch_acc[1] += bad;
        }
    }
    const float mx = block_max(w);
    const float e = t < K ? expf(w - mx) : 0.f;
    float se, dummy;
    block_sum2(e, 0.f, se, dummy);
    if (t < K) log_alpha[t] = (w - mx) - logf(se);
}

inline PeerView local_view(const float* stats) {
    PeerView pv{};
    pv.stats[0] = stats;
    pv.world = 1;
    return pv;
}

inline int weights_block(int K) { return K <= 256 ? 256 : (K <= 512 ? 512 : 1024); }

int check_shape(int D, int K) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1 || K > DGMM_MAX_K) return DGMM_ERR_UNSUPPORTED_K;
    return 0;
}

}  // namespace

extern "C" {

int dgmm_sgd_params_f32(const float* soft_weights, const float* l_diag, const float* l_lower, int D, int K,
                        float* log_alpha, float* Sigma, void* stream) {
    int rc = check_shape(D, K);
    if (rc) return rc;
    if (!soft_weights || !l_diag || (D > 1 && !l_lower) || !log_alpha || !Sigma) return DGMM_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    AdamCfg cfg{};
    sgd_component_kernel<<<K, D * D, 2 * D * D * sizeof(float), st>>>(
        0, local_view(nullptr), 0.f, 0.f, cfg, D, K, nullptr, const_cast<float*>(l_diag), const_cast<float*>(l_lower), nullptr,
        nullptr, nullptr, nullptr, nullptr, nullptr, Sigma);
    CUDA_TRY(cudaGetLastError());
    sgd_weights_kernel<<<1, weights_block(K), 0, st>>>(0, local_view(nullptr), 0.f, cfg, D, K, const_cast<float*>(soft_weights), nullptr, nullptr,
                                          log_alpha, nullptr);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static int adam_launch(const PeerView& pv, double n_batch, double n_total, float w, float lr, float beta1, float beta2,
                       float eps, int step, int D, int K, float* soft_weights, float* means, float* l_diag,
                       float* l_lower, float* const* exp_avg, float* const* exp_avg_sq, float* log_alpha, float* Sigma,
                       double* epoch_acc, cudaStream_t st) {
    int rc = check_shape(D, K);
    if (rc) return rc;
    if (!soft_weights || !means || !l_diag || (D > 1 && !l_lower) || !exp_avg || !exp_avg_sq || !log_alpha ||
        !Sigma || step < 1 || !(n_batch > 0) || !(n_total > 0))
        return DGMM_ERR_BAD_ARG;
    for (int i = 0; i < 4; ++i)
        if ((i != 3 || D > 1) && (!exp_avg[i] || !exp_avg_sq[i])) return DGMM_ERR_BAD_ARG;
    AdamCfg cfg;
    cfg.lr = lr; cfg.beta1 = beta1; cfg.beta2 = beta2; cfg.eps = eps;
    cfg.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
    cfg.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    const float inv_n = (float)(1.0 / n_batch);
    const float reg = (float)(n_batch / n_total * (double)w);
    // exp_avg / exp_avg_sq: [0] soft_weights, [1] means, [2] l_diag, [3] l_lower
    sgd_component_kernel<<<K, D * D, 2 * D * D * sizeof(float), st>>>(1, pv, inv_n, reg, cfg, D, K, means, l_diag,
                                                                      l_lower, exp_avg[1], exp_avg_sq[1], exp_avg[2],
                                                                      exp_avg_sq[2], exp_avg[3], exp_avg_sq[3], Sigma);
    CUDA_TRY(cudaGetLastError());
    sgd_weights_kernel<<<1, weights_block(K), 0, st>>>(1, pv, inv_n, cfg, D, K, soft_weights, exp_avg[0], exp_avg_sq[0],
                                                       log_alpha, epoch_acc);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int dgmm_sgd_adam_step_f32(const float* stats, double n_batch, double n_total, float w, float lr, float beta1,
                           float beta2, float eps, int step, int D, int K, float* soft_weights, float* means,
                           float* l_diag, float* l_lower, float* const* exp_avg, float* const* exp_avg_sq,
                           float* log_alpha, float* Sigma, double* epoch_acc, void* stream) {
    if (!stats) return DGMM_ERR_BAD_ARG;
    return adam_launch(local_view(stats), n_batch, n_total, w, lr, beta1, beta2, eps, step, D, K, soft_weights, means,
                       l_diag, l_lower, exp_avg, exp_avg_sq, log_alpha, Sigma, epoch_acc, (cudaStream_t)stream);
}

// The same step with the cross-GPU SUM of the statistics fused into the Adam kernels (no collective call):
// peer_bufs[r] = rank r's symmetric buffer as mapped in THIS process; this step's statistics of rank r sit at
// peer_bufs[r] + stats_off (floats), the hand-shake flags [world] of rank r at peer_bufs[r] + flags_off.  The local
// statistics must already be in peer_bufs[rank] + stats_off (written by the statistics pass on the same stream).
// `stamp` must grow by one per step; `timeout_flag` (device int, zero-initialised) reports a peer that never answered.
int dgmm_sgd_adam_step_peers_f32(void* const* peer_bufs, int world, int rank, int64_t stats_off, int64_t flags_off,
                                 uint32_t stamp, int* timeout_flag, double n_batch, double n_total, float w, float lr,
                                 float beta1, float beta2, float eps, int step, int D, int K, float* soft_weights,
                                 float* means, float* l_diag, float* l_lower, float* const* exp_avg,
                                 float* const* exp_avg_sq, float* log_alpha, float* Sigma, double* epoch_acc,
                                 void* stream) {
    if (!peer_bufs || world < 1 || world > kMaxPeers || rank < 0 || rank >= world || stats_off < 0 || flags_off < 0)
        return DGMM_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    PeerView pv{};
    PeerFlags pf{};
    for (int r = 0; r < world; ++r) {
        if (!peer_bufs[r]) return DGMM_ERR_BAD_ARG;
        pv.stats[r] = (const float*)peer_bufs[r] + stats_off;
        pf.flags[r] = (unsigned*)((float*)peer_bufs[r] + flags_off);
    }
    pv.flags = pf.flags[rank];
    pv.world = world;
    pv.stamp = stamp;
    pv.timeout = timeout_flag;
    peer_signal_kernel<<<1, 32, 0, st>>>(pf, world, rank, stamp);
    CUDA_TRY(cudaGetLastError());
    return adam_launch(pv, n_batch, n_total, w, lr, beta1, beta2, eps, step, D, K, soft_weights, means, l_diag, l_lower,
                       exp_avg, exp_avg_sq, log_alpha, Sigma, epoch_acc, st);
}

// One minibatch step of BaseSGDGMM.fit as ONE call (the host side of a step is otherwise three ctypes round trips
// plus their Python argument checks, which at D = 10, K = 64 cost as much as the kernels):
//   phase & 1: gather the rows perm[batch_lo .. batch_lo + B) of the resident table into bufX / bufS (perm == NULL:
//              the batch is the contiguous slice starting at batch_lo), then the fused statistics pass -> stats
//   phase & 2: chain rule + Adam + re-parametrisation from stats (dgmm_sgd_adam_step_f32)
// Single process: phase = 3.  Under torch.distributed: phase 1, all-reduce of `stats` by the caller, phase 2.
int dgmm_sgd_fused_step_f32(const float* X, const float* S, int noise_layout, int D, int K, const int64_t* perm,
                            int64_t batch_lo, int64_t B, float* bufX, float* bufS, float* ll_scratch, int phase,
                            float* stats, double n_batch, double n_total, float w, float lr, float beta1, float beta2,
                            float eps, int step, float* soft_weights, float* means, float* l_diag, float* l_lower,
                            float* const* exp_avg, float* const* exp_avg_sq, float* log_alpha, float* Sigma,
                            double* epoch_acc, void* workspace, size_t workspace_bytes, void* stream) {
    int rc = check_shape(D, K);
    if (rc) return rc;
    if (B < 1 || batch_lo < 0 || !stats) return DGMM_ERR_BAD_ARG;
    if (phase & 1) {
        const bool per_point = noise_layout == DGMM_NOISE_FULL || noise_layout == DGMM_NOISE_DIAG ||
                               noise_layout == DGMM_NOISE_PACKED;
        const size_t n_s = noise_layout == DGMM_NOISE_FULL ? (size_t)D * D
                           : (noise_layout == DGMM_NOISE_DIAG ? (size_t)D
                              : (noise_layout == DGMM_NOISE_PACKED ? (size_t)D * (D + 1) / 2 : 0));
        const float* bx;
        const float* bs;
        if (perm) {
            if (!bufX || (per_point && !bufS)) return DGMM_ERR_BAD_ARG;
            rc = dgmm_gather_points_f32(X, per_point ? S : nullptr, per_point ? noise_layout : DGMM_NOISE_NONE,
                                        perm + batch_lo, B, D, bufX, bufS, stream);
            if (rc) return rc;
            bx = bufX;
            bs = per_point ? bufS : S;
        } else {
            bx = X + (size_t)batch_lo * D;
            bs = per_point ? S + (size_t)batch_lo * n_s : S;
        }
        rc = dgmm_estep_stats_f32(bx, bs, noise_layout, B, D, K, log_alpha, means, Sigma, nullptr, 0.f, ll_scratch, stats,
                                  workspace, workspace_bytes, stream);
        if (rc) return rc;
    }
    if (phase & 2) {
        rc = dgmm_sgd_adam_step_f32(stats, n_batch, n_total, w, lr, beta1, beta2, eps, step, D, K, soft_weights, means,
                                    l_diag, l_lower, exp_avg, exp_avg_sq, log_alpha, Sigma, epoch_acc, stream);
        if (rc) return rc;
    }
    return 0;
}

}  // extern "C"
