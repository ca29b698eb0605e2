// dgmm_api.cu -- C ABI (include/dgmm_b200.h): argument checks, launch planning, the small
// parameter-side kernels (prepare / finalise / M-step update / FP32 probe) and the
// host-buffer streaming pipeline.  The hot kernels live in dgmm_pair*.cuh / dgmm_full*.cuh
// and are reached through the per-D lookup functions of dgmm_tables.cuh.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

#include "../../include/dgmm_b200.h"
#include "dgmm_tables.cuh"
#include "dgmm_team_geo.cuh"

// per-D lookup functions, one translation unit each (dgmm_inst_dXX.cu)
#define DGMM_DECL(D)                                                          \
    dgmm::PairVariant dgmm_pair_lookup_d##D(int mode, int which, int gcls, int nz); \
    const void* dgmm_full_lookup_d##D();                                      \
    const void* dgmm_sample_lookup_d##D();                                    \
    const void* dgmm_post_logprob_lookup_d##D();
DGMM_DECL(1) DGMM_DECL(2) DGMM_DECL(3) DGMM_DECL(4) DGMM_DECL(5) DGMM_DECL(6) DGMM_DECL(7) DGMM_DECL(8)
DGMM_DECL(9) DGMM_DECL(10) DGMM_DECL(11) DGMM_DECL(12) DGMM_DECL(13) DGMM_DECL(14) DGMM_DECL(15) DGMM_DECL(16)
DGMM_DECL(17) DGMM_DECL(18) DGMM_DECL(19) DGMM_DECL(20) DGMM_DECL(21) DGMM_DECL(22) DGMM_DECL(23) DGMM_DECL(24)
DGMM_DECL(25) DGMM_DECL(26) DGMM_DECL(27) DGMM_DECL(28) DGMM_DECL(29) DGMM_DECL(30) DGMM_DECL(31) DGMM_DECL(32)
#undef DGMM_DECL
// cooperative (team) kernels, D = 17..32 (dgmm_team_dXX.cu)
#define DGMM_TDECL(D) dgmm::TeamKernels dgmm_team_lookup_d##D();
DGMM_TDECL(17) DGMM_TDECL(18) DGMM_TDECL(19) DGMM_TDECL(20) DGMM_TDECL(21) DGMM_TDECL(22) DGMM_TDECL(23) DGMM_TDECL(24)
DGMM_TDECL(25) DGMM_TDECL(26) DGMM_TDECL(27) DGMM_TDECL(28) DGMM_TDECL(29) DGMM_TDECL(30) DGMM_TDECL(31) DGMM_TDECL(32)
#undef DGMM_TDECL

namespace {

using namespace dgmm;

typedef PairVariant (*pair_lookup_t)(int, int, int, int);
typedef const void* (*full_lookup_t)();

#define DGMM_ROW(D) {dgmm_pair_lookup_d##D, dgmm_full_lookup_d##D, dgmm_sample_lookup_d##D, dgmm_post_logprob_lookup_d##D}
struct DimRow {
    pair_lookup_t pair;
    full_lookup_t full;
    full_lookup_t sample;
    full_lookup_t post_logprob;
};
const DimRow kDims[DGMM_MAX_D + 1] = {
    {nullptr, nullptr, nullptr, nullptr},
    DGMM_ROW(1),  DGMM_ROW(2),  DGMM_ROW(3),  DGMM_ROW(4),  DGMM_ROW(5),  DGMM_ROW(6),  DGMM_ROW(7),  DGMM_ROW(8),
    DGMM_ROW(9),  DGMM_ROW(10), DGMM_ROW(11), DGMM_ROW(12), DGMM_ROW(13), DGMM_ROW(14), DGMM_ROW(15), DGMM_ROW(16),
    DGMM_ROW(17), DGMM_ROW(18), DGMM_ROW(19), DGMM_ROW(20), DGMM_ROW(21), DGMM_ROW(22), DGMM_ROW(23), DGMM_ROW(24),
    DGMM_ROW(25), DGMM_ROW(26), DGMM_ROW(27), DGMM_ROW(28), DGMM_ROW(29), DGMM_ROW(30), DGMM_ROW(31), DGMM_ROW(32)};

typedef TeamKernels (*team_lookup_t)();
const team_lookup_t kTeam[DGMM_MAX_D + 1] = {
    nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
    nullptr, nullptr, nullptr, nullptr,
    dgmm_team_lookup_d17, dgmm_team_lookup_d18, dgmm_team_lookup_d19, dgmm_team_lookup_d20, dgmm_team_lookup_d21,
    dgmm_team_lookup_d22, dgmm_team_lookup_d23, dgmm_team_lookup_d24, dgmm_team_lookup_d25, dgmm_team_lookup_d26,
    dgmm_team_lookup_d27, dgmm_team_lookup_d28, dgmm_team_lookup_d29, dgmm_team_lookup_d30, dgmm_team_lookup_d31,
    dgmm_team_lookup_d32};

#define CUDA_TRY(x)                                  \
    do {                                             \
        cudaError_t e__ = (x);                       \
        if (e__ != cudaSuccess) return (int)e__;     \
    } while (0)

struct DeviceInfo {
    int device = -1, sms = 0, smem_optin = 0, cc_major = 0;
};

int device_info(DeviceInfo* out) {
    static DeviceInfo cache[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return DGMM_ERR_NO_DEVICE;
    if (dev < 0 || dev >= 64) return DGMM_ERR_NO_DEVICE;
    if (cache[dev].device != dev) {
        DeviceInfo d;
        d.device = dev;
        if (cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return DGMM_ERR_NO_DEVICE;
        cudaDeviceGetAttribute(&d.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev);
        if (d.cc_major != 10) return DGMM_ERR_NO_DEVICE;  // sm_100a cubins only
        cache[dev] = d;
    }
    *out = cache[dev];
    return 0;
}

inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }
inline int group_size(int K) {
    if (K <= 32) {
        int g = 1;
        while (g < K) g <<= 1;
        return g;
    }
    return (K + 31) & ~31;
}
inline int nd_of(int D) { return D * (D + 1) / 2; }
inline int ns_of(int D) { return 1 + D + nd_of(D); }

// Upper bound of the grid used for (D, K): shared by the workspace query and the launcher.
inline int max_grid(int D, int K, int sms) {
    const size_t per_block = (size_t)ns_of(D) * std::max(kPairMaxThreads, group_size(K)) * 4;  // a column per thread, worst case
    size_t cap = ((size_t)256 << 20) / per_block;
    size_t g = (size_t)sms * 8;
    if (g > cap) g = cap;
    if (g < (size_t)sms) g = sms;
    return (int)g;
}

struct Workspace {
    float* sig_cols;
    float* mu_cols;
    float* cj;
    int* nonpd;
    double* part_ll;
    float* part;
    // cooperative path (D > 16): parameters in the register order of the team lanes
    float* team_sig;
    float* team_mu;
    float* team_cj;
    size_t bytes;
};

Workspace carve(void* base, int D, int K, int sms) {
    const int GS = group_size(K);
    Workspace w;
    size_t off = 0;
    char* b = (char*)base;
    w.sig_cols = (float*)(b + off); off += up256((size_t)nd_of(D) * GS * 4);
    w.mu_cols = (float*)(b + off);  off += up256((size_t)D * GS * 4);
    w.cj = (float*)(b + off);       off += up256((size_t)GS * 4);
    w.nonpd = (int*)(b + off);      off += 256;
    const int mg = max_grid(D, K, sms);
    // (the cooperative path keeps the partial sums of its two passes side by side)
    w.part_ll = (double*)(b + off); off += up256((size_t)mg * (D > 16 ? 4 : 2) * 8);
    w.part = (float*)(b + off);     off += up256((size_t)mg * ns_of(D) * std::max(kPairMaxThreads, GS) * 4);
    w.team_sig = w.team_mu = w.team_cj = nullptr;
    if (D > 16) {
        const size_t Kp = (size_t)K + team::kMaxTB;
        w.team_sig = (float*)(b + off); off += up256(Kp * team::geo_nreg(D) * team::kP * 4);
        w.team_mu = (float*)(b + off);  off += up256(Kp * team::geo_R(D) * team::kP * 4);
        w.team_cj = (float*)(b + off);  off += up256(Kp * 4);
    }
    w.bytes = off;
    return w;
}

// ---------------------------------------------------------------------------------------
// small kernels
// ---------------------------------------------------------------------------------------

// Transposes the mixture parameters into per-thread columns:
//   sig_cols[e][j] = lower(Sigma_j)[e] (+ shared noise), mu_cols[a][j], cj[j] = log alpha_j - D/2 log 2pi.
// Padding components j >= K get Sigma = I, mu = 0, cj = -inf.
__global__ void prepare_kernel(const float* __restrict__ log_alpha, const float* __restrict__ mu,
                               const float* __restrict__ Sigma, const float* __restrict__ S_shared, int D,
                               int K, int GS, float* __restrict__ sig_cols, float* __restrict__ mu_cols,
                               float* __restrict__ cj, int* __restrict__ nonpd) {
    const int ND = D * (D + 1) / 2;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int nth = gridDim.x * blockDim.x;
    if (tid == 0 && nonpd) *nonpd = 0;
    for (int idx = tid; idx < ND * GS; idx += nth) {
        const int e = idx / GS, j = idx - e * GS;
        int a = 0;
        while ((a + 1) * (a + 2) / 2 <= e) ++a;
        const int b = e - a * (a + 1) / 2;
        float v;
        if (j < K) {
            v = Sigma[((size_t)j * D + a) * D + b];
            if (S_shared) v += S_shared[a * D + b];
        } else {
            v = (a == b) ? 1.f : 0.f;
        }
        sig_cols[idx] = v;
    }
    for (int idx = tid; idx < D * GS; idx += nth) {
        const int a = idx / GS, j = idx - a * GS;
        mu_cols[idx] = (j < K) ? mu[(size_t)j * D + a] : 0.f;
    }
    for (int j = tid; j < GS; j += nth) cj[j] = (j < K) ? (log_alpha[j] - D * kHalfLog2Pi) : -INFINITY;
}

// Sums the per-block partials (double accumulation, fixed order: deterministic) into the
// packed statistics buffer.  beta = 1 accumulates onto the existing contents (host pipeline).
// One block per (packed statistic e, tile of 32 components): lane x -> component j (coalesced
// 128-byte reads of the partials), the 32 warps take the (block, slot) rows r = y, y + 32, ..., and
// their sums are combined in warp order.  The last block reduces the (sum ll, sum g) pairs.
constexpr int kFinY = 32;   // (8 warps: 15 us at c2's 444 x 66 x 64 partials; 32 warps: the rows are 4x shorter)
__global__ void __launch_bounds__(32 * kFinY) finalize_stats_kernel(
    const float* __restrict__ part, const double* __restrict__ part_ll, const int* __restrict__ nonpd, int nblk,
    int D, int K, int GS, int cols, int nslot, float* __restrict__ stats, int beta) {
    __shared__ double sh[2][kFinY][32];
    const int ND = D * (D + 1) / 2, NS = 1 + D + ND;
    const int njt = (K + 31) / 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int total = K * (1 + D + D * D);
    if ((int)blockIdx.x < NS * njt) {
        const int e = blockIdx.x / njt, j = (blockIdx.x - e * njt) * 32 + tx;
        // `cols` columns per block: GS when the block already summed its slots, TPB (= nslot * GS)
        // when the accumulators are per-thread global columns
        const int rows = nblk * nslot;
        double s = 0.0;
        if (j < K) {
#pragma unroll 4
            for (int r = ty; r < rows; r += kFinY) {
                const int blk = r / nslot, sl = r - blk * nslot;
                s += (double)part[((size_t)blk * NS + e) * cols + sl * GS + j];
            }
        }
        sh[0][ty][tx] = s;
        __syncthreads();
        if (ty == 0 && j < K) {
            double t = 0.0;
#pragma unroll
            for (int y = 0; y < kFinY; ++y) t += sh[0][y][tx];
            const float v = (float)t;
            if (e == 0) {
                stats[j] = beta ? stats[j] + v : v;
            } else if (e < 1 + D) {
                float* o = stats + K + j * D + (e - 1);
                *o = beta ? *o + v : v;
            } else {
                int a = 0;
                const int tt = e - 1 - D;
                while ((a + 1) * (a + 2) / 2 <= tt) ++a;
                const int b = tt - a * (a + 1) / 2;
                float* base = stats + K + K * D + (size_t)j * D * D;
                const float nv = beta ? base[a * D + b] + v : v;
                base[a * D + b] = nv;
                base[b * D + a] = nv;       // the full symmetric matrix is stored
            }
        }
    } else {
        const int t = ty * 32 + tx;
        double ll = 0.0, sg = 0.0;
        for (int blk = t; blk < nblk; blk += 32 * kFinY) { ll += part_ll[2 * blk]; sg += part_ll[2 * blk + 1]; }
        sh[0][ty][tx] = ll;
        sh[1][ty][tx] = sg;
        __syncthreads();
        if (t == 0) {
            ll = 0.0; sg = 0.0;
            for (int y = 0; y < kFinY; ++y)
                for (int x = 0; x < 32; ++x) { ll += sh[0][y][x]; sg += sh[1][y][x]; }
            float* tail = stats + total;
            if (beta) ll += (double)tail[0] + (double)tail[1];
            const float hi = (float)ll;
            tail[0] = hi;
            tail[1] = isfinite(hi) ? (float)(ll - (double)hi) : 0.f;
            tail[2] = beta ? tail[2] + (float)sg : (float)sg;
            tail[3] = beta ? tail[3] + (float)(*nonpd) : (float)(*nonpd);
        }
    }
}

__global__ void finalize_ll_kernel(const double* __restrict__ part_ll, const int* __restrict__ nonpd, int nblk,
                                   double* __restrict__ out_sum, int* __restrict__ status, int beta) {
    // one warp: lane-strided partial sums, then a fixed-order butterfly (deterministic)
    double ll = 0.0;
    for (int blk = threadIdx.x; blk < nblk; blk += 32) ll += part_ll[2 * blk];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ll += __shfl_xor_sync(0xffffffffu, ll, o);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (out_sum) {
            out_sum[0] = beta ? out_sum[0] + ll : ll;
            out_sum[1] = beta ? out_sum[1] + (double)(*nonpd) : (double)(*nonpd);
        }
        if (status) *status = *nonpd;
    }
}

// Closed-form EM update from the packed statistics; one block per component, D*D threads.
// deconv/gmm/deconv_gmm.py:81-109 via the identity of SURVEY.md Appendix C.
__global__ void mstep_kernel(const float* __restrict__ stats, double n_total, double w, int D, int K,
                             float* __restrict__ mu, float* __restrict__ Sigma, float* __restrict__ alpha,
                             float* __restrict__ log_alpha) {
    extern __shared__ double sh[];
    double* Sg = sh;            // [D*D]
    double* S2 = Sg + D * D;    // [D*D]
    double* W = S2 + D * D;     // [D*D]  Sigma S2
    double* v = W + D * D;      // [D]    Sigma s1
    const int j = blockIdx.x;
    const int t = threadIdx.x;
    const int a = t / D, b = t - a * D;
    const double s0 = (double)stats[j];
    const float* s1 = stats + K + (size_t)j * D;
    const float* s2 = stats + K + (size_t)K * D + (size_t)j * D * D;
    float* Sj = Sigma + (size_t)j * D * D;
    // the kernels only ever read the lower triangle; mirror it so that an asymmetric
    // caller-supplied Sigma cannot leak into the update
    Sg[t] = (double)Sj[a >= b ? a * D + b : b * D + a];
    S2[t] = (double)s2[t];
    __syncthreads();
    double acc = 0.0;
    for (int c = 0; c < D; ++c) acc += Sg[a * D + c] * S2[c * D + b];
    W[t] = acc;
    if (b == 0) {
        double vv = 0.0;
        for (int c = 0; c < D; ++c) vv += Sg[a * D + c] * (double)s1[c];
        v[a] = vv;
    }
    __syncthreads();
    double P = 0.0;
    for (int c = 0; c < D; ++c) P += W[a * D + c] * Sg[c * D + b];
    const double da = v[a] / s0, db = v[b] / s0;
    double nv = Sg[t] + P / s0 - da * db;
    if (a == b) nv += 2.0 * w / s0;
    Sj[t] = (float)nv;
    if (b == 0) mu[(size_t)j * D + a] = (float)((double)mu[(size_t)j * D + a] + da);
    if (t == 0) {
        const double al = s0 / n_total;
        alpha[j] = (float)al;
        log_alpha[j] = (float)log(al);
    }
}

// Online-EM M-step of one minibatch from the packed statistics; one block per component, D*D threads,
// double precision inside.  deconv/gmm/online_deconv_gmm.py:146-188 via SURVEY.md Appendix C:
//   batch: sr = s0 + floor, scm = s0 mu + Sigma s1, bm = scm / sr,
//          bc = [s0 (Sigma + c c^T) + Sigma S2 Sigma + c v^T + v c^T] / sr   (c = mu - bm, v = Sigma s1)
//   EMA:   SR' = (1-g) SR + g sr, SCM' = (1-g) SCM + g scm, mu' = SCM' / SR'
//   Sigma' = (1-g) adj(Sigma - wI, SR/SR', mu, mu') + g adj(bc, sr/SR', bm, mu') + wI,
//            adj(C, s, b, c) = s (C + b b^T) - c c^T   (_adjust, :134-144);   alpha' = SR' / global_batch.
// A minibatch with a non-PD pair leaves everything untouched and sets the sticky crash flag (:79-80).
__global__ void online_mstep_kernel(const float* __restrict__ stats, double floor_, double g, double w,
                                    double global_batch, int D, int K, float* __restrict__ mu,
                                    float* __restrict__ Sigma, float* __restrict__ SR, float* __restrict__ SCM,
                                    float* __restrict__ alpha, float* __restrict__ log_alpha,
                                    int* __restrict__ crash) {
    extern __shared__ double sh[];
    double* Sg = sh;            // [D*D]
    double* S2 = Sg + D * D;    // [D*D]
    double* W = S2 + D * D;     // [D*D]  Sigma S2
    double* v = W + D * D;      // [D]    Sigma s1
    const int j = blockIdx.x, t = threadIdx.x;
    const int a = t / D, b = t - a * D;
    const float* tail = stats + (size_t)K * (1 + D + D * D);
    if ((crash && *crash) || tail[3] > 0.f) {
        if (crash && j == 0 && t == 0) *crash = 1;
        return;
    }
    const double s0 = (double)stats[j];
    const float* s1 = stats + K + (size_t)j * D;
    const float* s2 = stats + K + (size_t)K * D + (size_t)j * D * D;
    float* Sj = Sigma + (size_t)j * D * D;
    float* muj = mu + (size_t)j * D;
    Sg[t] = (double)Sj[a >= b ? a * D + b : b * D + a];
    S2[t] = (double)s2[t];
    __syncthreads();
    double acc = 0.0;
    for (int c = 0; c < D; ++c) acc += Sg[a * D + c] * S2[c * D + b];
    W[t] = acc;
    if (b == 0) {
        double vv = 0.0;
        for (int c = 0; c < D; ++c) vv += Sg[a * D + c] * (double)s1[c];
        v[a] = vv;
    }
    __syncthreads();
    double P = 0.0;
    for (int c = 0; c < D; ++c) P += W[a * D + c] * Sg[c * D + b];
    const double sr = s0 + floor_;
    const double mua = (double)muj[a], mub = (double)muj[b];
    const double scm_a = s0 * mua + v[a], scm_b = s0 * mub + v[b];
    const double bm_a = scm_a / sr, bm_b = scm_b / sr;
    const double ca = mua - bm_a, cb = mub - bm_b;
    const double bc = (s0 * (Sg[t] + ca * cb) + P + ca * v[b] + v[a] * cb) / sr;
    const double sr_old = (double)SR[j];
    const double sr_new = (1.0 - g) * sr_old + g * sr;
    const double scmn_a = (1.0 - g) * (double)SCM[(size_t)j * D + a] + g * scm_a;
    const double scmn_b = (1.0 - g) * (double)SCM[(size_t)j * D + b] + g * scm_b;
    const double mn_a = scmn_a / sr_new, mn_b = scmn_b / sr_new;
    const double so = sr_old / sr_new, sb = sr / sr_new;
    const double wd = (a == b) ? w : 0.0;
    const double old_part = so * ((Sg[t] - wd) + mua * mub) - mn_a * mn_b;
    const double new_part = sb * (bc + bm_a * bm_b) - mn_a * mn_b;
    const double cov = (1.0 - g) * old_part + g * new_part + wd;
    __syncthreads();   // every read of mu / SCM / SR of this component is done
    Sj[t] = (float)cov;
    if (b == 0) {
        muj[a] = (float)mn_a;
        SCM[(size_t)j * D + a] = (float)scmn_a;
    }
    if (t == 0) {
        SR[j] = (float)sr_new;
        const double al = sr_new / global_batch;
        alpha[j] = (float)al;
        log_alpha[j] = (float)log(al);
    }
}

__global__ void log_kernel(const float* __restrict__ in, float* __restrict__ out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = logf(in[i]);
}

// FP32 FMA probe: 8 independent chains per thread, register operands only.
__global__ void __launch_bounds__(256) fp32_probe_kernel(float* sink, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f;
    float a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 0.999f + blockIdx.x * 1e-9f, c = 1e-3f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
        }
    }
    const float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123.456f) *sink = s;
}

// ---------------------------------------------------------------------------------------
// launch planning for the pair kernel
// ---------------------------------------------------------------------------------------
struct Tuning {
    // sigp/accp both ST_GLOBAL (2) force the fallback variant; rc selects the placement
    // (0 = main, 1 = alternative); the rest override the tile geometry.  -1 = automatic.
    int sigp = -1, accp = -1, nslot = -1, pt = -1, nst = -1, blocks_per_sm = -1, rc = -1, no_bulk = 0;
};
// Diagnostics state is per host thread: two estimators driven from two threads (each on its own stream) do not
// see each other's tuning overrides or timing records.
thread_local Tuning g_tuning;

struct Plan {
    const void* kernel = nullptr;
    int sigp, accp, rc, GS, NSLOT, TPB, PT, NST, grid, use_bulk, gcls, tmem_cols;
    size_t smem;
    long long n_tiles;
};

size_t pair_smem_bytes(int D, int noise, bool has_g, int sigp, int accp, int GS, int TPB, int PT, int NST,
                       int mode, int nz) {
    const StageLayout SL = stage_layout(D, noise, PT, has_g);
    size_t b = kSmemHeader + (size_t)NST * SL.bytes;
    b += (size_t)PT * repack_floats_per_point(D, nz) * 4;   // re-packed lower triangles of the tile
    if (sigp == ST_SMEM) b += (size_t)pad_col(nd_of(D)) * GS * 4;
    if (mode == MODE_STATS && accp == ST_SMEM) b += (size_t)pad_col(ns_of(D)) * TPB * 4;
    if (mode == MODE_LOGLIK) b += (size_t)2 * PT * (GS + 16) * 4;  // deferred logsumexp rows
    // the register-accumulator epilogue reuses the ring as TPB floats of scratch
    if (b < kSmemHeader + (size_t)TPB * 4) b = kSmemHeader + (size_t)TPB * 4;
    return b;
}

inline int group_class(int GS) {
    if (GS < 32) return GCLS_SUB;
    if (GS == 32) return GCLS_WARP;
    return GS <= 128 ? GCLS_MW4 : (GS <= kPairMaxThreads ? GCLS_MWN : GCLS_MWN512);
}

// Tile geometry for one variant; returns false when it cannot be made to fit.
bool fit_geometry(const DeviceInfo& dev, const PairVariant& v, int mode, int D, int noise, bool has_g, Plan* p) {
    const Tuning& t = g_tuning;
    const int GS = p->GS;
    int nslot = std::max(1, 128 / GS);
    if (t.nslot > 0) nslot = t.nslot;
    if (GS > 32 && nslot > 15) nslot = 15;            // named barriers 1..15
    while (GS * nslot > v.max_threads && nslot > 1) --nslot;
    if (GS * nslot > v.max_threads) return false;
    while (GS * nslot < 32) ++nslot;                   // whole warps only (full-mask shuffles)
    // D <= 10: 16 points per slot and tile (the per-tile costs -- barrier, ring refill, the
    // log-likelihood pass's logsumexp sweep -- are amortised over 16 trips of the point loop), as long
    // as three blocks still fit in an SM's shared memory; measured sweep: profiles/r01_pt_sweep.txt
    int pt0 = (D <= 10) ? std::min(16 * nslot, 64) : (D <= 16 ? 8 : 4);
    int nst = (D <= 16) ? 3 : 2;
    if (t.nst > 0) nst = std::min(t.nst, kMaxStages);
    const size_t smem_3blocks = (size_t)74 * 1024;
    for (;;) {
        int PT = t.pt > 0 ? t.pt : pt0;
        int lcm = nslot;                               // PT: a multiple of NSLOT and of 4
        while (lcm % 4) lcm += nslot;
        PT = ((PT + lcm - 1) / lcm) * lcm;
        const size_t smem = pair_smem_bytes(D, noise, has_g, v.sigp, v.accp, GS, GS * nslot, PT, nst, mode, v.nz);
        const bool tmem_ok = v.accp != ST_TMEM || mode != MODE_STATS ||
                             tmem_alloc_cols(ns_of(D), GS * nslot / 32) <= 512;
        if (D <= 10 && t.pt <= 0 && smem > smem_3blocks && v.min_blocks >= 3) {
            if (nst > 2 && t.nst <= 0) { nst = 2; continue; }
            if (pt0 > 16 && pt0 / 2 >= lcm) { pt0 /= 2; nst = (t.nst > 0) ? nst : 3; continue; }
        }
        if (smem <= (size_t)dev.smem_optin && tmem_ok) {
            p->NSLOT = nslot; p->TPB = GS * nslot; p->PT = PT; p->NST = nst; p->smem = smem;
            return true;
        }
        if (GS >= 32 && nslot > 1) { nslot = 1; continue; }
        if (GS < 32 && GS * nslot > 32) { nslot /= 2; continue; }
        if (nst > 1) { --nst; continue; }
        return false;
    }
}

int make_plan(const DeviceInfo& dev, int mode, int D, int K, long long N, int noise, bool has_g,
              const void* X, const void* S, const void* g, Plan* out) {
    Plan p;
    p.GS = group_size(K);
    p.gcls = group_class(p.GS);
    const Tuning& t = g_tuning;
    const bool force_fallback = (t.sigp == ST_GLOBAL && t.accp == ST_GLOBAL);
    int which = (t.rc == 1 || t.rc == 2) ? t.rc : 0;   // rc = 3 / 4 choose between the kernel families (use_team_path)
    PairVariant v{nullptr, 0, 0, 0, 0, 0, 0};
    const int nz = (noise == NOISE_FULL) ? NZ_FULL : (noise == NOISE_PACKED ? NZ_PACKED : NZ_DIAG);
    bool ok = false;
    if (!force_fallback) {
        v = kDims[D].pair(mode, which, p.gcls, nz);
        ok = v.fn && fit_geometry(dev, v, mode, D, noise, has_g, &p);
        if (!ok && which != 0) {
            which = 0;
            v = kDims[D].pair(mode, 0, p.gcls, nz);
            ok = v.fn && fit_geometry(dev, v, mode, D, noise, has_g, &p);
        }
    }
    if (ok && which == 0 && D > 16 && v.sigp == ST_SMEM && (t.rc < 0 || t.rc == 4) && 2 * (p.smem + 1024) > (size_t)dev.smem_optin) {
        // D = 17..24: Sigma_j columns in shared memory leave room for one block per SM only
        Plan q = p;
        const PairVariant alt = kDims[D].pair(mode, 1, p.gcls, nz);
        if (alt.fn && fit_geometry(dev, alt, mode, D, noise, has_g, &q) && 2 * (q.smem + 1024) <= (size_t)dev.smem_optin) {
            v = alt; p = q; which = 1;
        }
    }
    if (!ok) {
        which = 2;
        v = kDims[D].pair(mode, 2, GCLS_ANY, NZ_ANY);
        if (!v.fn) return DGMM_ERR_UNSUPPORTED_D;
        if (!fit_geometry(dev, v, mode, D, noise, has_g, &p)) return DGMM_ERR_UNSUPPORTED_K;
    }
    p.kernel = v.fn;
    p.sigp = v.sigp; p.accp = v.accp; p.rc = which;
    if (mode == MODE_LOGLIK) p.accp = ST_GLOBAL;

    CUDA_TRY(cudaFuncSetAttribute(p.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
    int occ = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, p.kernel, p.TPB, p.smem));
    if (occ < 1) return DGMM_ERR_UNSUPPORTED_K;
    p.tmem_cols = 0;
    const bool sig_tmem = (v.sigp == ST_TMEM);
    if ((mode == MODE_STATS && p.accp == ST_TMEM) || sig_tmem) {
        // The occupancy query answers 1 block/SM for any kernel that allocates tensor memory, yet
        // blocks whose columns fit side by side do run concurrently (tools/ubench_tmem.cu): derive
        // the residency from registers, shared memory and the 512 TMEM columns instead.  All
        // resident blocks must hold their columns at the same time (a block that cannot allocate
        // would spin until another one exits).
        p.tmem_cols = tmem_alloc_cols(sig_tmem ? nd_of(D) : ns_of(D), p.TPB / 32);
        if (p.tmem_cols > 512) return DGMM_ERR_UNSUPPORTED_K;
        cudaFuncAttributes fa;
        CUDA_TRY(cudaFuncGetAttributes(&fa, p.kernel));
        const int regs_per_block = ((fa.numRegs + 7) & ~7) * p.TPB;
        int smem_sm = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev.device));
        int o = 65536 / std::max(regs_per_block, 1);
        o = std::min<int>(o, smem_sm / (int)(p.smem + fa.sharedSizeBytes + 1024));
        o = std::min(o, 2048 / p.TPB);
        o = std::min(o, 512 / p.tmem_cols);
        occ = std::max(o, 1);
    }
    if (t.blocks_per_sm > 0) occ = std::min(occ, t.blocks_per_sm);
    const long long slots = std::min<long long>((long long)dev.sms * occ, max_grid(D, K, dev.sms));
    // Minibatch-sized launches (a block sees fewer than 16 tiles): the round-robin of whole tiles leaves the blocks
    // with ceil(tiles / blocks) or one tile less, e.g. 65 536 points on 444 blocks in tiles of 32 = 4 or 5 tiles (128 or
    // 160 points, 8 % above the mean; ncu: 87 % of the elapsed SM cycles active).  Pick the tile size <= the default one
    // (shared memory and occupancy can only improve) that minimises the points of the busiest block, with a
    // fixed cost of 2 points per tile (barrier, ring refill, per-tile pass): 30-point tiles give 5 x 30 = 150 points per
    // block there (measured, N = 65 536: D=10,K=64 statistics 0.239 -> 0.234 ms, log-lik 0.131 -> 0.126 ms).  Only for
    // the 24+-point tiles of D <= 10, K <= 64 and never below 3/4 of the default: smaller tiles lose more to the per-tile
    // costs than they win (D=9,K=128 with 8- instead of 16-point tiles: 0.339 -> 0.353 ms).
    // Tile sizes stay multiples of NSLOT and keep every sub-tile of a stage a whole number of 16-byte units (TMA).
    if (t.pt <= 0 && N > 0 && p.PT >= 24 && (N + p.PT - 1) / p.PT < 16 * slots && (N + p.PT - 1) / p.PT > slots) {
        const int n_s = noise_floats_per_point(D, noise);
        auto ok16 = [&](int c) {
            return (c % p.NSLOT) == 0 && ((long long)c * D * 4) % 16 == 0 && ((long long)c * n_s * 4) % 16 == 0 &&
                   (!has_g || (c * 4) % 16 == 0);
        };
        int best = p.PT;
        double best_cost = 1e300;
        for (int c = p.PT; c >= std::max(p.NSLOT, (3 * p.PT) / 4); --c) {
            if (!ok16(c)) continue;
            const long long tiles = (N + c - 1) / c;
            const long long per_block = (tiles + slots - 1) / slots;
            const double cost = (double)per_block * (c + 2.0);
            if (cost < best_cost - 1e-9) { best_cost = cost; best = c; }
        }
        if (best != p.PT) {
            p.PT = best;
            p.smem = pair_smem_bytes(D, noise, has_g, v.sigp, v.accp, p.GS, p.TPB, p.PT, p.NST, mode, v.nz);
        }
    }
    p.n_tiles = (N + p.PT - 1) / p.PT;
    long long grid = std::min<long long>(slots, std::max<long long>(p.n_tiles, 1));
    p.grid = (int)grid;

    const bool per_point = noise == NOISE_FULL || noise == NOISE_DIAG || noise == NOISE_PACKED;
    const bool aligned = (((uintptr_t)X | (uintptr_t)(per_point ? S : nullptr) | (uintptr_t)g) & 15) == 0;
    p.use_bulk = (aligned && !t.no_bulk) ? 1 : 0;
    *out = p;
    return 0;
}

int check_common(const void* X, const void* S, int noise, int64_t N, int D, int K, const void* la, const void* mu,
                 const void* Sg) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1 || K > DGMM_MAX_K) return DGMM_ERR_UNSUPPORTED_K;
    if (noise < 0 || noise > 4) return DGMM_ERR_LAYOUT;
    if (N < 0 || !la || !mu || !Sg) return DGMM_ERR_BAD_ARG;
    if (N > 0 && !X) return DGMM_ERR_BAD_ARG;
    if (N > 0 && noise != NOISE_NONE && !S) return DGMM_ERR_BAD_ARG;
    return 0;
}

// Optional per-launch timing of the pair kernel (bench.py's roofline line): CUDA events
// recorded on the launching stream directly around the launch, kept in a ring.
struct PairTiming {
    bool on = false;
    int count = 0;
    static constexpr int kRing = 512;
    cudaEvent_t a[kRing], b[kRing];
    bool made = false;
};
thread_local PairTiming g_timing;

int timing_begin(cudaStream_t st) {
    if (!g_timing.on) return -1;
    if (!g_timing.made) {
        for (int i = 0; i < PairTiming::kRing; ++i) {
            cudaEventCreate(&g_timing.a[i]);
            cudaEventCreate(&g_timing.b[i]);
        }
        g_timing.made = true;
    }
    const int slot = g_timing.count % PairTiming::kRing;
    cudaEventRecord(g_timing.a[slot], st);
    return slot;
}
void timing_end(int slot, cudaStream_t st) {
    if (slot < 0) return;
    cudaEventRecord(g_timing.b[slot], st);
    ++g_timing.count;
}

// ---------------------------------------------------------------------------------------
// cooperative (team) path, D > 16: parameter layout, launch plan, finalise
// ---------------------------------------------------------------------------------------

// Mixture parameters in the register order of the team lanes (dgmm_team.cuh):
//   team_sig[((j * NQ + q) * 8 + l) * 4 + c] = lower(Sigma_j + S_shared)[a][b],  4q + c -> (slot r, column b), a = l + 8r
// (0 outside the lower triangle / beyond D; identity for the padding components j >= K),
//   team_mu[(j * R + r) * 8 + l] = mu_j[a],  team_cj[j] = log alpha_j - D/2 log 2pi (-inf for j >= K).
__global__ void team_prepare_kernel(const float* __restrict__ log_alpha, const float* __restrict__ mu,
                                    const float* __restrict__ Sigma, const float* __restrict__ S_shared, int D, int K,
                                    int Kp, team::GeoRT geo, float* __restrict__ sig, float* __restrict__ muT,
                                    float* __restrict__ cj, int* __restrict__ nonpd) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    if (tid == 0 && nonpd) *nonpd = 0;
    const int per = geo.NREG * team::kP;
    for (long long idx = tid; idx < (long long)Kp * per; idx += nth) {
        const int j = (int)(idx / per), rem = (int)(idx - (long long)j * per);
        const int q = rem / 32, l = (rem % 32) / 4, c = rem % 4;
        const int flat = 4 * q + c;
        int r = 0;
        while (r + 1 < geo.R && flat >= geo.off[r + 1]) ++r;
        const int b = flat - geo.off[r], a = l + 8 * r;
        float v = 0.f;
        if (a < D && b <= a) {
            if (j < K) {
                v = Sigma[((size_t)j * D + a) * D + b];
                if (S_shared) v += S_shared[a * D + b];
            } else {
                v = (a == b) ? 1.f : 0.f;
            }
        }
        sig[idx] = v;
    }
    for (long long idx = tid; idx < (long long)Kp * geo.R * team::kP; idx += nth) {
        const int j = (int)(idx / (geo.R * team::kP)), rem = (int)(idx % (geo.R * team::kP));
        const int r = rem / team::kP, l = rem % team::kP, a = l + 8 * r;
        muT[idx] = (j < K && a < D) ? mu[(size_t)j * D + a] : 0.f;
    }
    for (long long j = tid; j < Kp; j += nth) cj[j] = (j < K) ? (log_alpha[j] - D * kHalfLog2Pi) : -INFINITY;
}

// Sums the per-block partials of the team statistics pass ([grid][TB][NS], block = slice * n_chunks + chunk)
// over the slices in a fixed order (double accumulation: deterministic) into the packed buffer; the last
// block assembles the tail from the two passes' (sum ll | sum g) partials.
__global__ void __launch_bounds__(128) team_finalize_kernel(const float* __restrict__ part, const double* __restrict__ pll1,
                                                            int grid1, const double* __restrict__ pll2, int grid2,
                                                            const int* __restrict__ nonpd, int D, int K, int TB,
                                                            int n_chunks, int n_slices, float* __restrict__ stats,
                                                            int beta) {
    const int ND = D * (D + 1) / 2, NS = 1 + D + ND;
    const int total = K * (1 + D + D * D);
    const int j = blockIdx.x;
    if (j < K) {
        const int chunk = j / TB, tm = j - chunk * TB;
        for (int e = threadIdx.x; e < NS; e += blockDim.x) {
            double s = 0.0;
            for (int sl = 0; sl < n_slices; ++sl)
                s += (double)part[(((size_t)sl * n_chunks + chunk) * TB + tm) * NS + e];
            const float v = (float)s;
            if (e == 0) {
                stats[j] = beta ? stats[j] + v : v;
            } else if (e < 1 + D) {
                float* o = stats + K + j * D + (e - 1);
                *o = beta ? *o + v : v;
            } else {
                int a = 0;
                const int tt = e - 1 - D;
                while ((a + 1) * (a + 2) / 2 <= tt) ++a;
                const int b = tt - a * (a + 1) / 2;
                float* base = stats + K + K * D + (size_t)j * D * D;
                const float nv = beta ? base[a * D + b] + v : v;
                base[a * D + b] = nv;
                base[b * D + a] = nv;
            }
        }
    } else {
        __shared__ double sh[2][128];
        double ll = 0.0, sg = 0.0;
        for (int b = threadIdx.x; b < grid1; b += blockDim.x) ll += pll1[2 * b];
        for (int b = threadIdx.x; b < grid2; b += blockDim.x) sg += pll2[2 * b];
        sh[0][threadIdx.x] = ll;
        sh[1][threadIdx.x] = sg;
        __syncthreads();
        if (threadIdx.x == 0) {
            ll = 0.0; sg = 0.0;
            for (int x = 0; x < (int)blockDim.x; ++x) { ll += sh[0][x]; sg += sh[1][x]; }
            float* tail = stats + total;
            if (beta) ll += (double)tail[0] + (double)tail[1];
            const float hi = (float)ll;
            tail[0] = hi;
            tail[1] = isfinite(hi) ? (float)(ll - (double)hi) : 0.f;
            tail[2] = beta ? tail[2] + (float)sg : (float)sg;
            tail[3] = beta ? tail[3] + (float)(*nonpd) : (float)(*nonpd);
        }
    }
}

struct TeamPlan {
    const void* kernel = nullptr;
    int TB, TPB, n_chunks, n_slices, grid, PT, NST, use_bulk, tmem_cols, nz;
    size_t smem;
    long long n_tiles;
};

// Which path serves (D, K): the cooperative kernels for D >= 21 and wherever the thread-per-pair family
// has no unrolled variant (K < 64 above D = 16); tuning rc = 3 forces the team path, rc = 4 the pair path.
bool use_team_path(int D, int K, int mode, int noise) {
    if (D <= 16) return false;
    if (g_tuning.rc == 3) return true;
    const int nz = (noise == NOISE_FULL) ? NZ_FULL : (noise == NOISE_PACKED ? NZ_PACKED : NZ_DIAG);
    const PairVariant v = kDims[D].pair(mode, 0, group_class(group_size(K)), nz);
    if (!v.fn) return true;
    if (g_tuning.rc == 4) return false;
    return D >= 21;
}

int make_team_plan(const DeviceInfo& dev, int mode, int D, int K, long long N, int noise, bool has_g, const void* X,
                   const void* S, const void* g, const void* ll, TeamPlan* out) {
    TeamPlan p;
    const Tuning& t = g_tuning;
    p.n_chunks = (K + team::kMaxTB - 1) / team::kMaxTB;
    p.TB = (((K + p.n_chunks - 1) / p.n_chunks) + 3) & ~3;
    p.TPB = p.TB * team::kP;
    p.PT = t.pt > 0 ? ((t.pt + 3) & ~3) : 4;
    p.NST = t.nst > 0 ? std::min(t.nst, kMaxStages) : 2;
    const bool stats = (mode == MODE_STATS);
    p.nz = (noise == NOISE_FULL || noise == NOISE_DIAG || noise == NOISE_PACKED) ? team::TNZ_MATRIX : team::TNZ_NONE;
    const TeamKernels tk = kTeam[D]();
    p.kernel = stats ? tk.stats[p.nz] : tk.loglik[p.nz];
    if (!p.kernel) return DGMM_ERR_UNSUPPORTED_D;
    for (;;) {
        p.smem = team::team_smem_bytes(D, noise, p.PT, p.NST, has_g, stats, p.TPB);
        if (p.smem <= (size_t)dev.smem_optin) break;
        if (p.NST > 1) { --p.NST; continue; }
        return DGMM_ERR_UNSUPPORTED_K;
    }
    CUDA_TRY(cudaFuncSetAttribute(p.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
    int occ = 0;
    p.tmem_cols = 0;
    cudaFuncAttributes fa;
    CUDA_TRY(cudaFuncGetAttributes(&fa, p.kernel));
    int smem_sm = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev.device));
    {
        // residency from registers, shared memory, threads and (statistics pass) the 512 tensor-memory columns;
        // the occupancy query answers 1 block/SM for any kernel that allocates tensor memory
        const int regs_per_block = ((fa.numRegs + 7) & ~7) * p.TPB;
        int o = 65536 / std::max(regs_per_block, 1);
        o = std::min<int>(o, smem_sm / (int)(p.smem + fa.sharedSizeBytes + 1024));
        o = std::min(o, 2048 / p.TPB);
        o = std::min(o, 32);
        if (stats) {
            p.tmem_cols = tmem_alloc_cols(team::geo_nacc(D), p.TPB / 32);
            if (p.tmem_cols > 512) return DGMM_ERR_UNSUPPORTED_D;
            o = std::min(o, 512 / p.tmem_cols);
        }
        occ = std::max(o, 1);
    }
    if (t.blocks_per_sm > 0) occ = std::min(occ, t.blocks_per_sm);
    p.n_tiles = (N + p.PT - 1) / p.PT;
    const long long cap = std::min<long long>((long long)dev.sms * occ, max_grid(D, K, dev.sms));
    if (stats) {
        long long sl = std::max<long long>(1, cap / p.n_chunks);
        sl = std::min<long long>(sl, std::max<long long>(p.n_tiles, 1));
        p.n_slices = (int)sl;
        p.grid = p.n_slices * p.n_chunks;
    } else {
        p.n_slices = 0;
        p.grid = (int)std::min<long long>(cap, std::max<long long>(p.n_tiles, 1));
    }
    const bool per_point = noise == NOISE_FULL || noise == NOISE_DIAG || noise == NOISE_PACKED;
    const bool aligned = (((uintptr_t)X | (uintptr_t)(per_point ? S : nullptr) | (uintptr_t)g | (uintptr_t)ll) & 15) == 0;
    p.use_bulk = (aligned && !t.no_bulk) ? 1 : 0;
    *out = p;
    return 0;
}

// Cooperative path of run_pair: prepare + log-likelihood pass (+ statistics pass).  `ll` [N] is required
// (it carries ll_i from the first pass to the second).  MODE_STATS leaves the partial sums of both passes in
// the workspace: call launch_team_finalize afterwards.
struct TeamRun {
    TeamPlan ll_plan, st_plan;
};
int run_team(int mode, const float* X, const float* S, int noise, int64_t N, int D, int K, const float* log_alpha,
             const float* mu, const float* Sigma, const float* g, float resp_floor, float* ll, const DeviceInfo& dev,
             const Workspace& ws, cudaStream_t st, TeamRun* run) {
    if (!ll) return DGMM_ERR_BAD_ARG;
    const team::GeoRT geo = team::geo_rt(D);
    int rc = make_team_plan(dev, MODE_LOGLIK, D, K, N, noise, false, X, S, nullptr, ll, &run->ll_plan);
    if (rc) return rc;
    if (mode == MODE_STATS) {
        rc = make_team_plan(dev, MODE_STATS, D, K, N, noise, g != nullptr, X, S, g, ll, &run->st_plan);
        if (rc) return rc;
    }
    const TeamPlan& lp = run->ll_plan;
    const int Kp = lp.n_chunks * lp.TB;
    team_prepare_kernel<<<64, 256, 0, st>>>(log_alpha, mu, Sigma, noise == NOISE_SHARED ? S : nullptr, D, K, Kp, geo,
                                            ws.team_sig, ws.team_mu, ws.team_cj, ws.nonpd);
    CUDA_TRY(cudaGetLastError());

    auto fill = [&](const TeamPlan& pl, bool stats, bool has_g) {
        team::TeamParams p;
        p.X = X; p.S = S; p.g = stats ? g : nullptr;
        p.sigT = ws.team_sig; p.muT = ws.team_mu; p.cj = ws.team_cj;
        p.ll = ll; p.out_ll = nullptr;
        p.part = ws.part;
        p.part_ll = stats ? ws.part_ll + 2 * (size_t)max_grid(D, K, dev.sms) : ws.part_ll;
        p.nonpd = ws.nonpd;
        p.N = N; p.n_tiles = pl.n_tiles;
        p.K = K; p.TB = pl.TB; p.n_chunks = pl.n_chunks; p.n_slices = pl.n_slices; p.PT = pl.PT; p.NST = pl.NST;
        p.noise = noise; p.use_bulk = pl.use_bulk;
        p.resp_floor = resp_floor;
        const team::TeamStage SL = team::team_stage(D, noise, pl.PT, has_g, stats);
        p.n_s = noise_floats_per_point(D, noise);
        p.off_X = SL.off_X; p.off_g = SL.off_g; p.off_ll = SL.off_ll; p.stage_bytes = SL.bytes;
        p.tmem_cols = pl.tmem_cols;
        return p;
    };
    const int tslot = timing_begin(st);
    {
        team::TeamParams p = fill(lp, false, false);
        void* args[] = {&p};
        CUDA_TRY(cudaLaunchKernel(lp.kernel, dim3(lp.grid), dim3(lp.TPB), args, lp.smem, st));
    }
    if (mode == MODE_STATS) {
        const TeamPlan& sp = run->st_plan;
        team::TeamParams p = fill(sp, true, g != nullptr);
        void* args[] = {&p};
        CUDA_TRY(cudaLaunchKernel(sp.kernel, dim3(sp.grid), dim3(sp.TPB), args, sp.smem, st));
    }
    timing_end(tslot, st);
    return 0;
}

int launch_team_finalize(const TeamRun& run, const DeviceInfo& dev, const Workspace& ws, int D, int K, float* out_stats,
                         int beta, cudaStream_t st) {
    const TeamPlan& sp = run.st_plan;
    team_finalize_kernel<<<K + 1, 128, 0, st>>>(ws.part, ws.part_ll, run.ll_plan.grid,
                                                ws.part_ll + 2 * (size_t)max_grid(D, K, dev.sms), sp.grid, ws.nonpd, D, K,
                                                sp.TB, sp.n_chunks, sp.n_slices, out_stats, beta);
    return (int)cudaGetLastError();
}

// Runs prepare + pair kernel; leaves partials in the workspace.  Returns the plan.
int run_pair(int mode, const float* X, const float* S, int noise, int64_t N, int D, int K, const float* log_alpha,
             const float* mu, const float* Sigma, const float* g, float resp_floor, float* out_ll, void* workspace,
             size_t workspace_bytes, cudaStream_t st, Plan* plan_out, Workspace* ws_out) {
    DeviceInfo dev;
    int rc = device_info(&dev);
    if (rc) return rc;
    if (!workspace) return DGMM_ERR_WORKSPACE;
    Workspace ws = carve(workspace, D, K, dev.sms);
    if (ws.bytes > workspace_bytes) return DGMM_ERR_WORKSPACE;
    Plan plan;
    rc = make_plan(dev, mode, D, K, N, noise, g != nullptr, X, S, g, &plan);
    if (rc) return rc;

    const int GS = plan.GS;
    prepare_kernel<<<8, 256, 0, st>>>(log_alpha, mu, Sigma, noise == NOISE_SHARED ? S : nullptr, D, K, GS,
                                      ws.sig_cols, ws.mu_cols, ws.cj, ws.nonpd);
    CUDA_TRY(cudaGetLastError());

    PairParams p;
    p.X = X; p.S = S; p.g = g;
    p.sig_cols = ws.sig_cols; p.mu_cols = ws.mu_cols; p.cj = ws.cj;
    p.out_ll = out_ll; p.part = ws.part; p.part_ll = ws.part_ll; p.nonpd = ws.nonpd;
    p.N = N; p.n_tiles = plan.n_tiles;
    p.K = K; p.GS = GS; p.NSLOT = plan.NSLOT; p.PT = plan.PT; p.NST = plan.NST; p.noise = noise;
    p.use_bulk = plan.use_bulk;
    p.resp_floor = resp_floor;
    {
        const StageLayout SL = stage_layout(D, noise, plan.PT, g != nullptr);
        p.n_s = noise_floats_per_point(D, noise);
        p.off_X = SL.off_X; p.off_g = SL.off_g; p.stage_bytes = SL.bytes;
        p.rounds = plan.PT / plan.NSLOT;
        p.tmem_cols = plan.tmem_cols;
    }
    void* args[] = {&p};
    const int tslot = timing_begin(st);
    CUDA_TRY(cudaLaunchKernel(plan.kernel, dim3(plan.grid), dim3(plan.TPB), args, plan.smem, st));
    timing_end(tslot, st);
    *plan_out = plan;
    *ws_out = ws;
    return 0;
}

// Log-likelihood pass on whichever kernel family serves (D, K); *nblk = per-block (sum ll) partials written.
int loglik_pass_1(const float* X, const float* S, int noise, int64_t N, int D, int K, const float* log_alpha,
                  const float* mu, const float* Sigma, float* out_ll, void* workspace, size_t workspace_bytes,
                  cudaStream_t st, Workspace* ws_out, int* nblk) {
    if (use_team_path(D, K, MODE_LOGLIK, noise)) {
        DeviceInfo dev;
        int rc = device_info(&dev);
        if (rc) return rc;
        if (!workspace) return DGMM_ERR_WORKSPACE;
        const Workspace ws = carve(workspace, D, K, dev.sms);
        if (ws.bytes > workspace_bytes) return DGMM_ERR_WORKSPACE;
        TeamRun run;
        rc = run_team(MODE_LOGLIK, X, S, noise, N, D, K, log_alpha, mu, Sigma, nullptr, 0.f, out_ll, dev, ws, st, &run);
        if (rc) return rc;
        *ws_out = ws;
        *nblk = run.ll_plan.grid;
        return 0;
    }
    Plan plan;
    int rc = run_pair(MODE_LOGLIK, X, S, noise, N, D, K, log_alpha, mu, Sigma, nullptr, 0.f, out_ll, workspace,
                      workspace_bytes, st, &plan, ws_out);
    if (rc) return rc;
    *nblk = plan.grid;
    return 0;
}

// Fused statistics pass + finalise into out_stats (beta = 1 accumulates) on whichever family serves (D, K).
// The cooperative path needs ll_scratch [N] (it may be the caller's out_ll).
int stats_pass_1(const float* X, const float* S, int noise, int64_t N, int D, int K, const float* log_alpha,
                 const float* mu, const float* Sigma, const float* g, float resp_floor, float* out_ll, float* out_stats,
                 int beta, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (use_team_path(D, K, MODE_STATS, noise)) {
        if (!out_ll) return DGMM_ERR_LL_REQUIRED;
        DeviceInfo dev;
        int rc = device_info(&dev);
        if (rc) return rc;
        if (!workspace) return DGMM_ERR_WORKSPACE;
        const Workspace ws = carve(workspace, D, K, dev.sms);
        if (ws.bytes > workspace_bytes) return DGMM_ERR_WORKSPACE;
        TeamRun run;
        rc = run_team(MODE_STATS, X, S, noise, N, D, K, log_alpha, mu, Sigma, g, resp_floor, out_ll, dev, ws, st, &run);
        if (rc) return rc;
        return launch_team_finalize(run, dev, ws, D, K, out_stats, beta, st);
    }
    Plan plan;
    Workspace ws;
    int rc = run_pair(MODE_STATS, X, S, noise, N, D, K, log_alpha, mu, Sigma, g, resp_floor, out_ll, workspace,
                      workspace_bytes, st, &plan, &ws);
    if (rc) return rc;
    const bool cols = (plan.accp == ST_GLOBAL || plan.accp == ST_TMEM);
    finalize_stats_kernel<<<ns_of(D) * ((K + 31) / 32) + 1, dim3(32, kFinY), 0, st>>>(
        ws.part, ws.part_ll, ws.nonpd, plan.grid, D, K, plan.GS, cols ? plan.TPB : plan.GS, cols ? plan.NSLOT : 1,
        out_stats, beta);
    return (int)cudaGetLastError();
}


// ---------------------------------------------------------------------------------------
// K > 256: component chunks
// ---------------------------------------------------------------------------------------
// "Threads are components" caps one launch at 256 components.  Larger mixtures (the reference's Gaia runs use
// K = 512: experiments/gmm/gaia/variable_k/fit_em.sh:18) are served in chunks of <= 256 components:
//   pass 1  log-likelihood kernel per chunk c            -> ll_c[i] = log sum_{j in c} alpha_j N(.)
//   combine ll_i = logsumexp_c ll_c[i],  g_c[i] = g_i exp(ll_c[i] - ll_i)
//   pass 2  statistics kernel per chunk with weights g_c -> omega_ij = g_c[i] r^c_ij = g_i r_ij  (exact: the
//           chunk-local responsibility times the chunk's share of the point is the global responsibility)
// The chunk statistics are scattered into the caller's packed buffer, the tail comes from the combine kernel.
// Cost: one extra log-likelihood pass (F_ll / F_em = 1/3 more flops than a single-launch E-step).  A
// responsibility floor (Online EM's 10 eps) is applied as g_c[i] * floor, i.e. scaled by the chunk's share.
// Scratch: (n_chunks + 1) N floats from the stream-ordered allocator (the only allocation the library makes
// on the device-pointer API, and only for chunked mixtures).  D <= 8 runs K <= 512 as ONE launch of 512-thread
// groups (dgmm_tables.cuh: pair_wide512) and chunks above that.
inline int single_launch_k(int D) { return D <= kPairWideMaxD ? kPairWideThreads : kPairMaxThreads; }
struct KChunks {
    int n, Kc;
};
inline KChunks k_chunks(int K, int D) {
    KChunks c;
    const int cap = single_launch_k(D);
    c.n = (K + cap - 1) / cap;
    c.Kc = (((K + c.n - 1) / c.n) + 31) & ~31;
    return c;
}
constexpr int kCombineGrid = 1024;

struct ChunkSpace {
    float* tmp_stats;     // packed statistics of one chunk
    double* part_ll;      // [kCombineGrid][2] partial (sum ll, sum g) of the combine kernel
    int* nonpd_acc;       // non-PD pairs over all chunk launches
    size_t inner_off;     // the single-chunk workspace starts here
    size_t bytes;
};
ChunkSpace chunk_space(void* base, int D, int Kc, size_t inner_bytes) {
    ChunkSpace c;
    char* b = (char*)base;
    size_t off = 0;
    c.tmp_stats = (float*)(b + off); off += up256((size_t)dgmm_stats_len(D, Kc) * 4);
    c.part_ll = (double*)(b + off);  off += up256((size_t)kCombineGrid * 2 * 8);
    c.nonpd_acc = (int*)(b + off);   off += 256;
    c.inner_off = off;
    c.bytes = off + inner_bytes;
    return c;
}

__global__ void nonpd_accum_kernel(const int* __restrict__ nonpd, int* __restrict__ acc, int reset) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *acc = (reset ? 0 : *acc) + (nonpd ? *nonpd : 0);
}

// llc [n][N]: in = chunk log-likelihoods, out = chunk weights g_i exp(ll_c - ll_i) (when want_w).
__global__ void __launch_bounds__(256) chunk_combine_kernel(float* __restrict__ llc, int n, long long N,
                                                            const float* __restrict__ g, float* __restrict__ out_ll,
                                                            int want_w, double* __restrict__ part) {
    __shared__ double sh[2][8];
    double ll_acc = 0.0, g_acc = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        float m = -INFINITY;
        bool nan = false;
        for (int c = 0; c < n; ++c) {
            const float v = llc[(size_t)c * N + i];
            nan = nan || !(v == v);
            m = fmaxf(m, v);
        }
        float ll;
        if (nan) {
            ll = -INFINITY;
        } else if (m == -INFINITY) {
            ll = -INFINITY;
        } else {
            float s = 0.f;
            for (int c = 0; c < n; ++c) s += expf(llc[(size_t)c * N + i] - m);
            ll = m + logf(s);
        }
        const float gi = g ? g[i] : 1.f;
        if (want_w) {
            for (int c = 0; c < n; ++c) {
                const float v = llc[(size_t)c * N + i];
                llc[(size_t)c * N + i] = (ll > -INFINITY) ? gi * expf(v - ll) : 0.f;
            }
        }
        if (out_ll) out_ll[i] = ll;
        ll_acc += (double)ll;
        g_acc += (double)gi;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ll_acc += __shfl_xor_sync(0xffffffffu, ll_acc, o);
        g_acc += __shfl_xor_sync(0xffffffffu, g_acc, o);
    }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = ll_acc; sh[1][threadIdx.x >> 5] = g_acc; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < 8; ++w) { a += sh[0][w]; b += sh[1][w]; }
        part[2 * blockIdx.x] = a;
        part[2 * blockIdx.x + 1] = b;
    }
}

// Chunk statistics (packed for Kc components) -> components [c0, c0 + Kc) of the caller's packed buffer.
__global__ void scatter_stats_kernel(const float* __restrict__ tmp, int Kc, int c0, int K, int D, float* __restrict__ out,
                                     int beta) {
    const int per = 1 + D + D * D;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < Kc * per; idx += gridDim.x * blockDim.x) {
        int src, dst;
        if (idx < Kc) {
            src = idx; dst = c0 + idx;
        } else if (idx < Kc + Kc * D) {
            const int r = idx - Kc;
            src = idx; dst = K + c0 * D + r;
        } else {
            const int r = idx - Kc - Kc * D;
            src = idx; dst = K + K * D + c0 * D * D + r;
        }
        out[dst] = beta ? out[dst] + tmp[src] : tmp[src];
    }
}

// Tail of the packed buffer from the combine kernel's partial sums (fixed order: deterministic).
__global__ void chunk_tail_kernel(const double* __restrict__ part, int nblk, const int* __restrict__ nonpd_acc,
                                  float* __restrict__ tail, int beta) {
    double ll = 0.0, sg = 0.0;
    for (int b = threadIdx.x; b < nblk; b += 32) { ll += part[2 * b]; sg += part[2 * b + 1]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ll += __shfl_xor_sync(0xffffffffu, ll, o);
        sg += __shfl_xor_sync(0xffffffffu, sg, o);
    }
    if (threadIdx.x == 0) {
        if (beta) ll += (double)tail[0] + (double)tail[1];
        const float hi = (float)ll;
        tail[0] = hi;
        tail[1] = isfinite(hi) ? (float)(ll - (double)hi) : 0.f;
        tail[2] = beta ? tail[2] + (float)sg : (float)sg;
        tail[3] = beta ? tail[3] + (float)(*nonpd_acc) : (float)(*nonpd_acc);
    }
}

// Chunked log-likelihood pass: fills llc [n][N] (chunk values, or chunk weights when want_w), out_ll (optional),
// the combine partials and the non-PD accumulator.
int chunked_loglik(const float* X, const float* S, int noise, int64_t N, int D, int K, const float* log_alpha,
                   const float* mu, const float* Sigma, const float* g, float* out_ll, int want_w, float* llc,
                   const ChunkSpace& cs, void* inner_ws, size_t inner_bytes, cudaStream_t st, int* nblk) {
    const KChunks kc = k_chunks(K, D);
    for (int c = 0; c < kc.n; ++c) {
        const int c0 = c * kc.Kc, k = std::min(kc.Kc, K - c0);
        Workspace ws;
        int nb = 0;
        int rc = loglik_pass_1(X, S, noise, N, D, k, log_alpha + c0, mu + (size_t)c0 * D, Sigma + (size_t)c0 * D * D,
                               llc + (size_t)c * N, inner_ws, inner_bytes, st, &ws, &nb);
        if (rc) return rc;
        nonpd_accum_kernel<<<1, 32, 0, st>>>(ws.nonpd, cs.nonpd_acc, c == 0);
        CUDA_TRY(cudaGetLastError());
    }
    const int grid = (int)std::min<long long>(kCombineGrid, std::max<long long>(1, (N + 255) / 256));
    chunk_combine_kernel<<<grid, 256, 0, st>>>(llc, kc.n, N, g, out_ll, want_w, cs.part_ll);
    CUDA_TRY(cudaGetLastError());
    *nblk = grid;
    return 0;
}

int loglik_pass(const float* X, const float* S, int noise, int64_t N, int D, int K, const float* log_alpha,
                const float* mu, const float* Sigma, float* out_ll, void* workspace, size_t workspace_bytes,
                cudaStream_t st, Workspace* ws_out, int* nblk) {
    if (K <= single_launch_k(D))
        return loglik_pass_1(X, S, noise, N, D, K, log_alpha, mu, Sigma, out_ll, workspace, workspace_bytes, st, ws_out, nblk);
    DeviceInfo dev;
    int rc = device_info(&dev);
    if (rc) return rc;
    if (!workspace) return DGMM_ERR_WORKSPACE;
    const KChunks kc = k_chunks(K, D);
    const size_t inner_bytes = carve(nullptr, D, kc.Kc, dev.sms).bytes;
    const ChunkSpace cs = chunk_space(workspace, D, kc.Kc, inner_bytes);
    if (cs.bytes > workspace_bytes) return DGMM_ERR_WORKSPACE;
    float* llc = nullptr;
    CUDA_TRY(cudaMallocAsync((void**)&llc, (size_t)kc.n * std::max<int64_t>(N, 1) * 4, st));
    rc = chunked_loglik(X, S, noise, N, D, K, log_alpha, mu, Sigma, nullptr, out_ll, 0, llc, cs,
                        (char*)workspace + cs.inner_off, inner_bytes, st, nblk);
    cudaFreeAsync(llc, st);
    if (rc) return rc;
    // what the callers' finalize_ll_kernel reads: per-block (sum ll) partials and the non-PD count
    Workspace ws = carve((char*)workspace + cs.inner_off, D, kc.Kc, dev.sms);
    ws.part_ll = cs.part_ll;
    ws.nonpd = cs.nonpd_acc;
    *ws_out = ws;
    return 0;
}

int stats_pass(const float* X, const float* S, int noise, int64_t N, int D, int K, const float* log_alpha,
               const float* mu, const float* Sigma, const float* g, float resp_floor, float* out_ll, float* out_stats,
               int beta, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (K <= single_launch_k(D))
        return stats_pass_1(X, S, noise, N, D, K, log_alpha, mu, Sigma, g, resp_floor, out_ll, out_stats, beta, workspace,
                            workspace_bytes, st);
    DeviceInfo dev;
    int rc = device_info(&dev);
    if (rc) return rc;
    if (!workspace) return DGMM_ERR_WORKSPACE;
    const KChunks kc = k_chunks(K, D);
    const size_t inner_bytes = carve(nullptr, D, kc.Kc, dev.sms).bytes;
    const ChunkSpace cs = chunk_space(workspace, D, kc.Kc, inner_bytes);
    if (cs.bytes > workspace_bytes) return DGMM_ERR_WORKSPACE;
    void* inner_ws = (char*)workspace + cs.inner_off;
    const int64_t Np = std::max<int64_t>(N, 1);
    float* llc = nullptr;
    CUDA_TRY(cudaMallocAsync((void**)&llc, (size_t)(kc.n + 1) * Np * 4, st));
    float* ll_scratch = llc + (size_t)kc.n * Np;      // ll carrier of the cooperative path (D > 16)
    int nblk = 0;
    rc = chunked_loglik(X, S, noise, N, D, K, log_alpha, mu, Sigma, g, out_ll, 1, llc, cs, inner_ws, inner_bytes, st, &nblk);
    for (int c = 0; c < kc.n && rc == 0; ++c) {
        const int c0 = c * kc.Kc, k = std::min(kc.Kc, K - c0);
        rc = stats_pass_1(X, S, noise, N, D, k, log_alpha + c0, mu + (size_t)c0 * D, Sigma + (size_t)c0 * D * D,
                          llc + (size_t)c * Np, resp_floor, ll_scratch, cs.tmp_stats, 0, inner_ws, inner_bytes, st);
        if (rc) break;
        scatter_stats_kernel<<<64, 256, 0, st>>>(cs.tmp_stats, k, c0, K, D, out_stats, beta);
        if (cudaGetLastError() != cudaSuccess) rc = DGMM_ERR_BAD_ARG;
    }
    cudaFreeAsync(llc, st);
    if (rc) return rc;
    chunk_tail_kernel<<<1, 32, 0, st>>>(cs.part_ll, nblk, cs.nonpd_acc, out_stats + (size_t)K * (1 + D + D * D), beta);
    return (int)cudaGetLastError();
}

}  // namespace

// =======================================================================================
// C ABI
// =======================================================================================
extern "C" {

int dgmm_abi_version(void) { return DGMM_ABI_VERSION; }

int dgmm_supported_dims(int* list, int* n) {
    if (!n) return DGMM_ERR_BAD_ARG;
    if (list)
        for (int d = 1; d <= DGMM_MAX_D; ++d) list[d - 1] = d;
    *n = DGMM_MAX_D;
    return 0;
}

size_t dgmm_workspace_bytes(int D, int K) {
    if (D < 1 || D > DGMM_MAX_D || K < 1 || K > DGMM_MAX_K) return 0;
    DeviceInfo dev;
    if (device_info(&dev)) return 0;
    if (K > single_launch_k(D)) {
        const KChunks kc = k_chunks(K, D);
        return chunk_space(nullptr, D, kc.Kc, carve(nullptr, D, kc.Kc, dev.sms).bytes).bytes;
    }
    return carve(nullptr, D, K, dev.sms).bytes;
}

int dgmm_stats_needs_ll(int D, int K, int noise_layout) {
    if (D < 1 || D > DGMM_MAX_D || K < 1 || K > DGMM_MAX_K) return 0;
    if (K > single_launch_k(D)) return 0;    // the chunked path carries ll_i in its own scratch
    return use_team_path(D, K, MODE_STATS, noise_layout) ? 1 : 0;
}

// Diagnostics / tuning hook (not part of the drop-in surface): -1 = automatic.
int dgmm_debug_set_tuning(int sigp, int accp, int nslot, int pt, int nst, int blocks_per_sm, int rc, int no_bulk) {
    g_tuning.sigp = sigp; g_tuning.accp = accp; g_tuning.nslot = nslot; g_tuning.pt = pt;
    g_tuning.nst = nst; g_tuning.blocks_per_sm = blocks_per_sm; g_tuning.rc = rc; g_tuning.no_bulk = no_bulk;
    return 0;
}

// Diagnostics: switch per-launch timing of the pair kernel on/off (resets the record).
int dgmm_debug_pair_timing(int enable) {
    g_timing.on = enable != 0;
    g_timing.count = 0;
    return 0;
}
// Diagnostics: device times (ms) of the pair-kernel launches recorded since timing was switched
// on (at most the last 512).  Synchronises on the recorded events.  *n in = capacity, out = count.
int dgmm_debug_pair_times(float* out_ms, int* n) {
    if (!out_ms || !n) return DGMM_ERR_BAD_ARG;
    const int have = g_timing.count < PairTiming::kRing ? g_timing.count : PairTiming::kRing;
    const int m = have < *n ? have : *n;
    for (int i = 0; i < m; ++i) {
        const int slot = (g_timing.count - m + i) % PairTiming::kRing;
        CUDA_TRY(cudaEventSynchronize(g_timing.b[slot]));
        CUDA_TRY(cudaEventElapsedTime(&out_ms[i], g_timing.a[slot], g_timing.b[slot]));
    }
    *n = m;
    return 0;
}

// Diagnostics: the launch plan the library would use (for bench.py / DESIGN.md reporting).
// out[0..9] = sigp, accp, rc, GS, NSLOT, TPB, PT, NST, grid, smem bytes ; out[10] = use_bulk
int dgmm_debug_get_plan(int mode, int D, int K, int64_t N, int noise, int has_g, int* out) {
    DeviceInfo dev;
    int rc = device_info(&dev);
    if (rc) return rc;
    if (D < 1 || D > DGMM_MAX_D || K < 1 || K > DGMM_MAX_K) return DGMM_ERR_BAD_ARG;
    if (K > single_launch_k(D)) K = k_chunks(K, D).Kc;      // component chunks: the plan of one chunk launch
    if (use_team_path(D, K, mode, noise)) {
        // cooperative path: sigp = 3 marks it; GS = teams per block, NSLOT = component chunks
        TeamPlan t;
        rc = make_team_plan(dev, mode, D, K, N, noise, has_g != 0, nullptr, nullptr, nullptr, nullptr, &t);
        if (rc) return rc;
        out[0] = 3; out[1] = ST_TMEM; out[2] = 3; out[3] = t.TB; out[4] = t.n_chunks; out[5] = t.TPB;
        out[6] = t.PT; out[7] = t.NST; out[8] = t.grid; out[9] = (int)t.smem; out[10] = t.use_bulk;
        return 0;
    }
    Plan p;
    rc = make_plan(dev, mode, D, K, N, noise, has_g != 0, nullptr, nullptr, nullptr, &p);
    if (rc) return rc;
    out[0] = p.sigp; out[1] = p.accp; out[2] = p.rc; out[3] = p.GS; out[4] = p.NSLOT; out[5] = p.TPB;
    out[6] = p.PT; out[7] = p.NST; out[8] = p.grid; out[9] = (int)p.smem; out[10] = p.use_bulk;
    return 0;
}

int dgmm_loglik_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                    const float* log_alpha, const float* mu, const float* Sigma, float* out_ll, double* out_sum,
                    int* status, void* workspace, size_t workspace_bytes, void* stream) {
    int rc = check_common(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma);
    if (rc) return rc;
    if (!out_ll) return DGMM_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    Workspace ws;
    int nblk = 0;
    rc = loglik_pass(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma, out_ll, workspace, workspace_bytes, st, &ws, &nblk);
    if (rc) return rc;
    if (out_sum || status) {
        finalize_ll_kernel<<<1, 32, 0, st>>>(ws.part_ll, ws.nonpd, nblk, out_sum, status, 0);
        CUDA_TRY(cudaGetLastError());
    }
    return 0;
}

int dgmm_estep_stats_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                         const float* log_alpha, const float* mu, const float* Sigma, const float* g,
                         float resp_floor, float* out_ll, float* out_stats, void* workspace,
                         size_t workspace_bytes, void* stream) {
    int rc = check_common(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma);
    if (rc) return rc;
    if (!out_stats) return DGMM_ERR_BAD_ARG;
    return stats_pass(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma, g, resp_floor, out_ll, out_stats, 0, workspace,
                      workspace_bytes, (cudaStream_t)stream);
}

int dgmm_estep_full_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                        const float* log_alpha, const float* mu, const float* Sigma, float* out_log_r, float* out_b,
                        float* out_B, float* out_ll, int* status, void* workspace, size_t workspace_bytes,
                        void* stream) {
    int rc = check_common(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma);
    if (rc) return rc;
    if (!out_ll) return DGMM_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    Workspace ws;
    int nblk = 0;
    rc = loglik_pass(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma, out_ll, workspace, workspace_bytes, st, &ws, &nblk);
    if (rc) return rc;
    if (N > 0 && (out_log_r || out_b || out_B)) {
        FullParams f;
        f.X = X; f.S = S; f.log_alpha = log_alpha; f.mu = mu; f.Sigma = Sigma; f.ll = out_ll;
        f.out_log_r = out_log_r; f.out_b = out_b; f.out_B = out_B; f.nonpd = ws.nonpd;
        f.N = N; f.K = K; f.noise = noise_layout;
        const void* k = kDims[D].full();
        // one block per (tile of 128 points, component)
        const long long blocks = ((long long)N + 127) / 128 * K;
        if (blocks > 0x7fffffffLL) return DGMM_ERR_BAD_ARG;
        void* args[] = {&f};
        const size_t dyn = out_B ? full_dyn_smem_bytes(D) : 0;
        if (dyn > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        CUDA_TRY(cudaLaunchKernel(k, dim3((unsigned)blocks), dim3(128), args, dyn, st));
    }
    if (status) {
        finalize_ll_kernel<<<1, 32, 0, st>>>(ws.part_ll, ws.nonpd, nblk, nullptr, status, 0);
        CUDA_TRY(cudaGetLastError());
    }
    return 0;
}

int dgmm_posterior_sample_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                              const float* log_alpha, const float* mu, const float* Sigma, int M, const float* u,
                              const float* z, float* out_samples, int32_t* out_comp, float* scratch_log_r,
                              float* out_ll, int* status, void* workspace, size_t workspace_bytes, void* stream) {
    if (M < 1 || !u || !z || !out_samples || !scratch_log_r) return DGMM_ERR_BAD_ARG;
    // ll_i and log r_ij [N,K] (no b, no B), then one thread per (point, draw)
    int rc = dgmm_estep_full_f32(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma, scratch_log_r, nullptr, nullptr,
                                 out_ll, nullptr, workspace, workspace_bytes, stream);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    DeviceInfo dev;
    rc = device_info(&dev);
    if (rc) return rc;
    const Workspace ws = carve(workspace, D, K, dev.sms);
    if (N > 0) {
        SampleParams sp;
        sp.X = X; sp.S = S; sp.mu = mu; sp.Sigma = Sigma; sp.log_r = scratch_log_r; sp.u = u; sp.z = z;
        sp.out = out_samples; sp.out_comp = out_comp; sp.nonpd = ws.nonpd; sp.N = N; sp.K = K; sp.M = M;
        sp.noise = noise_layout;
        const long long draws = (long long)N * M;
        const long long blocks = (draws + 127) / 128;
        if (blocks > 0x7fffffffLL) return DGMM_ERR_BAD_ARG;
        void* args[] = {&sp};
        CUDA_TRY(cudaLaunchKernel(kDims[D].sample(), dim3((unsigned)blocks), dim3(128), args, 0, st));
    }
    if (status) {
        finalize_ll_kernel<<<1, 32, 0, st>>>(ws.part_ll, ws.nonpd, 0, nullptr, status, 0);
        CUDA_TRY(cudaGetLastError());
    }
    return 0;
}

int dgmm_posterior_logprob_f32(const float* X, const float* S, int noise_layout, int64_t N, int D, int K,
                               const float* log_alpha, const float* mu, const float* Sigma, int M, const float* V,
                               float* out_logprob, float* scratch_log_r, float* out_ll, int* status, void* workspace,
                               size_t workspace_bytes, void* stream) {
    if (M < 1 || !V || !out_logprob || !scratch_log_r) return DGMM_ERR_BAD_ARG;
    int rc = dgmm_estep_full_f32(X, S, noise_layout, N, D, K, log_alpha, mu, Sigma, scratch_log_r, nullptr, nullptr,
                                 out_ll, nullptr, workspace, workspace_bytes, stream);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    DeviceInfo dev;
    rc = device_info(&dev);
    if (rc) return rc;
    const Workspace ws = carve(workspace, D, K, dev.sms);
    if (N > 0) {
        SampleParams sp;
        sp.X = X; sp.S = S; sp.mu = mu; sp.Sigma = Sigma; sp.log_r = scratch_log_r; sp.u = nullptr; sp.z = V;
        sp.out = out_logprob; sp.out_comp = nullptr; sp.nonpd = ws.nonpd; sp.N = N; sp.K = K; sp.M = M;
        sp.noise = noise_layout;
        const long long q = (long long)N * M;
        const long long blocks = (q + 127) / 128;
        if (blocks > 0x7fffffffLL) return DGMM_ERR_BAD_ARG;
        void* args[] = {&sp};
        CUDA_TRY(cudaLaunchKernel(kDims[D].post_logprob(), dim3((unsigned)blocks), dim3(128), args, 0, st));
    }
    if (status) {
        finalize_ll_kernel<<<1, 32, 0, st>>>(ws.part_ll, ws.nonpd, 0, nullptr, status, 0);
        CUDA_TRY(cudaGetLastError());
    }
    return 0;
}

int dgmm_mstep_update_f32(const float* stats, double n_total, float w, int D, int K, float* mu, float* Sigma,
                          float* alpha, float* log_alpha, void* stream) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1) return DGMM_ERR_UNSUPPORTED_K;
    if (!stats || !mu || !Sigma || !alpha || !log_alpha) return DGMM_ERR_BAD_ARG;
    const size_t sh = (size_t)(3 * D * D + D) * sizeof(double);
    mstep_kernel<<<K, D * D, sh, (cudaStream_t)stream>>>(stats, n_total, (double)w, D, K, mu, Sigma, alpha, log_alpha);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int dgmm_online_mstep_f32(const float* stats, float resp_floor, float step_size, float w, double global_batch,
                          int D, int K, float* mu, float* Sigma, float* sum_resps, float* sum_cond_means,
                          float* alpha, float* log_alpha, int* crash_flag, void* stream) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1) return DGMM_ERR_UNSUPPORTED_K;
    if (!stats || !mu || !Sigma || !sum_resps || !sum_cond_means || !alpha || !log_alpha || !(global_batch > 0))
        return DGMM_ERR_BAD_ARG;
    const size_t sh = (size_t)(3 * D * D + D) * sizeof(double);
    online_mstep_kernel<<<K, D * D, sh, (cudaStream_t)stream>>>(stats, (double)resp_floor, (double)step_size,
                                                                 (double)w, global_batch, D, K, mu, Sigma, sum_resps,
                                                                 sum_cond_means, alpha, log_alpha, crash_flag);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int dgmm_fp32_probe(float* sink, int iters, double* flops, void* stream) {
    DeviceInfo dev;
    int rc = device_info(&dev);
    if (rc) return rc;
    if (!sink || iters < 1) return DGMM_ERR_BAD_ARG;
    const int blocks = dev.sms * 8, threads = 256;
    fp32_probe_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(sink, iters);
    CUDA_TRY(cudaGetLastError());
    if (flops) *flops = (double)blocks * threads * (double)iters * 16.0 * 8.0 * 2.0;
    return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------
// host-buffer pipeline
// ---------------------------------------------------------------------------------------
struct dgmm_ctx {
    int device, D, K;
    int64_t chunk;
    cudaStream_t copy_st, comp_st;
    cudaEvent_t loaded[2], consumed[2];
    float* dX[2];
    float* dS[2];
    float* d_ll[2];
    float *d_alpha, *d_log_alpha, *d_mu, *d_Sigma, *d_stats;
    double* d_sum;
    double* h_sum;   // pinned, 2 doubles
    float* h_tail;   // pinned, 4 floats
    void* ws;
    size_t ws_bytes;
    // resident table (dgmm_ctx_upload)
    float* rX;
    float* rS;
    float* r_ll;
    int64_t rN;
    int r_layout;
};

// S_full [N,D,D] -> packed lower triangles [N,ND]
__global__ void pack_noise_kernel(const float* __restrict__ S, long long N, int D, float* __restrict__ out) {
    const int ND = D * (D + 1) / 2;
    const long long total = N * ND;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long i = idx / ND;
        const int e = (int)(idx - i * ND);
        int a = 0;
        while ((a + 1) * (a + 2) / 2 <= e) ++a;
        const int b = e - a * (a + 1) / 2;
        out[idx] = S[(size_t)i * D * D + a * D + b];
    }
}

extern "C" {

int dgmm_pack_noise_f32(const float* S_full, int64_t N, int D, float* S_packed, void* stream) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (N < 0 || (N > 0 && (!S_full || !S_packed))) return DGMM_ERR_BAD_ARG;
    if (N == 0) return 0;
    const long long total = (long long)N * (D * (D + 1) / 2);
    const int grid = (int)std::min<long long>((total + 255) / 256, 148 * 16);
    pack_noise_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(S_full, (long long)N, D, S_packed);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// Host packer: row a of the lower triangle is a + 1 contiguous floats in both layouts (one memcpy per row);
// the points are split over the host threads.
int dgmm_pack_noise_host(const float* S_full, int64_t N, int D, float* S_packed) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (N < 0 || (N > 0 && (!S_full || !S_packed))) return DGMM_ERR_BAD_ARG;
    const int ND = D * (D + 1) / 2;
    auto work = [=](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            const float* src = S_full + (size_t)i * D * D;
            float* dst = S_packed + (size_t)i * ND;
            for (int a = 0; a < D; ++a) memcpy(dst + a * (a + 1) / 2, src + a * D, (size_t)(a + 1) * 4);
        }
    };
    unsigned nt = std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > 32) nt = 32;
    if (N < 65536 || nt == 1) { work(0, N); return 0; }
    std::vector<std::thread> th;
    const int64_t per = (N + nt - 1) / nt;
    for (unsigned t = 0; t < nt; ++t) {
        const int64_t lo = (int64_t)t * per, hi = std::min<int64_t>(N, lo + per);
        if (lo < hi) th.emplace_back(work, lo, hi);
    }
    for (auto& t : th) t.join();
    return 0;
}

int dgmm_ctx_create(dgmm_ctx** out, int device, int D, int K, int64_t chunk_points) {
    if (!out || chunk_points < 1) return DGMM_ERR_BAD_ARG;
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1 || K > DGMM_MAX_K) return DGMM_ERR_UNSUPPORTED_K;
    CUDA_TRY(cudaSetDevice(device));
    dgmm_ctx* c = (dgmm_ctx*)calloc(1, sizeof(dgmm_ctx));
    if (!c) return DGMM_ERR_BAD_ARG;
    c->device = device; c->D = D; c->K = K; c->chunk = chunk_points;
    CUDA_TRY(cudaStreamCreateWithFlags(&c->copy_st, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&c->comp_st, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        CUDA_TRY(cudaEventCreateWithFlags(&c->loaded[i], cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&c->consumed[i], cudaEventDisableTiming));
        CUDA_TRY(cudaMalloc(&c->dX[i], (size_t)chunk_points * D * 4));
        CUDA_TRY(cudaMalloc(&c->dS[i], (size_t)chunk_points * D * D * 4));
        CUDA_TRY(cudaMalloc(&c->d_ll[i], (size_t)chunk_points * 4));
    }
    CUDA_TRY(cudaMalloc(&c->d_alpha, (size_t)K * 4));
    CUDA_TRY(cudaMalloc(&c->d_log_alpha, (size_t)K * 4));
    CUDA_TRY(cudaMalloc(&c->d_mu, (size_t)K * D * 4));
    CUDA_TRY(cudaMalloc(&c->d_Sigma, (size_t)K * D * D * 4));
    CUDA_TRY(cudaMalloc(&c->d_stats, dgmm_stats_len(D, K) * 4));
    CUDA_TRY(cudaMalloc(&c->d_sum, 2 * sizeof(double)));
    CUDA_TRY(cudaMallocHost(&c->h_sum, 2 * sizeof(double)));
    CUDA_TRY(cudaMallocHost(&c->h_tail, 4 * sizeof(float)));
    c->ws_bytes = dgmm_workspace_bytes(D, K);
    if (!c->ws_bytes) return DGMM_ERR_NO_DEVICE;
    CUDA_TRY(cudaMalloc(&c->ws, c->ws_bytes));
    *out = c;
    return 0;
}

void dgmm_ctx_destroy(dgmm_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < 2; ++i) {
        if (c->dX[i]) cudaFree(c->dX[i]);
        if (c->dS[i]) cudaFree(c->dS[i]);
        if (c->d_ll[i]) cudaFree(c->d_ll[i]);
        if (c->loaded[i]) cudaEventDestroy(c->loaded[i]);
        if (c->consumed[i]) cudaEventDestroy(c->consumed[i]);
    }
    cudaFree(c->d_alpha); cudaFree(c->d_log_alpha); cudaFree(c->d_mu); cudaFree(c->d_Sigma);
    cudaFree(c->d_stats); cudaFree(c->d_sum); cudaFree(c->ws);
    if (c->rX) cudaFree(c->rX);
    if (c->rS) cudaFree(c->rS);
    if (c->r_ll) cudaFree(c->r_ll);
    cudaFreeHost(c->h_sum); cudaFreeHost(c->h_tail);
    if (c->copy_st) cudaStreamDestroy(c->copy_st);
    if (c->comp_st) cudaStreamDestroy(c->comp_st);
    free(c);
}

}  // extern "C"

static size_t noise_floats(int D, int layout) {
    return layout == DGMM_NOISE_FULL ? (size_t)D * D
           : (layout == DGMM_NOISE_DIAG ? (size_t)D : (layout == DGMM_NOISE_PACKED ? (size_t)D * (D + 1) / 2 : 0));
}

// Streams the chunks of (X, S) through the two device buffers: copy stream uploads chunk c+1
// while the compute stream works on chunk c.  `body(chunk_index, n, dX, dS)` enqueues the
// kernels of one chunk on the compute stream.
template <class Body>
static int stream_chunks(dgmm_ctx* c, const float* X, const float* S, int layout, int64_t N, Body body) {
    const int D = c->D;
    const size_t nf = noise_floats(D, layout);
    const float* dShared = nullptr;
    if (layout == DGMM_NOISE_SHARED) {
        CUDA_TRY(cudaMemcpyAsync(c->dS[0], S, (size_t)D * D * 4, cudaMemcpyHostToDevice, c->comp_st));
        dShared = c->dS[0];
    }
    int64_t done = 0;
    int idx = 0;
    while (done < N) {
        const int64_t n = std::min<int64_t>(c->chunk, N - done);
        const int b = idx & 1;
        if (idx >= 2) CUDA_TRY(cudaStreamWaitEvent(c->copy_st, c->consumed[b], 0));
        CUDA_TRY(cudaMemcpyAsync(c->dX[b], X + (size_t)done * D, (size_t)n * D * 4, cudaMemcpyHostToDevice, c->copy_st));
        if (nf)
            CUDA_TRY(cudaMemcpyAsync(c->dS[b], S + (size_t)done * nf, (size_t)n * nf * 4, cudaMemcpyHostToDevice,
                                     c->copy_st));
        CUDA_TRY(cudaEventRecord(c->loaded[b], c->copy_st));
        CUDA_TRY(cudaStreamWaitEvent(c->comp_st, c->loaded[b], 0));
        int rc = body(idx, done, n, c->dX[b], layout == DGMM_NOISE_SHARED ? dShared : c->dS[b]);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->consumed[b], c->comp_st));
        done += n;
        ++idx;
    }
    return 0;
}

extern "C" {

static int upload_params(dgmm_ctx* c, const float* alpha, const float* mu, const float* Sigma) {
    const int D = c->D, K = c->K;
    cudaStream_t st = c->comp_st;
    CUDA_TRY(cudaMemcpyAsync(c->d_alpha, alpha, (size_t)K * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(c->d_mu, mu, (size_t)K * D * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(c->d_Sigma, Sigma, (size_t)K * D * D * 4, cudaMemcpyHostToDevice, st));
    log_kernel<<<(K + 127) / 128, 128, 0, st>>>(c->d_alpha, c->d_log_alpha, K);
    return (int)cudaGetLastError();
}

// Streams (X, S) through the chunk buffers and leaves the accumulated packed statistics in
// c->d_stats (compute stream, not synchronised).
static int stats_pass_host(dgmm_ctx* c, const float* X, const float* S, int layout, int64_t N, float resp_floor) {
    const int D = c->D, K = c->K;
    cudaStream_t st = c->comp_st;
    CUDA_TRY(cudaMemsetAsync(c->d_stats, 0, dgmm_stats_len(D, K) * 4, st));
    return stream_chunks(c, X, S, layout, N, [&](int idx, int64_t, int64_t n, const float* dX, const float* dS) -> int {
        // (d_ll: scratch of the cooperative path, D > 16; ordered on `st` like everything else of the chunk)
        return stats_pass(dX, dS, layout, n, D, K, c->d_log_alpha, c->d_mu, c->d_Sigma, nullptr, resp_floor,
                          c->d_ll[idx & 1], c->d_stats, 1, c->ws, c->ws_bytes, st);
    });
}

int dgmm_estep_stats_host(dgmm_ctx* c, const float* X, const float* S, int layout, int64_t N, const float* alpha,
                          const float* mu, const float* Sigma, float resp_floor, float* out_stats) {
    if (!c || !alpha || !mu || !Sigma || !out_stats) return DGMM_ERR_BAD_ARG;
    if (layout < 0 || layout > 4) return DGMM_ERR_LAYOUT;
    if (layout == DGMM_NOISE_SHARED && c->chunk < c->D) return DGMM_ERR_BAD_ARG;
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = upload_params(c, alpha, mu, Sigma);
    if (rc) return rc;
    rc = stats_pass_host(c, X, S, layout, N, resp_floor);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(out_stats, c->d_stats, dgmm_stats_len(c->D, c->K) * 4, cudaMemcpyDeviceToHost,
                             c->comp_st));
    CUDA_TRY(cudaStreamSynchronize(c->comp_st));
    return 0;
}

int dgmm_em_iteration_host(dgmm_ctx* c, const float* X, const float* S, int layout, int64_t N, float w, float* alpha,
                           float* mu, float* Sigma, double* ll_sum, int64_t* nonpd) {
    if (!c || !alpha || !mu || !Sigma) return DGMM_ERR_BAD_ARG;
    if (layout < 0 || layout > 4) return DGMM_ERR_LAYOUT;
    if (layout == DGMM_NOISE_SHARED && c->chunk < c->D) return DGMM_ERR_BAD_ARG;
    CUDA_TRY(cudaSetDevice(c->device));
    const int D = c->D, K = c->K;
    cudaStream_t st = c->comp_st;
    int rc = upload_params(c, alpha, mu, Sigma);
    if (rc) return rc;
    rc = stats_pass_host(c, X, S, layout, N, 0.f);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(c->h_tail, c->d_stats + (size_t)K * (1 + D + D * D), 4 * sizeof(float),
                             cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    const bool bad = c->h_tail[3] > 0.f;
    if (ll_sum) *ll_sum = bad ? -INFINITY : (double)c->h_tail[0] + (double)c->h_tail[1];
    if (nonpd) *nonpd = (int64_t)c->h_tail[3];
    if (bad) return 0;  // parameters stay as they were (deconv/gmm/base.py:48-49)
    rc = dgmm_mstep_update_f32(c->d_stats, (double)N, w, D, K, c->d_mu, c->d_Sigma, c->d_alpha, c->d_log_alpha, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(alpha, c->d_alpha, (size_t)K * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(mu, c->d_mu, (size_t)K * D * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(Sigma, c->d_Sigma, (size_t)K * D * D * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return 0;
}

int dgmm_score_host(dgmm_ctx* c, const float* X, const float* S, int layout, int64_t N, const float* alpha,
                    const float* mu, const float* Sigma, float* out_ll, double* ll_sum, int64_t* nonpd) {
    if (!c || !alpha || !mu || !Sigma) return DGMM_ERR_BAD_ARG;
    if (layout < 0 || layout > 4) return DGMM_ERR_LAYOUT;
    CUDA_TRY(cudaSetDevice(c->device));
    const int D = c->D, K = c->K;
    cudaStream_t st = c->comp_st;
    { int rcp = upload_params(c, alpha, mu, Sigma); if (rcp) return rcp; }
    CUDA_TRY(cudaMemsetAsync(c->d_sum, 0, 2 * sizeof(double), st));
    int rc = stream_chunks(c, X, S, layout, N, [&](int idx, int64_t off, int64_t n, const float* dX, const float* dS) -> int {
        float* dll = c->d_ll[idx & 1];  // safe to reuse: its D2H copy two chunks ago is ordered on `st`
        Workspace ws;
        int nblk = 0;
        int r = loglik_pass(dX, dS, layout, n, D, K, c->d_log_alpha, c->d_mu, c->d_Sigma, dll, c->ws, c->ws_bytes, st, &ws,
                            &nblk);
        if (r) return r;
        finalize_ll_kernel<<<1, 32, 0, st>>>(ws.part_ll, ws.nonpd, nblk, c->d_sum, nullptr, 1);
        CUDA_TRY(cudaGetLastError());
        if (out_ll) CUDA_TRY(cudaMemcpyAsync(out_ll + off, dll, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        return 0;
    });
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(c->h_sum, c->d_sum, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    const double total = c->h_sum[0];
    const int64_t bad = (int64_t)c->h_sum[1];
    if (ll_sum) *ll_sum = bad ? -INFINITY : total;
    if (nonpd) *nonpd = bad;
    return 0;
}

// ---- resident table -------------------------------------------------------------------------
int dgmm_ctx_upload(dgmm_ctx* c, const float* X, const float* S, int layout, int64_t N, int pack_full) {
    if (!c || N < 0 || (N > 0 && !X)) return DGMM_ERR_BAD_ARG;
    if (layout < 0 || layout > 4) return DGMM_ERR_LAYOUT;
    if (N > 0 && layout != DGMM_NOISE_NONE && !S) return DGMM_ERR_BAD_ARG;
    CUDA_TRY(cudaSetDevice(c->device));
    const int D = c->D;
    CUDA_TRY(cudaStreamSynchronize(c->comp_st));
    if (c->rX) { cudaFree(c->rX); c->rX = nullptr; }
    if (c->rS) { cudaFree(c->rS); c->rS = nullptr; }
    if (c->r_ll) { cudaFree(c->r_ll); c->r_ll = nullptr; }
    const bool pack = pack_full && layout == DGMM_NOISE_FULL;
    const int res_layout = pack ? DGMM_NOISE_PACKED : layout;
    const size_t nf_res = layout == DGMM_NOISE_SHARED ? 0 : noise_floats(D, res_layout);
    const int64_t Np = std::max<int64_t>(N, 1);
    CUDA_TRY(cudaMalloc(&c->rX, (size_t)Np * D * 4));
    CUDA_TRY(cudaMalloc(&c->r_ll, (size_t)Np * 4));
    if (layout == DGMM_NOISE_SHARED) {
        CUDA_TRY(cudaMalloc(&c->rS, (size_t)D * D * 4));
        CUDA_TRY(cudaMemcpyAsync(c->rS, S, (size_t)D * D * 4, cudaMemcpyHostToDevice, c->comp_st));
    } else if (nf_res) {
        CUDA_TRY(cudaMalloc(&c->rS, (size_t)Np * nf_res * 4));
    }
    c->rN = N;
    c->r_layout = res_layout;
    if (N == 0) return 0;
    if (!pack) {
        CUDA_TRY(cudaMemcpyAsync(c->rX, X, (size_t)N * D * 4, cudaMemcpyHostToDevice, c->comp_st));
        if (nf_res) CUDA_TRY(cudaMemcpyAsync(c->rS, S, (size_t)N * nf_res * 4, cudaMemcpyHostToDevice, c->comp_st));
    } else {
        // full matrices stream through the chunk buffers and are packed on the device as they arrive
        int rc = stream_chunks(c, X, S, layout, N, [&](int, int64_t off, int64_t n, const float* dX, const float* dS) -> int {
            CUDA_TRY(cudaMemcpyAsync(c->rX + (size_t)off * D, dX, (size_t)n * D * 4, cudaMemcpyDeviceToDevice, c->comp_st));
            return dgmm_pack_noise_f32(dS, n, D, c->rS + (size_t)off * nf_res, c->comp_st);
        });
        if (rc) return rc;
    }
    CUDA_TRY(cudaStreamSynchronize(c->comp_st));
    return 0;
}

static int resident_stats(dgmm_ctx* c, float resp_floor) {
    return stats_pass(c->rX, c->rS, c->r_layout, c->rN, c->D, c->K, c->d_log_alpha, c->d_mu, c->d_Sigma, nullptr,
                      resp_floor, c->r_ll, c->d_stats, 0, c->ws, c->ws_bytes, c->comp_st);
}

int dgmm_em_fit_resident(dgmm_ctx* c, int max_iters, float w, double tol, float* alpha, float* mu, float* Sigma,
                         int* iters_done, double* ll_trace, int64_t* nonpd) {
    if (!c || !alpha || !mu || !Sigma || max_iters < 0 || !c->rX) return DGMM_ERR_BAD_ARG;
    CUDA_TRY(cudaSetDevice(c->device));
    const int D = c->D, K = c->K;
    cudaStream_t st = c->comp_st;
    int rc = upload_params(c, alpha, mu, Sigma);
    if (rc) return rc;
    double prev = -INFINITY;
    int done = 0;
    int64_t bad_total = 0;
    for (int it = 0; it < max_iters; ++it) {
        rc = resident_stats(c, 0.f);
        if (rc) return rc;
        CUDA_TRY(cudaMemcpyAsync(c->h_tail, c->d_stats + (size_t)K * (1 + D + D * D), 4 * sizeof(float),
                                 cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));                       // the one host sync of the iteration (base.py:48)
        const bool bad = c->h_tail[3] > 0.f;
        const double ll = bad ? -INFINITY : (double)c->h_tail[0] + (double)c->h_tail[1];
        if (ll_trace) ll_trace[it] = ll;
        bad_total = (int64_t)c->h_tail[3];
        if (bad || ll != ll) break;                                // base.py:48-49: parameters stay as they were
        rc = dgmm_mstep_update_f32(c->d_stats, (double)c->rN, w, D, K, c->d_mu, c->d_Sigma, c->d_alpha, c->d_log_alpha, st);
        if (rc) return rc;
        ++done;
        if (fabs(ll - prev) < tol) break;                          // base.py:52-53
        prev = ll;
    }
    if (iters_done) *iters_done = done;
    if (nonpd) *nonpd = bad_total;
    CUDA_TRY(cudaMemcpyAsync(alpha, c->d_alpha, (size_t)K * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(mu, c->d_mu, (size_t)K * D * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(Sigma, c->d_Sigma, (size_t)K * D * D * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return 0;
}

int dgmm_em_iteration_resident(dgmm_ctx* c, float w, float* alpha, float* mu, float* Sigma, double* ll_sum,
                               int64_t* nonpd) {
    int done = 0;
    double ll = -INFINITY;
    const int rc = dgmm_em_fit_resident(c, 1, w, -1.0, alpha, mu, Sigma, &done, &ll, nonpd);
    if (rc) return rc;
    if (ll_sum) *ll_sum = ll;
    return 0;
}

int dgmm_score_resident(dgmm_ctx* c, const float* alpha, const float* mu, const float* Sigma, float* out_ll,
                        double* ll_sum, int64_t* nonpd) {
    if (!c || !alpha || !mu || !Sigma || !c->rX) return DGMM_ERR_BAD_ARG;
    CUDA_TRY(cudaSetDevice(c->device));
    cudaStream_t st = c->comp_st;
    { int rcp = upload_params(c, alpha, mu, Sigma); if (rcp) return rcp; }
    Workspace ws;
    int nblk = 0;
    int rc = loglik_pass(c->rX, c->rS, c->r_layout, c->rN, c->D, c->K, c->d_log_alpha, c->d_mu, c->d_Sigma, c->r_ll, c->ws,
                         c->ws_bytes, st, &ws, &nblk);
    if (rc) return rc;
    finalize_ll_kernel<<<1, 32, 0, st>>>(ws.part_ll, ws.nonpd, nblk, c->d_sum, nullptr, 0);
    CUDA_TRY(cudaGetLastError());
    if (out_ll && c->rN > 0) CUDA_TRY(cudaMemcpyAsync(out_ll, c->r_ll, (size_t)c->rN * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(c->h_sum, c->d_sum, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    const int64_t bad = (int64_t)c->h_sum[1];
    if (ll_sum) *ll_sum = bad ? -INFINITY : c->h_sum[0];
    if (nonpd) *nonpd = bad;
    return 0;
}

}  // extern "C"
