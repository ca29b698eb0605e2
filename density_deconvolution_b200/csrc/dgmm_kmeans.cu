// dgmm_kmeans.cu -- fused k-means assignment for the initialisation that precedes the hot loop
// in every estimator.  Replaces the reference's [B,K,D] difference / mask
// tensors and its Python loop over centroids:
//   deconv/gmm/util.py:19-24  (Lloyd:     distances, labels, per-cluster means)
//   deconv/gmm/util.py:51-64  (minibatch: distances, labels, bincount, masked shift)
// One pass over the points produces everything both variants need as [K,.] algebra:
//   counts_j = #{i : label_i = j},  sums_j = sum_{label_i = j} x_i,  dist = sum_i min_j |x_i - c_j|.
//
// Kernel (one block = 128 threads, a tile = 256 points, persistent grid striding over tiles):
//   phase 0  the tile's rows of X (contiguous in global memory) are copied, coalesced, into
//            padded shared-memory rows [x_0 .. x_{D-1}, 1, 0..]; the row stride is 4 * odd floats,
//            so the 128-bit row reads of phase 1 (one lane per point) are conflict-free;
//   phase 1  thread t owns points t and t + 128: squared distances to all K centroids (rows
//            broadcast from shared memory as float4), strict-less-than argmin (ties -> lowest
//            index, as torch.min), label -> shared memory (and global, if asked);
//   phase 2  thread e owns one float4 chunk of one centroid's accumulator row [sum x | count]
//            and adds the rows of the tile's points carrying its label, in point order.  The
//            count is the sum of the constant-1 column.  No atomics: every accumulator has
//            one owner and a fixed summation order, so the result is bit-reproducible.
// The per-block partials are summed in double, in block order, by kmeans_finalize_kernel.
// FP32-issue bound: 2*K*D instructions per point in phase 1 against 4*D bytes of input.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <algorithm>

#include "../../include/dgmm_b200.h"

namespace {

#define CUDA_TRY(x)                                  \
    do {                                             \
        cudaError_t e__ = (x);                       \
        if (e__ != cudaSuccess) return (int)e__;     \
    } while (0)

constexpr int kTPB = 128;          // threads per block
constexpr int kPPT = 2;            // points per thread in phase 1
constexpr int kTile = kTPB * kPPT; // points per tile
constexpr int kBlocksPerSM = 8;   // upper bound of resident blocks per SM (the grid follows the occupancy query)

__host__ __device__ constexpr int row_stride(int D) { return 4 * ((((D + 1) + 3) / 4) | 1); }  // floats, 4 * odd
__host__ __device__ constexpr int cen_stride(int D) { return (D + 3) & ~3; }
__host__ __device__ constexpr int n_chunks(int D) { return (D + 1 + 3) / 4; }                   // float4s holding [x | 1]

inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

struct KmWorkspace {
    float* part;      // [grid][K * NC * 4]
    double* part_d;   // [grid]
    size_t bytes;
};
inline int km_max_grid(int sms) { return sms * kBlocksPerSM; }
KmWorkspace km_carve(void* base, int D, int K, int sms) {
    KmWorkspace w;
    char* b = (char*)base;
    size_t off = 0;
    const int mg = km_max_grid(sms);
    w.part = (float*)(b + off);    off += up256((size_t)mg * K * n_chunks(D) * 4 * sizeof(float));
    w.part_d = (double*)(b + off); off += up256((size_t)mg * sizeof(double));
    w.bytes = off;
    return w;
}
size_t km_smem_bytes(int D, int K) {
    return (size_t)kTile * row_stride(D) * 4      // xs
           + (size_t)K * cen_stride(D) * 4       // centroids
           + (size_t)K * n_chunks(D) * 16        // accumulators
           + (size_t)kTile * 4                   // labels
           + 64 * 8;                             // double reduce
}

template <int D>
__global__ void __launch_bounds__(kTPB) kmeans_assign_kernel(const float* __restrict__ X, long long N, int K,
                                                             const float* __restrict__ centroids,
                                                             int* __restrict__ labels_out, float* __restrict__ part,
                                                             double* __restrict__ part_d) {
    constexpr int DS = row_stride(D);
    constexpr int DC = cen_stride(D);
    constexpr int NC = n_chunks(D);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* xs = reinterpret_cast<float*>(smem_raw);                // [kTile][DS]
    float* cs = xs + kTile * DS;                                    // [K][DC]
    float4* accs = reinterpret_cast<float4*>(cs + K * DC);          // [K * NC]
    int* labs = reinterpret_cast<int*>(accs + K * NC);              // [kTile]
    double* dred = reinterpret_cast<double*>(labs + kTile);         // [kTPB / 32]

    const int tid = threadIdx.x;
    for (int idx = tid; idx < K * DC; idx += kTPB) {
        const int j = idx / DC, a = idx - j * DC;
        cs[idx] = (a < D) ? centroids[(size_t)j * D + a] : 0.f;
    }
    for (int e = tid; e < K * NC; e += kTPB) accs[e] = make_float4(0.f, 0.f, 0.f, 0.f);

    const long long n_tiles = (N + kTile - 1) / kTile;
    double dsum = 0.0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long i0 = tile * kTile;
        const int npts = (int)((N - i0 < kTile) ? (N - i0) : kTile);
        __syncthreads();  // previous tile fully consumed (and cs / accs initialised)
        // ---- phase 0: coalesced copy of the tile into padded rows ---------------------------
        const float* src = X + (size_t)i0 * D;
        for (int idx = tid; idx < kTile * D; idx += kTPB) {
            const int p = idx / D, a = idx - p * D;
            xs[p * DS + a] = (p < npts) ? src[idx] : 0.f;
        }
        for (int idx = tid; idx < kTile * (DS - D); idx += kTPB) {
            const int p = idx / (DS - D), a = D + (idx - p * (DS - D));
            xs[p * DS + a] = (a == D && p < npts) ? 1.f : 0.f;
        }
        __syncthreads();
        // ---- phase 1: nearest centroid of my two points -----------------------------------
        float x0[D], x1[D];
        {
            const float4* r0 = reinterpret_cast<const float4*>(xs + tid * DS);
            const float4* r1 = reinterpret_cast<const float4*>(xs + (tid + kTPB) * DS);
#pragma unroll
            for (int q = 0; q < DC / 4; ++q) {
                const float4 v0 = r0[q], v1 = r1[q];
                const float a0[4] = {v0.x, v0.y, v0.z, v0.w};
                const float a1[4] = {v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                for (int t = 0; t < 4; ++t)
                    if (4 * q + t < D) { x0[4 * q + t] = a0[t]; x1[4 * q + t] = a1[t]; }
            }
        }
        float best0 = INFINITY, best1 = INFINITY;
        int lab0 = 0, lab1 = 0;
#pragma unroll 2
        for (int j = 0; j < K; ++j) {
            const float4* cr = reinterpret_cast<const float4*>(cs + j * DC);
            float d0 = 0.f, d1 = 0.f;
#pragma unroll
            for (int q = 0; q < DC / 4; ++q) {
                const float4 c = cr[q];
                const float cc[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    if (4 * q + t < D) {
                        const float e0 = x0[4 * q + t] - cc[t];
                        const float e1 = x1[4 * q + t] - cc[t];
                        d0 = fmaf(e0, e0, d0);
                        d1 = fmaf(e1, e1, d1);
                    }
                }
            }
            if (d0 < best0) { best0 = d0; lab0 = j; }
            if (d1 < best1) { best1 = d1; lab1 = j; }
        }
        const bool v0 = tid < npts, v1 = tid + kTPB < npts;
        labs[tid] = v0 ? lab0 : -1;
        labs[tid + kTPB] = v1 ? lab1 : -1;
        if (labels_out) {
            if (v0) labels_out[i0 + tid] = lab0;
            if (v1) labels_out[i0 + tid + kTPB] = lab1;
        }
        dsum += (v0 ? (double)sqrtf(best0) : 0.0) + (v1 ? (double)sqrtf(best1) : 0.0);
        __syncthreads();
        // ---- phase 2: owner-accumulates, point order ---------------------------------------
        const float4* xs4 = reinterpret_cast<const float4*>(xs);
        const int4* labs4 = reinterpret_cast<const int4*>(labs);
        for (int e = tid; e < K * NC; e += kTPB) {
            const int j = e / NC, c = e - j * NC;
            float4 a = accs[e];
            const float4* col = xs4 + c;
#pragma unroll 2
            for (int p4 = 0; p4 < kTile / 4; ++p4) {
                const int4 l = labs4[p4];
                if (l.x == j) { const float4 v = col[(4 * p4 + 0) * (DS / 4)]; a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }
                if (l.y == j) { const float4 v = col[(4 * p4 + 1) * (DS / 4)]; a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }
                if (l.z == j) { const float4 v = col[(4 * p4 + 2) * (DS / 4)]; a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }
                if (l.w == j) { const float4 v = col[(4 * p4 + 3) * (DS / 4)]; a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }
            }
            accs[e] = a;
        }
    }
    __syncthreads();
    // ---- epilogue: per-block partials ----------------------------------------------------------
    float4* out = reinterpret_cast<float4*>(part) + (size_t)blockIdx.x * K * NC;
    for (int e = tid; e < K * NC; e += kTPB) out[e] = accs[e];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, o);
    if ((tid & 31) == 0) dred[tid >> 5] = dsum;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < kTPB / 32; ++w) s += dred[w];
        part_d[blockIdx.x] = s;
    }
}

// Sums the per-block partials (double, fixed order: deterministic) into the packed buffer
//   [counts (K) | sums (K*D) | dist_hi | dist_lo].
// One warp per output: lanes take blocks b = lane, lane + 32, ..., then a fixed butterfly.
__global__ void kmeans_finalize_kernel(const float* __restrict__ part, const double* __restrict__ part_d, int nblk,
                                       int D, int K, float* __restrict__ stats) {
    const int NC = n_chunks(D);
    const int row = NC * 4;
    const int lane = threadIdx.x & 31;
    const int o = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);   // output handled by this warp
    const int total = K * (D + 1);
    if (o > total) return;
    double s = 0.0;
    if (o < total) {
        const int j = o / (D + 1), a = o - j * (D + 1);  // a == D: the count column
        for (int b = lane; b < nblk; b += 32) s += (double)part[((size_t)b * K + j) * row + a];
    } else {
        for (int b = lane; b < nblk; b += 32) s += part_d[b];
    }
#pragma unroll
    for (int x = 16; x > 0; x >>= 1) s += __shfl_xor_sync(0xffffffffu, s, x);
    if (lane != 0) return;
    if (o < total) {
        const int j = o / (D + 1), a = o - j * (D + 1);
        if (a == D) stats[j] = (float)s;
        else stats[K + j * D + a] = (float)s;
    } else {
        const float hi = (float)s;
        stats[total] = hi;
        stats[total + 1] = isfinite(hi) ? (float)(s - (double)hi) : 0.f;
    }
}

// Centroid update from the (all-reduced) packed buffer; one block per centroid, D threads.
//   mode 0 (Lloyd, util.py:23-24):      c_j = sums_j / counts_j   (an empty cluster keeps its centroid)
//   mode 1 (minibatch, util.py:56-64):  n_j += counts_j;  c_j += (sums_j - counts_j c_j) / n_j
__global__ void kmeans_update_kernel(int mode, const float* __restrict__ stats, int D, int K,
                                     float* __restrict__ centroids, float* __restrict__ counts_run) {
    const int j = blockIdx.x, a = threadIdx.x;
    const float cnt = stats[j];
    const float sum = stats[K + j * D + a];
    float* c = centroids + (size_t)j * D + a;
    if (mode == 0) {
        if (cnt > 0.f) *c = sum / cnt;
    } else {
        const float n_new = counts_run[j] + cnt;
        __syncthreads();
        const float eta = 1.f / n_new;
        *c = *c + eta * (sum - cnt * *c);
        if (a == 0) counts_run[j] = n_new;
    }
}

// Minibatch gather for the device-resident batch iterator: rows idx[b] of X [N,D] and
// of the noise table [N,n_s] (n_s = D*D, D or 0) -> contiguous outX [B,D], outS [B,n_s], one launch.
// Replaces the per-sample __getitem__ + collate of the reference's DataLoader
// (deconv/gmm/data.py:7-17, sgd_gmm.py:139-145).  128-bit copies when the rows allow.
// Flat grid-stride copy: element e of the batch's noise block maps to (point b = e / n4, 16-byte chunk c = e % n4);
// four independent loads are in flight per thread before the first store (the copy is latency-bound otherwise:
// one warp per point left 25 of 32 lanes with a single 16-byte load each at D = 10).
__global__ void __launch_bounds__(256) gather_points_kernel(const float* __restrict__ X, const float* __restrict__ S,
                                                            const long long* __restrict__ idx, long long B, int D, int n_s,
                                                            float* __restrict__ outX, float* __restrict__ outS) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nth = (long long)gridDim.x * blockDim.x;
    const bool vec = n_s > 0 && (n_s & 3) == 0 && ((((uintptr_t)S) | ((uintptr_t)outS)) & 15) == 0;
    if (vec) {
        const int n4 = n_s >> 2;
        const long long total = B * n4;
        const float4* S4 = reinterpret_cast<const float4*>(S);
        float4* O4 = reinterpret_cast<float4*>(outS);
        long long e = tid;
        for (; e + 3 * nth < total; e += 4 * nth) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long long ee = e + u * nth;
                const long long b = ee / n4;
                const int c = (int)(ee - b * n4);
                v[u] = __ldg(S4 + idx[b] * n4 + c);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) O4[e + u * nth] = v[u];
        }
        for (; e < total; e += nth) {
            const long long b = e / n4;
            const int c = (int)(e - b * n4);
            O4[e] = __ldg(S4 + idx[b] * n4 + c);
        }
    } else if (n_s > 0) {
        const long long total = B * n_s;
        for (long long e = tid; e < total; e += nth) {
            const long long b = e / n_s;
            const int c = (int)(e - b * n_s);
            outS[e] = __ldg(S + idx[b] * n_s + c);
        }
    }
    const long long totalX = B * D;
    for (long long e = tid; e < totalX; e += nth) {
        const long long b = e / D;
        const int a = (int)(e - b * D);
        outX[e] = __ldg(X + idx[b] * D + a);
    }
}

typedef void (*assign_fn)(const float*, long long, int, const float*, int*, float*, double*);
template <int D>
assign_fn assign_entry() { return &kmeans_assign_kernel<D>; }
#define KM_ROW4(D) assign_entry<D>(), assign_entry<D + 1>(), assign_entry<D + 2>(), assign_entry<D + 3>()
const assign_fn kAssign[DGMM_MAX_D + 1] = {nullptr,     KM_ROW4(1),  KM_ROW4(5),  KM_ROW4(9), KM_ROW4(13),
                                           KM_ROW4(17), KM_ROW4(21), KM_ROW4(25), KM_ROW4(29)};

// ---- scratch of the allocation-free entry point (dgmm_kmeans_assign_static_f32) -----------------------------------
// A file-scope device array: every GPU that loads the library holds these 8 MiB until the process exits.  In HOST code
// the name `g_km_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per
// device with cudaGetSymbolAddress (that device current) and cached.
constexpr size_t kKmStaticBytes = (size_t)8 << 20;
static __device__ __align__(256) unsigned char g_km_scratch[kKmStaticBytes];

int km_static_scratch(void** out) {
    static void* cache[64] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DGMM_ERR_NO_DEVICE;
    if (!cache[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_km_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!ptr) return DGMM_ERR_NO_DEVICE;
        cache[dev] = ptr;
    }
    *out = cache[dev];
    return 0;
}

int km_device(int* sms, int* smem_optin) {
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return DGMM_ERR_NO_DEVICE;
    if (cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return DGMM_ERR_NO_DEVICE;
    cudaDeviceGetAttribute(smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (major != 10) return DGMM_ERR_NO_DEVICE;
    return 0;
}

}  // namespace

extern "C" {

size_t dgmm_kmeans_workspace_bytes(int D, int K) {
    if (D < 1 || D > DGMM_MAX_D || K < 1 || K > DGMM_MAX_K) return 0;
    int sms = 0, optin = 0;
    if (km_device(&sms, &optin)) return 0;
    return km_carve(nullptr, D, K, sms).bytes;
}

int dgmm_kmeans_assign_f32(const float* X, int64_t N, int D, int K, const float* centroids, int32_t* labels,
                           float* out_stats, void* workspace, size_t workspace_bytes, void* stream) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1 || K > DGMM_MAX_K) return DGMM_ERR_UNSUPPORTED_K;
    if (N < 0 || !centroids || !out_stats || (N > 0 && !X)) return DGMM_ERR_BAD_ARG;
    int sms = 0, optin = 0;
    int rc = km_device(&sms, &optin);
    if (rc) return rc;
    if (!workspace) return DGMM_ERR_WORKSPACE;
    const KmWorkspace ws = km_carve(workspace, D, K, sms);
    if (ws.bytes > workspace_bytes) return DGMM_ERR_WORKSPACE;
    const size_t smem = km_smem_bytes(D, K);
    if (smem > (size_t)optin) return DGMM_ERR_UNSUPPORTED_K;
    cudaStream_t st = (cudaStream_t)stream;
    const assign_fn fn = kAssign[D];
    CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long n_tiles = (N + kTile - 1) / kTile;
    int occ = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kTPB, smem));
    occ = std::max(1, std::min(occ, kBlocksPerSM));
    int grid = (int)std::min<long long>((long long)sms * occ, n_tiles > 0 ? n_tiles : 1);
    fn<<<grid, kTPB, smem, st>>>(X, (long long)N, K, centroids, labels, ws.part, ws.part_d);
    CUDA_TRY(cudaGetLastError());
    const int total = K * (D + 1) + 1;                       // one warp per output, 8 warps per block
    kmeans_finalize_kernel<<<(total + 7) / 8, 256, 0, st>>>(ws.part, ws.part_d, grid, D, K, out_stats);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

size_t dgmm_kmeans_static_scratch_bytes(void) { return kKmStaticBytes; }

// dgmm_kmeans_assign_f32 with its workspace in the library's own static device array: no cudaMalloc / cudaFree / allocating
// routine anywhere on this path.  Shapes whose workspace (dgmm_kmeans_workspace_bytes) exceeds the array are rejected.
// All calls on a device share the array: calls that overlap on the GPU (two streams) race.  Nothing in it has to start at
// zero: every partial the finalise kernel reads is written by the assignment kernel of the same call.
int dgmm_kmeans_assign_static_f32(const float* X, int64_t N, int D, int K, const float* centroids, int32_t* labels,
                                  float* out_stats, void* stream) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1 || K > DGMM_MAX_K) return DGMM_ERR_UNSUPPORTED_K;
    int sms = 0, optin = 0;
    int rc = km_device(&sms, &optin);
    if (rc) return rc;
    if (km_carve(nullptr, D, K, sms).bytes > kKmStaticBytes) return DGMM_ERR_STATIC_SCRATCH;
    void* scratch = nullptr;
    rc = km_static_scratch(&scratch);
    if (rc) return rc;
    return dgmm_kmeans_assign_f32(X, N, D, K, centroids, labels, out_stats, scratch, kKmStaticBytes, stream);
}

int dgmm_gather_points_f32(const float* X, const float* S, int noise_layout, const int64_t* idx, int64_t B, int D,
                           float* outX, float* outS, void* stream) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (noise_layout != DGMM_NOISE_FULL && noise_layout != DGMM_NOISE_DIAG && noise_layout != DGMM_NOISE_NONE &&
        noise_layout != DGMM_NOISE_PACKED)
        return DGMM_ERR_LAYOUT;
    const int n_s = noise_layout == DGMM_NOISE_FULL ? D * D
                    : (noise_layout == DGMM_NOISE_DIAG ? D : (noise_layout == DGMM_NOISE_PACKED ? D * (D + 1) / 2 : 0));
    if (B < 0 || (B > 0 && (!X || !idx || !outX || (n_s && (!S || !outS))))) return DGMM_ERR_BAD_ARG;
    if (B == 0) return 0;
    int sms = 0, optin = 0;
    int rc = km_device(&sms, &optin);
    if (rc) return rc;
    const long long work = B * std::max(n_s / 4, D);          // ~ one 16-byte chunk per thread and trip
    const int grid = (int)std::max<long long>(1, std::min<long long>((work + 1023) / 1024, (long long)sms * 8));
    gather_points_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, S, (const long long*)idx, (long long)B, D, n_s,
                                                                 outX, outS);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int dgmm_kmeans_update_f32(int mode, const float* stats, int D, int K, float* centroids, float* counts_run,
                           void* stream) {
    if (D < 1 || D > DGMM_MAX_D) return DGMM_ERR_UNSUPPORTED_D;
    if (K < 1 || K > DGMM_MAX_K) return DGMM_ERR_UNSUPPORTED_K;
    if (!stats || !centroids || (mode != 0 && mode != 1) || (mode == 1 && !counts_run)) return DGMM_ERR_BAD_ARG;
    kmeans_update_kernel<<<K, D, 0, (cudaStream_t)stream>>>(mode, stats, D, K, centroids, counts_run);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

}  // extern "C"
