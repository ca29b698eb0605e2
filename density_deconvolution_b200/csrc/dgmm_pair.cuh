// dgmm_pair.cuh -- the fused log-likelihood / E-step-statistics kernel for sm_100a.
//
// Mapping ("threads are components"): a group of GS threads owns one data point at a time;
// thread `jl` of the group owns mixture component j = jl for the whole kernel.  A block
// holds NSLOT such groups.  For the point in its slot every thread
//   1. forms T = Sigma_j + S_i (Sigma_j: its private column in shared memory or global;
//      S_i: broadcast reads from the TMA-staged tile),
//   2. factorises T in registers (dgmm_core.cuh) and gets log p_ij,
//   3. takes part in the logsumexp over the group (warp shuffles; for GS > 32 one shared
//      exchange + a named barrier per point) -> ll_i, r_ij,
//   4. (MODE_STATS) inverts the factor and adds omega_ij * {1, u, u u^T - T^{-1}} to its
//      PRIVATE accumulators (registers / shared columns / global columns).
// Because component j never leaves the thread there is no cross-lane reduction of the
// K(1+D+D(D+1)/2) statistics in the hot loop, and because the group covers all K components
// at once, ll_i is known before the statistics are weighted: one pass, no recomputation
// (SURVEY.md 7.2 "two normalisations").
//
// Input streaming: the points of a tile are contiguous in X [N,D] and S [N,D,D], so one
// tile is three 1-D bulk copies (cp.async.bulk -> UBLKCP, the TMA engine) into a NST-deep
// shared-memory ring signalled through mbarriers; unaligned bases and the ragged last tile
// take a cooperative LDG->STS path in the same kernel.
//
// Replaces deconv/gmm/deconv_gmm.py:39-107 and deconv/gmm/sgd_deconv_gmm.py:14-28 of the
// reference (see include/dgmm_b200.h for the per-entry-point map).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "dgmm_core.cuh"

namespace dgmm {

enum : int { MODE_LOGLIK = 0, MODE_STATS = 1 };
// Largest D whose log-likelihood kernel re-forms the next point's T inside the factorisation (refill).
#ifndef DGMM_LL_ROTATE_D
#define DGMM_LL_ROTATE_D 12
#endif
enum : int { ST_REG = 0, ST_SMEM = 1, ST_GLOBAL = 2, ST_TMEM = 3 };  // where Sigma_j / accumulators live
enum : int { NOISE_FULL = 0, NOISE_DIAG = 1, NOISE_NONE = 2, NOISE_SHARED = 3, NOISE_PACKED = 4 };

constexpr int kMaxStages = 4;
constexpr int kPairMaxThreads = 256;  // largest block the pair kernel is compiled for (D <= 8 also has 512-thread variants)
constexpr int kPairWideThreads = 512;
constexpr int kPairWideMaxD = 8;      // up to here K = 257..512 runs as ONE group of 512 threads (<= 128 registers each)
constexpr int kSmemHeader = 1152;  // mbarriers (64 B) + logsumexp exchange (2 x 256 B) + double reduce (512 B)

struct PairParams {
    const float* X;          // [N,D]
    const float* S;          // [N,D,D] | [N,D] | [N,D(D+1)/2] (packed lower triangles) | unused
    const float* g;          // [N] or null
    const float* sig_cols;   // [ND][GS]  transposed packed lower triangles (+ shared noise folded in)
    const float* mu_cols;    // [D][GS]
    const float* cj;         // [GS]  log alpha_j - D/2 log 2pi  (-inf for j >= K)
    float* out_ll;           // [N] or null
    float* part;             // [grid][NS][GS]  per-block partial statistics
    double* part_ll;         // [grid][2]       per-block sum ll, sum g
    int* nonpd;              // [1] device counter
    long long N;
    long long n_tiles;
    int K, GS, NSLOT, PT, NST, noise;
    int use_bulk;
    float resp_floor;
    // derived on the host (constant-bank operands in the kernel): ring-stage layout and loop counts
    int n_s;          // noise floats per point in the tile (D*D | D | D(D+1)/2 | 0)
    int off_X, off_g; // byte offsets of the X and g sub-tiles inside a stage (S sits at 0)
    int stage_bytes;
    int rounds;       // PT / NSLOT
    int tmem_cols;    // columns to allocate for ST_TMEM accumulators (power of two, 32..512)
};

// ---------------------------------------------------------------------------------------
// PTX helpers (mbarrier + 1-D bulk copy)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return done != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------------------------------
// Tensor memory as a per-thread scratchpad (accumulator placement ST_TMEM).
// With the 32x32b access shape lane i of warp w owns TMEM lane 32*(w % 4) + i, so every thread has
// a private row of up to 512 fp32 columns that costs neither registers nor shared memory; the
// K x (1 + D + D(D+1)/2) statistics live there and are updated with tcgen05.ld / tcgen05.st
// (8 columns per instruction).  Measured (tools/ubench_tmem.cu): 150-300 B/cycle/SM of
// read-modify-write, next to the 128 B/cycle of shared memory that the Sigma_j / S_i reads keep.
// ---------------------------------------------------------------------------------------
__host__ __device__ constexpr int tmem_cols_per_thread(int ns) { return (ns + 7) & ~7; }
__host__ __device__ constexpr int tmem_alloc_cols(int ns, int warps) {
    const int need = tmem_cols_per_thread(ns) * ((warps + 3) / 4);
    int c = 32;
    while (c < need) c <<= 1;
    return c;  // > 512: does not fit
}
__device__ __forceinline__ void tmem_alloc(uint32_t* slot_in_smem, int cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_in_smem)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t base, int cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(cols) : "memory");
}
__device__ __forceinline__ void tmem_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "f"(v[0]),
                 "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
          "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(
            taddr),
        "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]), "f"(v[8]), "f"(v[9]),
        "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]));
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// group classes of the pair kernel (see dgmm_pair_body.inc)
enum : int { GCLS_SUB = 0, GCLS_WARP = 1, GCLS_MW4 = 2, GCLS_MWN = 3, GCLS_ANY = 4,
             GCLS_MWN512 = 5 };   // host-side only: GCLS_MWN compiled for blocks of 512 threads (GS = 257..512, D <= 8)
// noise classes of the pair kernel
enum : int { NZ_FULL = 0, NZ_DIAG = 1, NZ_ANY = 2, NZ_PACKED = 3 };

// predicated 32-bit global store (no branch in the instruction stream)
__device__ __forceinline__ void st_global_if(float* ptr, float v, bool pred) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %2, 0;\n"
        "@p st.global.f32 [%0], %1;\n"
        "}\n" ::"l"(ptr), "f"(v), "r"((int)pred)
        : "memory");
}
// an empty use of a register value
__device__ __forceinline__ void keep_alive(float v) { asm volatile("" ::"f"(v)); }


__device__ __forceinline__ float warp_max_f32(float v) {
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));  // CREDUX.MAX.F32 (sm_100a)
    return r;
}

// Lanes that share one point in the per-tile logsumexp pass of the log-likelihood kernel: the
// largest power of two <= min(32, TPB / PT).
__host__ __device__ inline int lse_lanes(int TPB, int PT) {
    int r = 1;
    while (2 * r <= 32 && 2 * r * PT <= TPB) r <<= 1;
    return r;
}

// The pair kernel re-packs a tile's noise into 16-byte aligned rows (128-bit shared-memory reads) where the
// input layout does not give them (compiled placements with a fixed noise class only):
//   full noise with odd D          -> lower triangles as rows of RS = roundup4(D) floats, [PT][D][RS]
//   packed noise with ND % 4 != 0  -> packed triangles at a stride of roundup4(ND) floats, [PT][NDp]
__host__ __device__ constexpr bool repack_noise(int D, int nz) {
    return (nz == NZ_FULL && (D & 1) && D > 1 && D <= DGMM_UNROLL_D) ||
           (nz == NZ_PACKED && ((D * (D + 1) / 2) & 3) != 0 && D <= DGMM_UNROLL_D);
}
__host__ __device__ constexpr int repack_floats_per_point(int D, int nz) {
    return !repack_noise(D, nz) ? 0 : (nz == NZ_PACKED ? ((D * (D + 1) / 2 + 3) & ~3) : D * ((D + 3) & ~3));
}

// Bytes of one ring stage and the offsets of its three sub-tiles (all multiples of 16).
__host__ __device__ inline int noise_floats_per_point(int D, int noise) {
    return noise == NOISE_FULL ? D * D : (noise == NOISE_DIAG ? D : (noise == NOISE_PACKED ? D * (D + 1) / 2 : 0));
}
__host__ __device__ inline int align16(int b) { return (b + 15) & ~15; }
struct StageLayout {
    int off_S, off_X, off_g, bytes;
};
__host__ __device__ inline StageLayout stage_layout(int D, int noise, int PT, bool has_g) {
    StageLayout L;
    L.off_S = 0;
    L.off_X = align16(PT * noise_floats_per_point(D, noise) * 4);
    L.off_g = L.off_X + align16(PT * D * 4);
    L.bytes = L.off_g + (has_g ? align16(PT * 4) : 0);
    L.bytes = (L.bytes + 127) & ~127;
    return L;
}

}  // namespace dgmm

#define DGMM_LOOP _Pragma("unroll")
namespace dgmm {
namespace unrolled {
#include "dgmm_pair_body.inc"
}  // namespace unrolled
}  // namespace dgmm
#undef DGMM_LOOP

