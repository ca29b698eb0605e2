// dgmm_full.cuh -- materialising E-step (small N): log r [N,K], b [N,K,D], B [N,K,D,D].
//
// One thread per (point, component) pair, flat over N*K.  Replaces the posterior-moment
// part of DeconvGMM._e_step (deconv/gmm/deconv_gmm.py:64-79) and
// SGDDeconvGMM.posterior_params (deconv/gmm/sgd_deconv_gmm.py:71-94):
//   b_ij = mu_j + Sigma_j u_ij,   B_ij = Sigma_j - Sigma_j T_ij^{-1} Sigma_j.
// ll_i comes from a preceding log-likelihood pass (pair kernel), so log r = log p - ll.
// The output is N*K*D*D floats: this kernel is bound by its own HBM writes and is meant
// for the API-parity call sites, not for the N = 1e7 path (256 GB at c2).
#pragma once

#include "dgmm_core.cuh"
#include "dgmm_pair.cuh"

// Largest D whose materialising E-step forms B_ij = Sigma_j - V^T V from V = L^{-1} Sigma_j in registers
#ifndef DGMM_FULL_V_D
#define DGMM_FULL_V_D 10
#endif
// resident blocks per SM of that variant: 3 x 128 threads at <= 168 registers (48 bytes of spills at D = 10) beat 2 at 198
// (gpurun A/B at the c2 shape: 1.02 -> 0.95 ms)
#ifndef DGMM_FULL_MINB
#define DGMM_FULL_MINB 3
#endif

namespace dgmm {

// dynamic shared memory of full_kernel<D>: the per-warp D x D staging tiles of the small-D route
__host__ __device__ constexpr size_t full_dyn_smem_bytes(int D) {
    return D <= DGMM_FULL_V_D ? (size_t)4 * 32 * ((D * D) | 1) * sizeof(float)
           : (D <= DGMM_UNROLL_D ? (size_t)D * ((D + 3) & ~3) * 128 * sizeof(float) : 0);   // D = 11..16: V per thread
}

struct FullParams {
    const float* X;
    const float* S;
    const float* log_alpha;  // [K]
    const float* mu;         // [K,D]
    const float* Sigma;      // [K,D,D]
    const float* ll;         // [N]
    float* out_log_r;        // [N,K] or null
    float* out_b;            // [N,K,D] or null
    float* out_B;            // [N,K,D,D] or null
    int* nonpd;
    long long N;
    int K, noise;
};

// posterior sampling (sample_kernel): one thread per (point, draw)
struct SampleParams {
    const float* X;
    const float* S;
    const float* mu;         // [K,D]
    const float* Sigma;      // [K,D,D]
    const float* log_r;      // [N,K]  log responsibilities (full_kernel)
    const float* u;          // [N,M]  uniforms in [0,1)
    const float* z;          // [N,M,D] standard normals
    float* out;              // [N,M,D]
    int* out_comp;           // [N,M] picked component, or null
    int* nonpd;              // counts draws whose B_ij was not positive definite
    long long N;
    int K, M, noise;
};

}  // namespace dgmm

// the loop over the rows of B_ij: fully unrolled with the rest (compile-time row offsets and divisors in the
// staged stores) for D <= 16, rolled above
#define DGMM_FULL_CLOOP _Pragma("unroll")
#define DGMM_LOOP _Pragma("unroll")
namespace dgmm {
namespace unrolled {
#include "dgmm_full_body.inc"
}  // namespace unrolled
}  // namespace dgmm
#undef DGMM_LOOP
#undef DGMM_FULL_CLOOP

#define DGMM_FULL_CLOOP _Pragma("unroll 1")
#define DGMM_LOOP _Pragma("unroll 4")
namespace dgmm {
namespace rolled {
#include "dgmm_full_body.inc"
}  // namespace rolled
}  // namespace dgmm
#undef DGMM_LOOP
#undef DGMM_FULL_CLOOP
