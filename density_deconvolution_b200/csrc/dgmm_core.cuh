// dgmm_core.cuh -- per-(point, component) small-D linear algebra, thread-private.
//
// One thread owns one pair (i, j) and keeps the packed lower triangle of
// T_ij = Sigma_j + S_i in registers (D <= DGMM_UNROLL_D: every loop fully unrolled so that
// all indices are static) or in local memory (larger D: rolled loops, slower, same maths).
//
// What the reference does with batched LAPACK calls on materialised [N,K,D,D] tensors
// (deconv/gmm/deconv_gmm.py:43-62: cholesky, cholesky_solve(I), matmuls) is done here in
// place on D(D+1)/2 floats:
// Pair kernel (hot path) -- square-root-free, inverse-free formulation:
//   ldl_forward     : T = L D L^T (unit lower L, right-looking), y = L^{-1} d carried along,
//                     sum log2 of the pivots; the diagonal is left holding 1/D_kk.
//   ldl_back_solve  : u = L^{-T} D^{-1} y = T^{-1} d (back substitution, no inverse needed).
//   ldl_inverse     : L, 1/D -> T^{-1} (lower triangle) in place by the Takahashi recurrence
//                     Z = D^{-1} L^{-1} + (I - L^T) Z, bottom row upwards; L^{-1} is never formed.
// Materialising kernel (API-parity path, dgmm_full.cuh) -- classic Cholesky route:
//   chol_forward : T = L L^T (right-looking), y = L^{-1} d carried as an extra row,
//                  sum log2 of the pivots; the diagonal is left holding 1/L_kk.
//   tri_inverse  : L -> L^{-1} in place.
//   back_apply   : u = L^{-T} y = T^{-1} d.
//   gram_lower   : L^{-1} -> T^{-1} = L^{-T} L^{-1} (lower triangle) in place.
//
// The functions are __host__ __device__ so that tests/ can compile them for the CPU and
// check the index arithmetic against the oracle without a GPU (test tool only; the
// product has no CPU path).
#pragma once

#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define DGMM_HD __host__ __device__ __forceinline__
#else
#define DGMM_HD inline
#endif

#ifndef DGMM_UNROLL_D
#define DGMM_UNROLL_D 16
#endif

namespace dgmm {

DGMM_HD constexpr int tri(int a, int b) { return a * (a + 1) / 2 + b; }  // requires b <= a
// row of packed-triangle position t (inverse of tri); folds to a constant for constant t
DGMM_HD constexpr int tri_row(int t) {
    int a = 0;
    while ((a + 1) * (a + 2) / 2 <= t) ++a;
    return a;
}
// per-thread column stride: a multiple of 4 floats holding an odd number of float4s, so that
// 32 threads reading their own column with 128-bit accesses never conflict on a bank
DGMM_HD constexpr int pad_col(int n) {
    int m = (n + 3) / 4;
    if (m % 2 == 0) ++m;
    return 4 * m;
}

// Work-balanced slicing of the rows of ldl_inverse into P slices: rows [inv_slice_lo(D,k,P),
// inv_slice_lo(D,k-1,P)) form slice k (k = 0..P-1; slice 0 starts at row D-1).  Row i costs about
// (D-i)(D-1-i)+2 instructions.
DGMM_HD constexpr int inv_slice_lo(int D, int k, int P) {
    if (k < 0) return D;
    if (k >= P - 1) return 0;
    int total = 0;
    for (int i = 0; i < D; ++i) total += (D - i) * (D - 1 - i) + 2;
    const int target = total * (k + 1) / P;
    int cum = 0, i = D;
    while (i > 0) {
        const int w = (D - (i - 1)) * (D - 1 - (i - 1)) + 2;
        if (cum + w > target && cum > 0) break;
        cum += w;
        --i;
        if (cum >= target) break;
    }
    return i;
}

template <int D>
struct Dim {
    static constexpr int ND = D * (D + 1) / 2;  // packed lower triangle
    static constexpr int NS = 1 + D + ND;       // per-component packed statistics s0|s1|S2(lower)
};

constexpr float kLn2 = 0.6931471805599453f;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kHalfLog2Pi = 0.9189385332046727f;

DGMM_HD float mufu_rsqrt(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / sqrtf(x);
#endif
}

DGMM_HD float mufu_rcp(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / x;
#endif
}

DGMM_HD float mufu_lg2(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return log2f(x);
#endif
}

DGMM_HD float mufu_ex2(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return exp2f(x);
#endif
}

// 1/sqrt(p): MUFU.RSQ plus one Newton step (full fp32 accuracy; pivots may span
// 1e-4 .. 1e12, SURVEY.md 7.2 "dynamic range of S_i").
DGMM_HD float rsqrt_refined(float p) {
    float r = mufu_rsqrt(p);
    float h = 0.5f * p * r;
    return fmaf(r, fmaf(-h, r, 0.5f), r);
}

}  // namespace dgmm

#define DGMM_LOOP _Pragma("unroll")
namespace dgmm {
namespace unrolled {
#include "dgmm_core_body.inc"
}  // namespace unrolled
}  // namespace dgmm
#undef DGMM_LOOP

#define DGMM_LOOP _Pragma("unroll 4")
namespace dgmm {
namespace rolled {
#include "dgmm_core_body.inc"
}  // namespace rolled
}  // namespace dgmm
#undef DGMM_LOOP

// ---------------------------------------------------------------------------------------
// Row slices of ldl_inverse with every index a compile-time constant (template expansion
// instead of `#pragma unroll`: nvcc stops honouring the pragma on the larger slices of
// D = 13..16, which would put the triangle into local memory).  Rows I_HI, I_HI-1, ..., I_LO;
// callers run the slices in descending order, D-1 first: the pair kernel weaves them between
// the steps of the logsumexp.
// ---------------------------------------------------------------------------------------
#include <type_traits>
#include <utility>

namespace dgmm {
namespace unrolled {

template <int... Is, class F>
DGMM_HD void static_for_impl(std::integer_sequence<int, Is...>, F&& f) {
    (f(std::integral_constant<int, Is>{}), ...);
}
// f(integral_constant<int, 0>) ... f(integral_constant<int, N-1>)
template <int N, class F>
DGMM_HD void static_for(F&& f) {
    if constexpr (N > 0) static_for_impl(std::make_integer_sequence<int, N>{}, f);
}

template <int D, int I>
DGMM_HD void ldl_inverse_row(float* A) {
    float z[D];
    static_for<D - 1 - I>([&](auto jc) {
        constexpr int j = I + 1 + decltype(jc)::value;
        float s = -A[tri(D - 1, I)] * A[tri(D - 1, j)];
        static_for<D - 2 - I>([&](auto kc) {
            constexpr int k = D - 2 - decltype(kc)::value;  // D-2 down to I+1
            s = fmaf(-A[tri(k, I)], A[k >= j ? tri(k, j) : tri(j, k)], s);
        });
        z[j] = s;
    });
    float s0 = A[tri(I, I)], s1 = 0.f;  // two partial sums over the fresh z[k]
    static_for<D - 1 - I>([&](auto kc) {
        constexpr int k = I + 1 + decltype(kc)::value;
        if constexpr ((k - I) & 1) s0 = fmaf(-A[tri(k, I)], z[k], s0);
        else s1 = fmaf(-A[tri(k, I)], z[k], s1);
    });
    A[tri(I, I)] = s0 + s1;
    static_for<D - 1 - I>([&](auto jc) {
        constexpr int j = I + 1 + decltype(jc)::value;
        A[tri(j, I)] = z[j];
    });
}

template <int D, int I_HI, int I_LO>
DGMM_HD void ldl_inverse_rows(float* A) {
    static_for<I_HI - I_LO + 1>([&](auto ic) { ldl_inverse_row<D, I_HI - decltype(ic)::value>(A); });
}

}  // namespace unrolled

namespace rolled {
template <int D, int I_HI, int I_LO>
DGMM_HD void ldl_inverse_rows(float* A) {
    static_assert(I_HI == D - 1 && I_LO == 0, "the rolled path only runs the whole inverse");
    ldl_inverse<D>(A);
}
}  // namespace rolled
}  // namespace dgmm
