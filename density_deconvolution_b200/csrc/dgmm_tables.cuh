// dgmm_tables.cuh -- per-D kernel tables.  Each dgmm_inst_dXX.cu instantiates the kernels
// of one dimension (so that the 32 dimensions compile in parallel) and exports one lookup
// function; dgmm_api.cu launches through the returned pointers with cudaLaunchKernel.
#pragma once

#include "dgmm_full.cuh"
#include "dgmm_pair.cuh"

namespace dgmm {

// One compiled pair-kernel variant: the entry point plus the placement / launch bounds it was
// instantiated with (the launch planner sizes shared memory and blocks from these).
struct PairVariant {
    const void* fn;
    int sigp, accp, gcls, nz, max_threads, min_blocks;
};

// Returns the variant for (mode, which, gcls); fn == nullptr when that combination is not compiled
// for this D.  `which`: 0 = main placement, 1 = alternative placement (tuning sweeps),
// 2 = universal fallback (Sigma and accumulators in global memory, any group size).
typedef PairVariant (*pair_lookup_fn)(int mode, int which, int gcls, int nz);
typedef const void* (*full_lookup_fn)();

template <int D, int MODE, int SIGP, int ACCP, int GCLS, int NZ, int MAXT, int MINB>
inline PairVariant pair_variant() {
    static_assert(D <= DGMM_UNROLL_D, "thread-per-pair kernels above D = 16 exist only as pair_variant_wide_unrolled");
    PairVariant v;
    v.fn = reinterpret_cast<const void*>(&unrolled::pair_kernel<D, MODE, SIGP, ACCP, GCLS, NZ, MAXT, MINB>);
    v.sigp = SIGP; v.accp = ACCP; v.gcls = GCLS; v.nz = NZ; v.max_threads = MAXT; v.min_blocks = MINB;
    return v;
}

// D = 17..DGMM_PAIR_UNROLL_D (= 20): the fully unrolled body still fits 255 registers with a few spilled words
// (ptxas: <= 480 bytes of stack at D = 20), which is an order of magnitude faster than the rolled fallback.
// Compiled only for the group class of K = 64..128 (one block of 128 threads per SM) to keep build time down.
#ifndef DGMM_PAIR_UNROLL_D
#define DGMM_PAIR_UNROLL_D 20   // (21..24 existed in round 1: 0.10-0.34 of peak; the cooperative kernels serve them now)
#endif
template <int D, int MODE, int SIGP, int ACCP, int GCLS, int NZ>
inline PairVariant pair_variant_wide_unrolled() {
    PairVariant v;
    v.fn = reinterpret_cast<const void*>(&unrolled::pair_kernel<D, MODE, SIGP, ACCP, GCLS, NZ, kPairMaxThreads, 1>);
    v.sigp = SIGP; v.accp = ACCP; v.gcls = GCLS; v.nz = NZ; v.max_threads = kPairMaxThreads; v.min_blocks = 1;
    return v;
}
// K = 64..128: Sigma_j in shared-memory columns; K = 129..256 (one 256-thread group): the columns no longer
// fit beside the ring, Sigma_j is read through L1 instead.
template <int D, int MODE, int ACCP>
inline PairVariant pair_wide_unrolled(int gcls, int nz, bool sigma_global = false) {
    if (gcls == GCLS_MW4 && sigma_global) {
        if (nz == NZ_FULL) return pair_variant_wide_unrolled<D, MODE, ST_GLOBAL, ACCP, GCLS_MW4, NZ_FULL>();
        if (nz == NZ_DIAG) return pair_variant_wide_unrolled<D, MODE, ST_GLOBAL, ACCP, GCLS_MW4, NZ_DIAG>();
    } else if (gcls == GCLS_MW4) {
        if (nz == NZ_FULL) return pair_variant_wide_unrolled<D, MODE, ST_SMEM, ACCP, GCLS_MW4, NZ_FULL>();
        if (nz == NZ_DIAG) return pair_variant_wide_unrolled<D, MODE, ST_SMEM, ACCP, GCLS_MW4, NZ_DIAG>();
    } else if (gcls == GCLS_MWN) {
        if (nz == NZ_FULL) return pair_variant_wide_unrolled<D, MODE, ST_GLOBAL, ACCP, GCLS_MWN, NZ_FULL>();
        if (nz == NZ_DIAG) return pair_variant_wide_unrolled<D, MODE, ST_GLOBAL, ACCP, GCLS_MWN, NZ_DIAG>();
    }
    return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
}

// D = 17..20 log-likelihood pass: Sigma_j column (up to 210 floats) in the thread's tensor-memory row instead of global
// memory through L1 (K = 64..128: two 128-thread blocks per SM of 256 columns each; K = 129..256: one 256-thread block).
template <int D>
inline PairVariant pair_wide_loglik_tmem(int gcls, int nz) {
    if (gcls == GCLS_MW4) {
        if (nz == NZ_FULL) return pair_variant_wide_unrolled<D, MODE_LOGLIK, ST_TMEM, ST_GLOBAL, GCLS_MW4, NZ_FULL>();
        if (nz == NZ_DIAG) return pair_variant_wide_unrolled<D, MODE_LOGLIK, ST_TMEM, ST_GLOBAL, GCLS_MW4, NZ_DIAG>();
    } else if (gcls == GCLS_MWN) {
        if (nz == NZ_FULL) return pair_variant_wide_unrolled<D, MODE_LOGLIK, ST_TMEM, ST_GLOBAL, GCLS_MWN, NZ_FULL>();
        if (nz == NZ_DIAG) return pair_variant_wide_unrolled<D, MODE_LOGLIK, ST_TMEM, ST_GLOBAL, GCLS_MWN, NZ_DIAG>();
    }
    return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
}

// The four group classes of one placement.  Groups wider than 128 lanes (GCLS_MWN) need blocks of
// 256 threads, so that class is always compiled with the (256, 1) bounds.
template <int D, int MODE, int SIGP, int ACCP, int NZ, int MAXT, int MINB>
inline PairVariant pair_by_gcls_nz(int gcls) {
    switch (gcls) {
        case GCLS_SUB:  return pair_variant<D, MODE, SIGP, ACCP, GCLS_SUB, NZ, MAXT, MINB>();
        case GCLS_WARP: return pair_variant<D, MODE, SIGP, ACCP, GCLS_WARP, NZ, MAXT, MINB>();
        case GCLS_MW4:  return pair_variant<D, MODE, SIGP, ACCP, GCLS_MW4, NZ, MAXT, MINB>();
        case GCLS_MWN:  return pair_variant<D, MODE, SIGP, ACCP, GCLS_MWN, NZ, kPairMaxThreads, 1>();
        default: break;
    }
    return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
}

template <int D, int MODE, int SIGP, int ACCP, int MAXT, int MINB>
inline PairVariant pair_by_gcls(int gcls, int nz) {
    if (nz == NZ_FULL) return pair_by_gcls_nz<D, MODE, SIGP, ACCP, NZ_FULL, MAXT, MINB>(gcls);
    if (nz == NZ_PACKED) return pair_by_gcls_nz<D, MODE, SIGP, ACCP, NZ_PACKED, MAXT, MINB>(gcls);
    return pair_by_gcls_nz<D, MODE, SIGP, ACCP, NZ_DIAG, MAXT, MINB>(gcls);
}

// Placements (measured on B200, profiles/r01_placement_ab.txt; fraction of the FP32 peak, STATS):
//   D <= 10      : Sigma_j in shared-memory columns, accumulators in registers, 128 threads x 3
//                  blocks/SM (168 registers)                       [D=10,K=64: 0.57 vs 0.55 TMEM]
//   D = 11, 12   : accumulators in tensor memory, 4 resp. 3 blocks/SM  [D=12: 0.59 vs 0.50 registers]
//   13 <= D <= 16: Sigma_j in shared-memory columns, accumulators in tensor memory, 255 registers,
//                  2 blocks/SM       [D=16,K=128: 0.61 vs 0.44 with shared-memory accumulators]
//   LOGLIK       : D <= 6: Sigma_j in shared-memory columns + refill, 3 blocks/SM; D = 7..12: Sigma_j in tensor memory,
//                  3 blocks/SM; D = 13..16: Sigma_j in tensor memory, 255 registers
//   D = 17..24   : K >= 64: as 13..16 (unrolled, a few spilled words; Sigma_j through L1 when its columns
//                  would leave one block per SM, and for K > 128); otherwise the fallback
//   D = 21..32, and D = 17..20 with K < 64: the cooperative 8-lane-team kernels (dgmm_team.cuh)
// K = 257..512 at D <= 8 (the reference's Gaia runs: D = 7, K = 512): one group of 512 threads, i.e. the whole mixture in
// ONE launch (no component chunks, no second pass): Sigma_j in shared-memory columns, accumulators in tensor memory.
template <int D, int MODE, int ACCP>
inline PairVariant pair_wide512(int nz) {
    if (nz == NZ_FULL) return pair_variant<D, MODE, ST_SMEM, ACCP, GCLS_MWN, NZ_FULL, kPairWideThreads, 1>();
    if (nz == NZ_PACKED) return pair_variant<D, MODE, ST_SMEM, ACCP, GCLS_MWN, NZ_PACKED, kPairWideThreads, 1>();
    return pair_variant<D, MODE, ST_SMEM, ACCP, GCLS_MWN, NZ_DIAG, kPairWideThreads, 1>();
}

template <int D>
inline PairVariant pair_lookup(int mode, int which, int gcls, int nz) {
    if (gcls == GCLS_MWN512) {
        if constexpr (D <= kPairWideMaxD) {
            if (which != 0) return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
            if (mode == MODE_LOGLIK) return pair_wide512<D, MODE_LOGLIK, ST_GLOBAL>(nz);
            return pair_wide512<D, MODE_STATS, ST_TMEM>(nz);
        }
        return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
    }
    if (which == 2) {
        // universal fallback of the thread-per-pair family; above D = 16 the cooperative kernels
        // (dgmm_team.cuh) serve everything the unrolled variants do not
        if constexpr (D <= DGMM_UNROLL_D) {
            if (mode == MODE_LOGLIK)
                return pair_variant<D, MODE_LOGLIK, ST_GLOBAL, ST_GLOBAL, GCLS_ANY, NZ_ANY, kPairMaxThreads, 1>();
            return pair_variant<D, MODE_STATS, ST_GLOBAL, ST_GLOBAL, GCLS_ANY, NZ_ANY, kPairMaxThreads, 1>();
        }
        return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
    }
    if (which == 1) {  // alternative placements (tuning sweeps, tools/gpu_ab.sh); not compiled for packed noise
        if (nz == NZ_PACKED) return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
        if constexpr (D <= 12) {
            // (one variant, for A/B runs: the shared-memory / in-factorisation-refill placement that D >= 7 left)
            if (mode == MODE_LOGLIK && gcls == GCLS_MW4 && nz == NZ_FULL && D >= 7)
                return pair_variant<D, MODE_LOGLIK, ST_SMEM, ST_GLOBAL, GCLS_MW4, NZ_FULL, 128, 3>();
        } else if constexpr (D <= DGMM_UNROLL_D) {
            // D = 13..16 log-likelihood pass with the Sigma_j columns in shared memory (the placement before tensor memory)
            // (one variant, for A/B runs: K = 64..128, full noise)
            if (mode == MODE_LOGLIK && gcls == GCLS_MW4 && nz == NZ_FULL)
                return pair_variant<D, MODE_LOGLIK, ST_SMEM, ST_GLOBAL, GCLS_MW4, NZ_FULL, kPairMaxThreads, 1>();
        } else if constexpr (D > DGMM_UNROLL_D && D <= DGMM_PAIR_UNROLL_D) {
            // Sigma_j through L1 instead of shared-memory columns: the planner takes this one when the
            // columns would leave room for only one block per SM (two blocks of 4 warps beat one)
            if (mode == MODE_LOGLIK && gcls == GCLS_MW4 && nz == NZ_FULL)   // (A/B: Sigma_j through L1, the placement before)
                return pair_variant_wide_unrolled<D, MODE_LOGLIK, ST_GLOBAL, ST_GLOBAL, GCLS_MW4, NZ_FULL>();
            if (mode == MODE_LOGLIK) return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
            if constexpr (D <= 20) return pair_wide_unrolled<D, MODE_STATS, ST_TMEM>(gcls, nz, true);
        }
        return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
    }
    if (which != 0) return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
    // Accumulators in tensor memory need warp-uniform control flow: sub-warp groups (K < 32) keep
    // the register / shared-memory placements.
    if constexpr (D <= 12) {
        // log-likelihood pass, D = 7..12: Sigma_j column in tensor memory, no in-factorisation refill (gpurun A/B against the
        // shared-memory + refill placement: D=9,K=128 +9 %, D=10,K=64 +7 %, D=12,K=128 +9 %); D <= 6 and sub-warp groups keep it
        if constexpr (D >= 7) {
            if (mode == MODE_LOGLIK && gcls != GCLS_SUB) return pair_by_gcls<D, MODE_LOGLIK, ST_TMEM, ST_GLOBAL, 128, 3>(gcls, nz);
        }
        if (mode == MODE_LOGLIK) return pair_by_gcls<D, MODE_LOGLIK, ST_SMEM, ST_GLOBAL, 128, 3>(gcls, nz);
        if constexpr (D == 11) {
            if (gcls != GCLS_SUB) return pair_by_gcls<D, MODE_STATS, ST_SMEM, ST_TMEM, 128, 4>(gcls, nz);
        } else if constexpr (D == 12) {
            if (gcls != GCLS_SUB) return pair_by_gcls<D, MODE_STATS, ST_SMEM, ST_TMEM, 128, 3>(gcls, nz);
        }
        return pair_by_gcls<D, MODE_STATS, ST_SMEM, ST_REG, 128, 3>(gcls, nz);
    } else if constexpr (D <= DGMM_UNROLL_D) {
        // log-likelihood pass: Sigma_j column in the thread's tensor-memory row (tcgen05.ld: 9 x16-loads per point at D = 16
        // instead of 34 LDS.128; D=16,K=128: 21.5 -> 18.4 ms per 3e6 points, bit-identical); sub-warp groups keep shared memory
        if (mode == MODE_LOGLIK && gcls != GCLS_SUB)
            return pair_by_gcls<D, MODE_LOGLIK, ST_TMEM, ST_GLOBAL, kPairMaxThreads, 1>(gcls, nz);
        if (mode == MODE_LOGLIK) return pair_by_gcls<D, MODE_LOGLIK, ST_SMEM, ST_GLOBAL, kPairMaxThreads, 1>(gcls, nz);
        if (gcls != GCLS_SUB) return pair_by_gcls<D, MODE_STATS, ST_SMEM, ST_TMEM, kPairMaxThreads, 1>(gcls, nz);
        return pair_by_gcls<D, MODE_STATS, ST_GLOBAL, ST_SMEM, kPairMaxThreads, 1>(gcls, nz);
    } else if constexpr (D <= DGMM_PAIR_UNROLL_D) {
        if (mode == MODE_LOGLIK) return pair_wide_loglik_tmem<D>(gcls, nz);
        return pair_wide_unrolled<D, MODE_STATS, ST_TMEM>(gcls, nz);
    }
    return PairVariant{nullptr, 0, 0, 0, 0, 0, 0};
}

template <int D>
inline const void* full_lookup() {
    if constexpr (D <= DGMM_UNROLL_D)
        return reinterpret_cast<const void*>(&unrolled::full_kernel<D>);
    else
        return reinterpret_cast<const void*>(&rolled::full_kernel<D>);
}

template <int D>
inline const void* sample_lookup() {
    if constexpr (D <= DGMM_UNROLL_D)
        return reinterpret_cast<const void*>(&unrolled::sample_kernel<D>);
    else
        return reinterpret_cast<const void*>(&rolled::sample_kernel<D>);
}

template <int D>
inline const void* post_logprob_lookup() {
    if constexpr (D <= DGMM_UNROLL_D)
        return reinterpret_cast<const void*>(&unrolled::post_logprob_kernel<D>);
    else
        return reinterpret_cast<const void*>(&rolled::post_logprob_kernel<D>);
}

}  // namespace dgmm

#define DGMM_DECLARE_DIM(D)                                                  \
    extern "C++" dgmm::PairVariant dgmm_pair_lookup_d##D(int mode, int which, int gcls, int nz); \
    extern "C++" const void* dgmm_full_lookup_d##D();                        \
    extern "C++" const void* dgmm_sample_lookup_d##D();                      \
    extern "C++" const void* dgmm_post_logprob_lookup_d##D();

#define DGMM_DEFINE_DIM(D)                                                                      \
    dgmm::PairVariant dgmm_pair_lookup_d##D(int mode, int which, int gcls, int nz) {             \
        return dgmm::pair_lookup<D>(mode, which, gcls, nz);                                      \
    }                                                                                           \
    const void* dgmm_full_lookup_d##D() { return dgmm::full_lookup<D>(); }                      \
    const void* dgmm_sample_lookup_d##D() { return dgmm::sample_lookup<D>(); }                  \
    const void* dgmm_post_logprob_lookup_d##D() { return dgmm::post_logprob_lookup<D>(); }
