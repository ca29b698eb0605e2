// dgmm_tables.cuh -- per-D kernel tables.  Each dgmm_inst_dXX.cu instantiates the kernels
// of one dimension (so that the 32 dimensions compile in parallel) and exports one lookup
// function; dgmm_api.cu launches through the returned pointers with cudaLaunchKernel.
#pragma once

#include "dgmm_full.cuh"
#include "dgmm_pair.cuh"

namespace dgmm {


// Returns the kernel for (mode, sigma placement, accumulator placement) or nullptr when
// that combination is not compiled for this D.
typedef const void* (*pair_lookup_fn)(int mode, int sigp, int accp, int rc);
typedef const void* (*full_lookup_fn)();

// Register-cap classes (launch bounds): rc 0 = <=256 threads, 255 registers;
// rc 1 = <=128 threads, 3 blocks/SM (168 registers); rc 2 = <=128 threads, 4 blocks/SM (128).
template <int D, int MODE, int SIGP, int ACCP, int RC>
inline const void* pair_ptr() {
    constexpr int MAXT = RC == 0 ? kPairMaxThreads : 128;
    constexpr int MINB = RC == 0 ? 1 : (RC == 1 ? 3 : 4);
    if constexpr (D <= DGMM_UNROLL_D)
        return reinterpret_cast<const void*>(&unrolled::pair_kernel<D, MODE, SIGP, ACCP, MAXT, MINB>);
    else
        return reinterpret_cast<const void*>(&rolled::pair_kernel<D, MODE, SIGP, ACCP, MAXT, MINB>);
}

template <int D, int MODE, int SIGP, int ACCP>
inline const void* pair_ptr_rc(int rc, int max_rc) {
    if (rc > max_rc) return nullptr;
    if (rc == 0) return pair_ptr<D, MODE, SIGP, ACCP, 0>();
    if constexpr (D <= DGMM_UNROLL_D) {
        if (rc == 1) return pair_ptr<D, MODE, SIGP, ACCP, 1>();
    }
    if constexpr (D <= 12) {
        if (rc == 2) return pair_ptr<D, MODE, SIGP, ACCP, 2>();
    }
    return nullptr;
}

// Which placements exist for which D (keep in sync with make_plan() in dgmm_api.cu):
//   D <= 12      : Sigma {SMEM, REG} x acc {REG (Sigma SMEM only), SMEM}; rc 0..2
//   13 <= D <= 16: Sigma {SMEM, GLOBAL} x acc {SMEM};                     rc 0..1
//   every D      : GLOBAL/GLOBAL fallback (rc 0); D > 16 has nothing else (rolled loops)
template <int D>
inline const void* pair_lookup(int mode, int sigp, int accp, int rc) {
    if (mode == MODE_LOGLIK) {
        if constexpr (D <= 12) {
            if (sigp == ST_REG) return pair_ptr_rc<D, MODE_LOGLIK, ST_REG, ST_GLOBAL>(rc, 2);
        }
        if constexpr (D <= DGMM_UNROLL_D) {
            if (sigp == ST_SMEM) return pair_ptr_rc<D, MODE_LOGLIK, ST_SMEM, ST_GLOBAL>(rc, 2);
        }
        if (sigp == ST_GLOBAL) return pair_ptr_rc<D, MODE_LOGLIK, ST_GLOBAL, ST_GLOBAL>(rc, 0);
        return nullptr;
    }
    if constexpr (D <= 12) {
        if (sigp == ST_SMEM && accp == ST_REG) return pair_ptr_rc<D, MODE_STATS, ST_SMEM, ST_REG>(rc, 2);
        if (sigp == ST_REG && accp == ST_SMEM) return pair_ptr_rc<D, MODE_STATS, ST_REG, ST_SMEM>(rc, 1);
    }
    if constexpr (D <= DGMM_UNROLL_D) {
        if (sigp == ST_SMEM && accp == ST_SMEM) return pair_ptr_rc<D, MODE_STATS, ST_SMEM, ST_SMEM>(rc, 2);
    }
    if constexpr (D > 12 && D <= DGMM_UNROLL_D) {
        if (sigp == ST_GLOBAL && accp == ST_SMEM) return pair_ptr_rc<D, MODE_STATS, ST_GLOBAL, ST_SMEM>(rc, 1);
    }
    if (sigp == ST_GLOBAL && accp == ST_GLOBAL) return pair_ptr_rc<D, MODE_STATS, ST_GLOBAL, ST_GLOBAL>(rc, 0);
    return nullptr;
}

template <int D>
inline const void* full_lookup() {
    if constexpr (D <= DGMM_UNROLL_D)
        return reinterpret_cast<const void*>(&unrolled::full_kernel<D>);
    else
        return reinterpret_cast<const void*>(&rolled::full_kernel<D>);
}

}  // namespace dgmm

#define DGMM_DECLARE_DIM(D)                                                  \
    extern "C++" const void* dgmm_pair_lookup_d##D(int mode, int sigp, int accp, int rc); \
    extern "C++" const void* dgmm_full_lookup_d##D();

#define DGMM_DEFINE_DIM(D)                                                                      \
    const void* dgmm_pair_lookup_d##D(int mode, int sigp, int accp, int rc) {                   \
        return dgmm::pair_lookup<D>(mode, sigp, accp, rc);                                       \
    }                                                                                           \
    const void* dgmm_full_lookup_d##D() { return dgmm::full_lookup<D>(); }
