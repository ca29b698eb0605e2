// dgmm_team.cuh -- the cooperative ("team") kernels for D = 17..32.
//
// Above D = 16 the packed triangle of T_ij = Sigma_j + S_i (up to 528 floats) no longer fits one
// thread's registers, so a TEAM of 8 lanes owns one (point, component) pair: lane l keeps rows
// a = l, l + 8, l + 16, l + 24 of the lower triangle (row-cyclic: <= 80 registers at D = 32, balanced to
// 82 % useful slots), four teams per warp, up to 32 teams (= 32 components) per block, every team of a
// block on the same data point so that S_i / x_i are shared-memory broadcasts.
//
// The factorisation is written so that one step is a uniform rank-1 update whose only communication
// is ONE vector per team published through shared memory (no per-element shuffles, no transposes):
//
//   log-likelihood pass (team_loglik_kernel): right-looking LDL^T.  Step k: the lanes publish column k
//     (their rows' elements A[a][k]), one __syncwarp, every lane reads the column back as broadcast
//     128-bit loads and updates its rows' trailing elements.  Pivot and right-hand side travel by shuffle.
//     log p_ij from the pivots (log-det) and the forward-eliminated right-hand side (Mahalanobis form).
//
//   statistics pass (team_stats_kernel): the SWEEP operator (Gauss-Jordan on the symmetric matrix):
//     sweeping k = 0..D-1 turns T into -T^{-1} in place with the same D^3/2 multiply-adds as
//     LDL^T + triangular inverse, but every step updates ALL stored elements with the same formula
//       A[a][b] -= (A[a][k] / p) * A[k][b],   row/column k := A[.][k] / p,   A[k][k] := -1/p
//     so the work per step is identical for every lane (no shrinking trailing block, no distributed
//     back substitution, no symmetric copy).  The pivots are bit-identical to the LDL^T ones (same
//     trailing updates), the right-hand side swept along ends as u = T^{-1} d.  Then
//     omega * {1, u, u u^T + A} goes to the team's accumulators in TENSOR MEMORY (tcgen05.ld/st as a
//     per-thread scratchpad: 88 columns per lane at D = 32), which keeps the kernel at <= 128 registers.
//
// A block of the statistics pass covers a chunk of <= 32 components, so the logsumexp over ALL K
// components cannot be formed inside it: the statistics pass reads ll_i written by the log-likelihood
// pass (two passes; the first costs F_ll of the F_em + F_ll total, SURVEY.md 7.2 option (a)).
//
// Input streaming as in dgmm_pair.cuh: a tile of PT points is 3-4 contiguous runs -> cp.async.bulk (TMA)
// into an mbarrier ring; the tile's noise matrices are then re-packed once into rows of RS floats
// (RS/4 odd) so that the 8 lanes of a team read their rows with conflict-free 128-bit loads.
//
// Replaces deconv/gmm/deconv_gmm.py:39-107 / deconv/gmm/sgd_deconv_gmm.py:14-28 for D > 16.
#pragma once

#include "dgmm_pair.cuh"
#include "dgmm_team_geo.cuh"

namespace dgmm {
namespace team {

using unrolled::static_for;

// ---------------------------------------------------------------------------------------------------
// per-pair pieces
// ---------------------------------------------------------------------------------------------------

// A := Sigma_j (global, register order) [+ S_i rows from the padded tile], y := x_i - mu_j for my rows
template <int D, int NZ>
__device__ __forceinline__ void form_T(float (&A)[Geo<D>::NREG], float (&y)[Geo<D>::R], const float4* __restrict__ sig4,
                                       const float* __restrict__ mu_r, const float* __restrict__ Sp,
                                       const float* __restrict__ Xp, int l) {
    using G = Geo<D>;
#pragma unroll
    for (int q = 0; q < G::NQ; ++q) {
        const float4 v = __ldg(sig4 + q * kP);
        A[4 * q] = v.x; A[4 * q + 1] = v.y; A[4 * q + 2] = v.z; A[4 * q + 3] = v.w;
    }
    if constexpr (NZ == TNZ_MATRIX) {
        static_for<G::R>([&](auto rc) {
            constexpr int r = decltype(rc)::value;
            const float4* row = reinterpret_cast<const float4*>(Sp + (l + 8 * r) * G::RS);
#pragma unroll
            for (int q = 0; q < geo_wp(D, r) / 4; ++q) {
                const float4 v = row[q];
                A[geo_off(D, r) + 4 * q] += v.x; A[geo_off(D, r) + 4 * q + 1] += v.y;
                A[geo_off(D, r) + 4 * q + 2] += v.z; A[geo_off(D, r) + 4 * q + 3] += v.w;
            }
        });
    }
    static_for<G::R>([&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const int a = l + 8 * r;
        y[r] = (8 * r + 8 <= D || a < D) ? Xp[a < D ? a : 0] - mu_r[r] : 0.f;
    });
}

// One step of the right-looking LDL^T (log-likelihood pass).  vb: this team's published vector of
// parity k & 1.  On return the trailing rows are updated, lg / maha carry the pivot's contribution.
template <int D, int k>
__device__ __forceinline__ void ldl_step(float (&A)[Geo<D>::NREG], float (&y)[Geo<D>::R], float& lg, float& maha,
                                         float* __restrict__ vb, int l, int src_lane) {
    using G = Geo<D>;
    constexpr int rk = k >> 3;
    // publish column k: my rows a > k
    static_for<G::R - rk>([&](auto rc) {
        constexpr int r = rk + decltype(rc)::value;
        if constexpr (k < geo_w(D, r)) {
            const int a = l + 8 * r;
            if ((r > rk || a > k) && (8 * r + 8 <= D || a < D)) vb[a] = A[geo_off(D, r) + k];
        }
    });
    __syncwarp();
    const float p = __shfl_sync(0xffffffffu, A[geo_off(D, rk) + k], src_lane);
    const float yk = __shfl_sync(0xffffffffu, y[rk], src_lane);
    const float inv = mufu_rcp(p);
    lg += mufu_lg2(p);
    const float t = yk * inv;
    maha = fmaf(t, yk, maha);
    if constexpr (k + 1 < D) {
        float g[G::R];
        static_for<G::R - rk>([&](auto rc) {
            constexpr int r = rk + decltype(rc)::value;
            const float c = vb[l + 8 * r];
            g[r] = -c * inv;
            y[r] = fmaf(-t, c, y[r]);
        });
        // trailing update: rows of slots >= rk, columns k+1 .. w(r)-1, column values as broadcast float4
        constexpr int q0 = (k + 1) / 4;
        constexpr int q1 = (D + 3) / 4;
        static_for<q1 - q0>([&](auto qc) {
            constexpr int q = q0 + decltype(qc)::value;
            const float4 v4 = *reinterpret_cast<const float4*>(vb + 4 * q);
            const float v[4] = {v4.x, v4.y, v4.z, v4.w};
            static_for<G::R - rk>([&](auto rc) {
                constexpr int r = rk + decltype(rc)::value;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int b = 4 * q + c;   // constant after unrolling
                    if (b > k && b < geo_w(D, r)) A[geo_off(D, r) + b] = fmaf(g[r], v[c], A[geo_off(D, r) + b]);
                }
            });
        });
    }
}

// One sweep step (statistics pass): the whole stored triangle is updated; row / column k become
// A[.][k] / p, the pivot position -1/p, the right-hand side is swept along.
template <int D, int k>
__device__ __forceinline__ void sweep_step(float (&A)[Geo<D>::NREG], float (&y)[Geo<D>::R], float& lg, float& maha,
                                           float* __restrict__ vb, int l, int src_lane) {
    using G = Geo<D>;
    constexpr int rk = k >> 3;
    constexpr int lo = k & 7;
    const bool owner = (l == lo);
    // publish the symmetric row k: the owner of row k stores A[k][0..k] (vector stores for whole groups) ...
    {
        constexpr int nq = (k + 1) / 4;
        if (owner) {
#pragma unroll
            for (int q = 0; q < nq; ++q)
                *reinterpret_cast<float4*>(vb + 4 * q) =
                    make_float4(A[geo_off(D, rk) + 4 * q], A[geo_off(D, rk) + 4 * q + 1], A[geo_off(D, rk) + 4 * q + 2],
                                A[geo_off(D, rk) + 4 * q + 3]);
#pragma unroll
            for (int b = 4 * nq; b <= k; ++b) vb[b] = A[geo_off(D, rk) + b];
        }
    }
    // ... and every lane its rows' elements of column k (rows a > k)
    static_for<G::R - rk>([&](auto rc) {
        constexpr int r = rk + decltype(rc)::value;
        if constexpr (k < geo_w(D, r)) {
            const int a = l + 8 * r;
            if ((r > rk || a > k) && (8 * r + 8 <= D || a < D)) vb[a] = A[geo_off(D, r) + k];
        }
    });
    __syncwarp();
    const float p = __shfl_sync(0xffffffffu, A[geo_off(D, rk) + k], src_lane);
    const float yk = __shfl_sync(0xffffffffu, y[rk], src_lane);
    const float inv = mufu_rcp(p);
    lg += mufu_lg2(p);
    const float t = yk * inv;
    maha = fmaf(t, yk, maha);
    float g[G::R];    // -A[a][k] / p for my rows
    float gk[G::R];   // +A[a][k] / p (the new column k)
    static_for<G::R>([&](auto rc) {
        constexpr int r = decltype(rc)::value;
        const float c = vb[l + 8 * r];
        gk[r] = c * inv;
        g[r] = -gk[r];
        y[r] = fmaf(-t, c, y[r]);
    });
    y[rk] = owner ? t : y[rk];
    const float cf = owner ? inv : g[rk];   // the owner's row k becomes A[k][b] / p
    constexpr int q1 = (D + 3) / 4;
    static_for<q1>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        const float4 v4 = *reinterpret_cast<const float4*>(vb + 4 * q);
        const float v[4] = {v4.x, v4.y, v4.z, v4.w};
        static_for<G::R>([&](auto rc) {
            constexpr int r = decltype(rc)::value;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int b = 4 * q + c;   // constant after unrolling
                if (b < geo_w(D, r) && b != k) {
                    if (r == rk) {
                        const float base = owner ? 0.f : A[geo_off(D, r) + b];
                        A[geo_off(D, r) + b] = fmaf(cf, v[c], base);
                    } else {
                        A[geo_off(D, r) + b] = fmaf(g[r], v[c], A[geo_off(D, r) + b]);
                    }
                }
            }
        });
    });
    // column k
    static_for<G::R - rk>([&](auto rc) {
        constexpr int r = rk + decltype(rc)::value;
        if constexpr (k < geo_w(D, r)) {
            if (r == rk) A[geo_off(D, r) + k] = owner ? -inv : gk[r];
            else A[geo_off(D, r) + k] = gk[r];
        }
    });
}

// Shared set-up of both kernels: ring, padded noise buffer, published vectors.
struct TeamSmem {
    uint64_t* full_bar;
    uint32_t* tmem_slot;
    unsigned char* ring;
    float* spad;    // [PT][ROWS][RS]
    float* vbuf;    // [warps][2][kTW * kVS]
    float* lse;     // [PT][kMaxTB][2]   (log-likelihood pass)
};

template <int D>
__device__ __forceinline__ TeamSmem team_carve(unsigned char* smem, const TeamParams& p, int warps) {
    TeamSmem s;
    s.full_bar = reinterpret_cast<uint64_t*>(smem);
    s.tmem_slot = reinterpret_cast<uint32_t*>(smem + 64);
    s.ring = smem + kTeamHeader;
    s.spad = reinterpret_cast<float*>(s.ring + (size_t)p.NST * p.stage_bytes);
    const bool has_mat = (p.noise == NOISE_FULL || p.noise == NOISE_DIAG || p.noise == NOISE_PACKED);
    s.vbuf = s.spad + (has_mat ? p.PT * Geo<D>::ROWS * Geo<D>::RS : 0);
    s.lse = s.vbuf + warps * 2 * kTW * kVS;
    return s;
}

// Tile staging shared by both kernels: waits for (or cooperatively loads) the tile, then re-packs the
// noise matrices into padded rows.  Returns pointers to the tile's X / g / ll runs.
template <int D>
__device__ __forceinline__ void team_issue_tile(const TeamParams& p, const TeamSmem& sm, long long tile, int stage,
                                                bool has_g, bool has_ll) {
    const long long i0 = tile * p.PT;
    unsigned char* st = sm.ring + (size_t)stage * p.stage_bytes;
    const uint32_t bS = (uint32_t)(p.PT * p.n_s * 4), bX = (uint32_t)(p.PT * D * 4), bG = (uint32_t)(p.PT * 4);
    mbar_expect_tx(&sm.full_bar[stage], bS + bX + (has_g ? bG : 0u) + (has_ll ? bG : 0u));
    if (bS) bulk_g2s(st, p.S + (size_t)i0 * p.n_s, bS, &sm.full_bar[stage]);
    bulk_g2s(st + p.off_X, p.X + (size_t)i0 * D, bX, &sm.full_bar[stage]);
    if (has_g) bulk_g2s(st + p.off_g, p.g + i0, bG, &sm.full_bar[stage]);
    if (has_ll) bulk_g2s(st + p.off_ll, p.ll + i0, bG, &sm.full_bar[stage]);
}

template <int D>
__device__ __forceinline__ void team_stage_tile(const TeamParams& p, const TeamSmem& sm, long long tile, long long it,
                                                bool bulk, bool has_g, bool has_ll, int npts) {
    const int tid = threadIdx.x, TPB = blockDim.x;
    const int stage = (int)(it % p.NST);
    unsigned char* st = sm.ring + (size_t)stage * p.stage_bytes;
    const long long i0 = tile * p.PT;
    if (bulk) {
        mbar_wait(&sm.full_bar[stage], (uint32_t)((it / p.NST) & 1));
    } else {
        float* wS = reinterpret_cast<float*>(st);
        float* wX = reinterpret_cast<float*>(st + p.off_X);
        float* wG = reinterpret_cast<float*>(st + p.off_g);
        float* wL = reinterpret_cast<float*>(st + p.off_ll);
        for (int idx = tid; idx < p.PT * p.n_s; idx += TPB) wS[idx] = idx < npts * p.n_s ? p.S[(size_t)i0 * p.n_s + idx] : 0.f;
        for (int idx = tid; idx < p.PT * D; idx += TPB) wX[idx] = idx < npts * D ? p.X[(size_t)i0 * D + idx] : 0.f;
        if (has_g)
            for (int idx = tid; idx < p.PT; idx += TPB) wG[idx] = idx < npts ? p.g[i0 + idx] : 0.f;
        if (has_ll)
            for (int idx = tid; idx < p.PT; idx += TPB) wL[idx] = idx < npts ? p.ll[i0 + idx] : 0.f;
        __syncthreads();
    }
    // re-pack: row a of point pt -> spad[(pt * ROWS + a) * RS + 0..D)
    const float* raw = reinterpret_cast<const float*>(st);
    if (p.noise == NOISE_FULL) {
        for (int idx = tid; idx < p.PT * (D * D); idx += TPB) {
            const int pt = idx / (D * D), rem = idx - pt * (D * D);
            const int a = rem / D, b = rem - a * D;
            sm.spad[(pt * Geo<D>::ROWS + a) * Geo<D>::RS + b] = raw[idx];
        }
    } else if (p.noise == NOISE_DIAG) {
        for (int idx = tid; idx < p.PT * D; idx += TPB) {
            const int pt = idx / D, a = idx - pt * D;
            sm.spad[(pt * Geo<D>::ROWS + a) * Geo<D>::RS + a] = raw[idx];
        }
    } else if (p.noise == NOISE_PACKED) {
        constexpr int ND = D * (D + 1) / 2;
        for (int idx = tid; idx < p.PT * ND; idx += TPB) {
            const int pt = idx / ND, e = idx - pt * ND;
            const int a = tri_row(e), b = e - tri(a, 0);
            sm.spad[(pt * Geo<D>::ROWS + a) * Geo<D>::RS + b] = raw[idx];
        }
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------
// log-likelihood pass
// ---------------------------------------------------------------------------------------------------
template <int D, int NZ>
__global__ void __launch_bounds__(256, 2) team_loglik_kernel(const TeamParams p) {
    using G = Geo<D>;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, TPB = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, warps = TPB >> 5;
    const int l = lane & 7, tw = lane >> 3;
    const int team_id = warp * kTW + tw;
    const int src_base = lane & 24;
    const TeamSmem sm = team_carve<D>(smem, p, warps);
    float* vb0 = sm.vbuf + (warp * 2) * (kTW * kVS) + tw * kVS;
    float* vb1 = vb0 + kTW * kVS;

    // zero the padded noise buffer (pad columns / rows stay zero) and the published vectors
    if (NZ == TNZ_MATRIX)
        for (int idx = tid; idx < p.PT * G::ROWS * G::RS; idx += TPB) sm.spad[idx] = 0.f;
    for (int idx = tid; idx < warps * 2 * kTW * kVS; idx += TPB) sm.vbuf[idx] = 0.f;
    if (tid == 0 && p.use_bulk) {
        for (int s = 0; s < p.NST; ++s) mbar_init(&sm.full_bar[s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    auto tile_is_bulk = [&](long long tile) -> bool { return p.use_bulk && (tile + 1) * (long long)p.PT <= p.N; };
    const long long first = blockIdx.x, step = gridDim.x;
    if (tid == 0)
        for (int s = 0; s < p.NST; ++s) {
            const long long t = first + (long long)s * step;
            if (t < p.n_tiles && tile_is_bulk(t)) team_issue_tile<D>(p, sm, t, s, false, false);
        }

    double ll_acc = 0.0, n_acc = 0.0;
    int bad_pairs = 0;
    long long it = 0;
    for (long long tile = first; tile < p.n_tiles; tile += step, ++it) {
        const int stage = (int)(it % p.NST);
        const long long i0 = tile * p.PT;
        const int npts = (int)((p.N - i0 < p.PT) ? (p.N - i0) : p.PT);
        team_stage_tile<D>(p, sm, tile, it, tile_is_bulk(tile), false, false, npts);
        const float* tX = reinterpret_cast<const float*>(sm.ring + (size_t)stage * p.stage_bytes + p.off_X);

        for (int pt = 0; pt < npts; ++pt) {
            float m_run = -1e30f, s_run = 0.f;
            const float* Sp = sm.spad + pt * (G::ROWS * G::RS);
            const float* Xp = tX + pt * D;
#pragma unroll 1
            for (int ch = 0; ch < p.n_chunks; ++ch) {
                const int j = ch * p.TB + team_id;   // team_id < TB for every team of the block
                float A[G::NREG];
                float y[G::R];
                float mu_r[G::R];
#pragma unroll
                for (int r = 0; r < G::R; ++r) mu_r[r] = __ldg(p.muT + ((size_t)j * G::R + r) * kP + l);
                form_T<D, NZ>(A, y, reinterpret_cast<const float4*>(p.sigT) + (size_t)j * G::NQ * kP + l, mu_r, Sp, Xp, l);
                float lg = 0.f, maha = 0.f;
                static_for<D>([&](auto kc) {
                    constexpr int k = decltype(kc)::value;
                    ldl_step<D, k>(A, y, lg, maha, (k & 1) ? vb1 : vb0, l, src_base | (k & 7));
                });
                const float c2 = __ldg(p.cj + j) * kLog2e;
                float lp2 = fmaf(-0.5f * kLog2e, maha, fmaf(-0.5f, lg, c2));
                if (j >= p.K) lp2 = -INFINITY;   // padding component of the last chunk
                if (l == 0 && j < p.K && !(lp2 < INFINITY)) ++bad_pairs;
                // online logsumexp over this team's components (NaN / +inf poison the sum)
                const float m_new = fmaxf(m_run, lp2);
                s_run = fmaf(s_run, mufu_ex2(m_run - m_new), mufu_ex2(lp2 - m_new));
                m_run = m_new;
                __syncwarp();
            }
            if (l == 0) {
                sm.lse[(pt * kMaxTB + team_id) * 2] = m_run;
                sm.lse[(pt * kMaxTB + team_id) * 2 + 1] = s_run;
            }
        }
        __syncthreads();
        if (tid == 0) {
            const long long nt = tile + (long long)p.NST * step;
            if (nt < p.n_tiles && tile_is_bulk(nt)) team_issue_tile<D>(p, sm, nt, stage, false, false);
        }
        // per-tile logsumexp over the teams: one warp per point
        for (int pt = warp; pt < npts; pt += warps) {
            const bool have = lane < p.TB;
            const float m = have ? sm.lse[(pt * kMaxTB + lane) * 2] : -1e30f;
            const float s = have ? sm.lse[(pt * kMaxTB + lane) * 2 + 1] : 0.f;
            float M = m;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
            float tot = s * mufu_ex2(m - M);   // NaN in s propagates
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            const float ll2 = M + mufu_lg2(tot);
            const bool fin = fabsf(ll2) <= FLT_MAX;
            const float ll = fin ? ll2 * kLn2 : -INFINITY;
            if (lane == 0) {
                p.ll[i0 + pt] = ll;
                if (p.out_ll != nullptr) p.out_ll[i0 + pt] = ll;
                ll_acc += (double)ll;
                n_acc += 1.0;
            }
        }
        __syncthreads();   // lse / spad are rewritten by the next tile
    }
    if (bad_pairs) atomicAdd(p.nonpd, bad_pairs);
    // block sums (lane 0 of every warp holds partials)
    __shared__ double dred[2][8];
    if (lane == 0) { dred[0][warp] = ll_acc; dred[1][warp] = n_acc; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < warps; ++w) { a += dred[0][w]; b += dred[1][w]; }
        p.part_ll[2 * blockIdx.x] = a;
        p.part_ll[2 * blockIdx.x + 1] = b;
    }
}

// ---------------------------------------------------------------------------------------------------
// statistics pass
// ---------------------------------------------------------------------------------------------------
// Register cap of the statistics pass: two blocks per SM at 128 registers up to D = 26; from D = 27 the 80+ matrix
// registers per lane spill at that cap (4.3e8 local loads per launch at D = 32) and ONE block at 255 registers is
// faster (gpurun A/B, N = 262 144: D=32,K=256 179.8 -> 164.6 ms; D=28,K=128 74.0 -> 71.5; D=24,K=128 42.7 -> 46.4).
constexpr int team_stats_min_blocks(int D) { return D >= 27 ? 1 : 2; }
template <int D, int NZ>
__global__ void __launch_bounds__(256, team_stats_min_blocks(D)) team_stats_kernel(const TeamParams p) {
    using G = Geo<D>;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, TPB = blockDim.x;
    const int lane = tid & 31, warp = tid >> 5, warps = TPB >> 5;
    const int l = lane & 7, tw = lane >> 3;
    const int team_id = warp * kTW + tw;
    const int src_base = lane & 24;
    const TeamSmem sm = team_carve<D>(smem, p, warps);
    float* vb0 = sm.vbuf + (warp * 2) * (kTW * kVS) + tw * kVS;
    float* vb1 = vb0 + kTW * kVS;
    const bool has_g = p.g != nullptr;

    const int chunk = blockIdx.x % p.n_chunks;
    const int slice = blockIdx.x / p.n_chunks;
    const int j = chunk * p.TB + team_id;
    const bool live = j < p.K;

    if (NZ == TNZ_MATRIX)
        for (int idx = tid; idx < p.PT * G::ROWS * G::RS; idx += TPB) sm.spad[idx] = 0.f;
    for (int idx = tid; idx < warps * 2 * kTW * kVS; idx += TPB) sm.vbuf[idx] = 0.f;
    if (tid == 0 && p.use_bulk) {
        for (int s = 0; s < p.NST; ++s) mbar_init(&sm.full_bar[s], 1);
        mbar_fence_init();
    }
    // accumulators in tensor memory: my row = TMEM lane 32 * (warp % 4) + lane, NACC columns per warp quad
    if (warp == 0) tmem_alloc(sm.tmem_slot, p.tmem_cols);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tacc = *sm.tmem_slot + ((uint32_t)(warp & 3) << 21) + (uint32_t)((warp >> 2) * G::NACC);
    {
        float z8[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) z8[c] = 0.f;
#pragma unroll
        for (int c8 = 0; c8 < G::NACC; c8 += 8) tmem_st8(tacc + c8, z8);
        tmem_wait_st();
    }

    float mu_r[G::R];
#pragma unroll
    for (int r = 0; r < G::R; ++r) mu_r[r] = __ldg(p.muT + ((size_t)j * G::R + r) * kP + l);
    const float c2 = __ldg(p.cj + j) * kLog2e;
    const float my_floor = live ? p.resp_floor : 0.f;
    const float4* sig4 = reinterpret_cast<const float4*>(p.sigT) + (size_t)j * G::NQ * kP + l;

    auto tile_is_bulk = [&](long long tile) -> bool { return p.use_bulk && (tile + 1) * (long long)p.PT <= p.N; };
    const long long first = slice, step = p.n_slices;
    if (tid == 0)
        for (int s = 0; s < p.NST; ++s) {
            const long long t = first + (long long)s * step;
            if (t < p.n_tiles && tile_is_bulk(t)) team_issue_tile<D>(p, sm, t, s, has_g, true);
        }

    double g_acc = 0.0;
    long long it = 0;
    for (long long tile = first; tile < p.n_tiles; tile += step, ++it) {
        const int stage = (int)(it % p.NST);
        const long long i0 = tile * p.PT;
        const int npts = (int)((p.N - i0 < p.PT) ? (p.N - i0) : p.PT);
        team_stage_tile<D>(p, sm, tile, it, tile_is_bulk(tile), has_g, true, npts);
        unsigned char* st = sm.ring + (size_t)stage * p.stage_bytes;
        const float* tX = reinterpret_cast<const float*>(st + p.off_X);
        const float* tG = reinterpret_cast<const float*>(st + p.off_g);
        const float* tL = reinterpret_cast<const float*>(st + p.off_ll);

        for (int pt = 0; pt < npts; ++pt) {
            const float ll2 = tL[pt] * kLog2e;
            const float gi = has_g ? tG[pt] : 1.0f;
            if (tid == 0) g_acc += (double)gi;
            if (!(fabsf(ll2) <= FLT_MAX)) continue;   // non-PD pair or total underflow at this point (block-uniform)
            float A[G::NREG];
            float y[G::R];
            form_T<D, NZ>(A, y, sig4, mu_r, sm.spad + pt * (G::ROWS * G::RS), tX + pt * D, l);
            float lg = 0.f, maha = 0.f;
            static_for<D>([&](auto kc) {
                constexpr int k = decltype(kc)::value;
                sweep_step<D, k>(A, y, lg, maha, (k & 1) ? vb1 : vb0, l, src_base | (k & 7));
            });
            const float lp2 = fmaf(-0.5f * kLog2e, maha, fmaf(-0.5f, lg, c2));
            const float om = live ? gi * (mufu_ex2(lp2 - ll2) + my_floor) : 0.f;
            // publish u = T^{-1} d (my rows), then omega * {u u^T + A, u, 1} into my accumulator columns
            float* vu = (D & 1) ? vb1 : vb0;
            static_for<G::R>([&](auto rc) {
                constexpr int r = decltype(rc)::value;
                const int a = l + 8 * r;
                if (8 * r + 8 <= D || a < D) vu[a] = y[r];
            });
            __syncwarp();
            float ty[G::R];
#pragma unroll
            for (int r = 0; r < G::R; ++r) ty[r] = om * y[r];
            static_for<G::NACC / 8>([&](auto cc) {
                constexpr int c8 = 8 * decltype(cc)::value;
                float v[8];
                tmem_ld8(tacc + c8, v);
                // u[b] of the two float4 groups of these columns (broadcast reads)
                float ub[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int e0 = c8 + 4 * h;   // constant after unrolling
                    if (e0 < G::NREG) {
                        const int b0 = e0 - geo_off(D, geo_slot(D, e0));
                        const float4 u4 = *reinterpret_cast<const float4*>(vu + b0);
                        ub[4 * h] = u4.x; ub[4 * h + 1] = u4.y; ub[4 * h + 2] = u4.z; ub[4 * h + 3] = u4.w;
                    } else {
                        ub[4 * h] = ub[4 * h + 1] = ub[4 * h + 2] = ub[4 * h + 3] = 0.f;
                    }
                }
                tmem_wait_ld();
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int e = c8 + c;   // constant after unrolling
                    if (e < G::NREG) {
                        const int r = geo_slot(D, e);
                        const int b = e - geo_off(D, r);
                        if (b < geo_w(D, r)) v[c] = fmaf(ty[r], ub[c], fmaf(om, A[e], v[c]));
                    } else if (e < G::NREG + G::R) {
                        v[c] += ty[e - G::NREG];
                    } else if (e == G::NREG + G::R) {
                        v[c] += om;
                    }
                }
                tmem_st8(tacc + c8, v);
            });
            __syncwarp();
        }
        __syncthreads();
        if (tid == 0) {
            const long long nt = tile + (long long)p.NST * step;
            if (nt < p.n_tiles && tile_is_bulk(nt)) team_issue_tile<D>(p, sm, nt, stage, has_g, true);
        }
    }

    if (tid == 0) {
        p.part_ll[2 * blockIdx.x] = (chunk == 0) ? g_acc : 0.0;
        p.part_ll[2 * blockIdx.x + 1] = 0.0;
    }
    // accumulators -> part[blk][team][canonical packed order s0 | s1 | S2 lower]
    float* out = p.part + ((size_t)blockIdx.x * p.TB + team_id) * G::NS;
    tmem_wait_st();
    static_for<G::NACC / 8>([&](auto cc) {
        constexpr int c8 = 8 * decltype(cc)::value;
        float v[8];
        tmem_ld8(tacc + c8, v);
        tmem_wait_ld();
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int e = c8 + c;
            if (e < G::NREG) {
                const int r = geo_slot(D, e);
                const int b = e - geo_off(D, r);
                const int a = l + 8 * r;
                if (b < geo_w(D, r) && b <= a && a < D) out[1 + D + a * (a + 1) / 2 + b] = v[c];
            } else if (e < G::NREG + G::R) {
                const int a = l + 8 * (e - G::NREG);
                if (a < D) out[1 + a] = v[c];
            } else if (e == G::NREG + G::R) {
                if (l == 0) out[0] = v[c];
            }
        }
    });
    tmem_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(*sm.tmem_slot, p.tmem_cols);
}

}  // namespace team

}  // namespace dgmm

#define DGMM_DEFINE_TEAM_DIM(D)                                                                            \
    dgmm::TeamKernels dgmm_team_lookup_d##D() {                                                            \
        dgmm::TeamKernels t;                                                                               \
        t.loglik[0] = reinterpret_cast<const void*>(&dgmm::team::team_loglik_kernel<D, dgmm::team::TNZ_MATRIX>); \
        t.loglik[1] = reinterpret_cast<const void*>(&dgmm::team::team_loglik_kernel<D, dgmm::team::TNZ_NONE>);   \
        t.stats[0] = reinterpret_cast<const void*>(&dgmm::team::team_stats_kernel<D, dgmm::team::TNZ_MATRIX>);   \
        t.stats[1] = reinterpret_cast<const void*>(&dgmm::team::team_stats_kernel<D, dgmm::team::TNZ_NONE>);     \
        return t;                                                                                          \
    }
