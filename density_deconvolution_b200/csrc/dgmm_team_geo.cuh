// dgmm_team_geo.cuh -- geometry, parameter block and lookup type of the cooperative (team) kernels
// (dgmm_team.cuh); shared by the kernels and the launcher in dgmm_api.cu.
#pragma once

#include "dgmm_pair.cuh"

namespace dgmm {
namespace team {

constexpr int kP = 8;        // lanes per team
constexpr int kTW = 4;       // teams per warp
constexpr int kVS = 40;      // floats between the published vectors of a warp's teams (banks 0, 8, 16, 24)
constexpr int kMaxTB = 32;   // teams per block (blocks of <= 256 threads)

// Geometry of the row-cyclic register layout for dimension D (host + device).
struct GeoRT {
    int R, NREG, RS, ROWS, NACC;
    int w[4], wp[4], off[4];
};
__host__ __device__ constexpr int geo_R(int D) { return (D + 7) / 8; }
__host__ __device__ constexpr int geo_w(int D, int r) { return 8 * (r + 1) < D ? 8 * (r + 1) : D; }   // columns of slot r
__host__ __device__ constexpr int geo_wp(int D, int r) { return (geo_w(D, r) + 3) & ~3; }                // padded to float4
__host__ __device__ constexpr int geo_off(int D, int r) {
    int o = 0;
    for (int i = 0; i < r; ++i) o += geo_wp(D, i);
    return o;
}
__host__ __device__ constexpr int geo_nreg(int D) { return geo_off(D, geo_R(D)); }
// slot of register e (e < geo_nreg(D))
__host__ __device__ constexpr int geo_slot(int D, int e) {
    int r = 0;
    while (r + 1 < geo_R(D) && e >= geo_off(D, r + 1)) ++r;
    return r;
}
// padded row stride of the re-packed noise matrices: a multiple of 4 floats with RS/4 odd
__host__ __device__ constexpr int geo_rs(int D) {
    int m = (D + 3) / 4;
    if (m % 2 == 0) ++m;
    return 4 * m;
}
__host__ __device__ constexpr int geo_rows(int D) { return 8 * geo_R(D); }
// accumulator columns per lane: [A regs | s1 (R) | s0], rounded up to 8
__host__ __device__ constexpr int geo_nacc(int D) { return (geo_nreg(D) + geo_R(D) + 1 + 7) & ~7; }

inline GeoRT geo_rt(int D) {
    GeoRT g;
    g.R = geo_R(D); g.NREG = geo_nreg(D); g.RS = geo_rs(D); g.ROWS = geo_rows(D); g.NACC = geo_nacc(D);
    for (int r = 0; r < 4; ++r) { g.w[r] = geo_w(D, r); g.wp[r] = geo_wp(D, r); g.off[r] = geo_off(D, r); }
    return g;
}

template <int D>
struct Geo {
    static constexpr int R = geo_R(D);
    static constexpr int NREG = geo_nreg(D);
    static constexpr int NQ = NREG / 4;
    static constexpr int RS = geo_rs(D);
    static constexpr int ROWS = geo_rows(D);
    static constexpr int NACC = geo_nacc(D);
    static constexpr int NS = Dim<D>::NS;
};

enum : int { TNZ_MATRIX = 0, TNZ_NONE = 1 };   // noise class: padded matrices in shared memory | no per-point noise

struct TeamParams {
    const float* X;        // [N,D]
    const float* S;        // [N,D,D] | [N,D] | unused
    const float* g;        // [N] or null (statistics pass)
    const float* sigT;     // [Kp][NQ][8] float4: Sigma_j (+ shared noise) in the register order of the lanes
    const float* muT;      // [Kp][R][8]
    const float* cj;       // [Kp] log alpha_j - D/2 log 2pi (-inf for padding components)
    float* ll;             // [N]: written by the log-likelihood pass, read by the statistics pass
    float* out_ll;         // [N] or null: user-visible copy (log-likelihood pass)
    float* part;           // statistics pass: [grid][TB][NS]
    double* part_ll;       // [grid][2]: log-likelihood pass: sum ll, #points ; statistics pass: sum g, unused
    int* nonpd;
    long long N, n_tiles;
    int K, TB, n_chunks, n_slices, PT, NST, noise, use_bulk;
    float resp_floor;
    int n_s, off_X, off_g, off_ll, stage_bytes;   // ring-stage layout (bytes)
    int tmem_cols;
};

// Bytes of one ring stage: S | X | g | ll  (each 16-byte aligned)
struct TeamStage {
    int off_X, off_g, off_ll, bytes;
};
__host__ __device__ inline TeamStage team_stage(int D, int noise, int PT, bool has_g, bool has_ll) {
    TeamStage L;
    L.off_X = align16(PT * noise_floats_per_point(D, noise) * 4);
    L.off_g = L.off_X + align16(PT * D * 4);
    L.off_ll = L.off_g + (has_g ? align16(PT * 4) : 0);
    L.bytes = L.off_ll + (has_ll ? align16(PT * 4) : 0);
    L.bytes = (L.bytes + 127) & ~127;
    return L;
}
constexpr int kTeamHeader = 128;   // mbarriers (4 x 8 B), TMEM slot
__host__ __device__ inline size_t team_smem_bytes(int D, int noise, int PT, int NST, bool has_g, bool stats, int TPB) {
    const TeamStage L = team_stage(D, noise, PT, has_g, stats);
    size_t b = kTeamHeader + (size_t)NST * L.bytes;
    if (noise == NOISE_FULL || noise == NOISE_DIAG || noise == NOISE_PACKED) b += (size_t)PT * geo_rows(D) * geo_rs(D) * 4;  // re-packed noise
    b += (size_t)(TPB / 32) * 2 * kTW * kVS * 4;                                                     // published vectors
    if (!stats) b += (size_t)PT * kMaxTB * 2 * 4;                                                    // (max, sum) per point and team
    return b;
}

}  // namespace team

// per-D lookup of the team kernels (dgmm_team_dXX.cu)
struct TeamKernels {
    const void* loglik[2];   // [TNZ_MATRIX | TNZ_NONE]
    const void* stats[2];
};

}  // namespace dgmm
