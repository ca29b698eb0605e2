// Host-side check of the thread-private small-D linear algebra in csrc/dgmm_core.cuh
// (test tool only: the product has no CPU path).  Reads nothing; for D in a list builds
// random SPD matrices T and right-hand sides d in double, runs the float routines and
// prints, per D, the max relative errors against a double-precision Gauss-Jordan inverse:
//   D  err(logdet)  err(maha)  err(u)  err(c*T^{-1})  [classic route: err(u) err(T^{-1})]
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <cmath>
#include <random>

#include "../../density_deconvolution_b200/csrc/dgmm_core.cuh"

using namespace dgmm;

static void inverse_double(int D, const std::vector<double>& T, std::vector<double>& Z, double& logdet) {
    std::vector<double> M(T);
    Z.assign(D * D, 0.0);
    for (int i = 0; i < D; ++i) Z[i * D + i] = 1.0;
    logdet = 0.0;
    for (int k = 0; k < D; ++k) {
        const double p = M[k * D + k];  // SPD: no pivoting needed
        logdet += std::log(p);
        for (int c = 0; c < D; ++c) { M[k * D + c] /= p; Z[k * D + c] /= p; }
        for (int r = 0; r < D; ++r) {
            if (r == k) continue;
            const double f = M[r * D + k];
            for (int c = 0; c < D; ++c) { M[r * D + c] -= f * M[k * D + c]; Z[r * D + c] -= f * Z[k * D + c]; }
        }
    }
}

template <int D, bool ROLLED>
static int check(std::mt19937_64& rng) {
    constexpr int ND = D * (D + 1) / 2;
    std::normal_distribution<double> nd(0.0, 1.0);
    double e_ld = 0, e_m = 0, e_u = 0, e_z = 0, e_u2 = 0, e_z2 = 0;
    for (int rep = 0; rep < 20; ++rep) {
        std::vector<double> Q(D * D), T(D * D), d(D), Z;
        for (auto& q : Q) q = nd(rng);
        for (int a = 0; a < D; ++a)
            for (int b = 0; b < D; ++b) {
                double s = (a == b) ? 0.5 : 0.0;
                for (int c = 0; c < D; ++c) s += Q[c * D + a] * Q[c * D + b];
                T[a * D + b] = s;
            }
        for (auto& v : d) v = 3.0 * nd(rng);
        double logdet;
        inverse_double(D, T, Z, logdet);
        std::vector<double> u(D, 0.0);
        double maha = 0.0, zmax = 0.0, umax = 0.0;
        for (int a = 0; a < D; ++a) {
            for (int b = 0; b < D; ++b) { u[a] += Z[a * D + b] * d[b]; zmax = std::fmax(zmax, std::fabs(Z[a * D + b])); }
        }
        for (int a = 0; a < D; ++a) { maha += d[a] * u[a]; umax = std::fmax(umax, std::fabs(u[a])); }
        const float c = 1.0f;

        float A[ND], y[D], uf[D], lg2;
        for (int a = 0; a < D; ++a) {
            y[a] = (float)d[a];
            for (int b = 0; b <= a; ++b) A[tri(a, b)] = (float)T[a * D + b];
        }
        float mf;
        if (ROLLED) {
            rolled::ldl_forward<D>(A, y, lg2);
            mf = rolled::ldl_scale_rhs<D>(A, y);
            rolled::ldl_back_solve<D>(A, y, uf);
            rolled::ldl_inverse<D>(A);
        } else {
            unrolled::ldl_forward<D>(A, y, lg2);
            mf = unrolled::ldl_scale_rhs<D>(A, y);
            unrolled::ldl_back_solve<D>(A, y, uf);
            unrolled::ldl_inverse<D>(A);
        }
        e_ld = std::fmax(e_ld, std::fabs(lg2 * 0.6931471805599453 - logdet) / std::fmax(1.0, std::fabs(logdet)));
        e_m = std::fmax(e_m, std::fabs(mf - maha) / maha);
        for (int a = 0; a < D; ++a) {
            e_u = std::fmax(e_u, std::fabs(uf[a] - u[a]) / umax);
            for (int b = 0; b <= a; ++b) e_z = std::fmax(e_z, std::fabs(A[tri(a, b)] - c * Z[a * D + b]) / (c * zmax));
        }
        // classic route (materialising kernel)
        for (int a = 0; a < D; ++a) {
            y[a] = (float)d[a];
            for (int b = 0; b <= a; ++b) A[tri(a, b)] = (float)T[a * D + b];
        }
        unrolled::chol_forward<D>(A, y, lg2);
        unrolled::tri_inverse<D>(A);
        unrolled::back_apply<D>(A, y, uf);
        unrolled::gram_lower<D>(A);
        for (int a = 0; a < D; ++a) {
            e_u2 = std::fmax(e_u2, std::fabs(uf[a] - u[a]) / umax);
            for (int b = 0; b <= a; ++b) e_z2 = std::fmax(e_z2, std::fabs(A[tri(a, b)] - Z[a * D + b]) / zmax);
        }
    }
    printf("%d %d %.3e %.3e %.3e %.3e %.3e %.3e\n", D, (int)ROLLED, e_ld, e_m, e_u, e_z, e_u2, e_z2);
    const double tol = 2e-4;
    return (e_ld < tol && e_m < tol && e_u < tol && e_z < tol && e_u2 < tol && e_z2 < tol) ? 0 : 1;
}

int main() {
    std::mt19937_64 rng(12345);
    int bad = 0;
    bad += check<1, false>(rng);
    bad += check<2, false>(rng);
    bad += check<3, false>(rng);
    bad += check<5, false>(rng);
    bad += check<7, false>(rng);
    bad += check<10, false>(rng);
    bad += check<11, false>(rng);
    bad += check<16, false>(rng);
    bad += check<10, true>(rng);
    bad += check<17, true>(rng);
    bad += check<32, true>(rng);
    printf(bad ? "FAIL\n" : "OK\n");
    return bad;
}
