// Small dense solvers of the ego-velocity fit (kernels_vel.cu); __host__ __device__ so tests/native/host_vel_solve_check.cu
// can run them on the CPU.
#pragma once
#include <cmath>
#ifdef __CUDACC__
#define RB_HD __host__ __device__
#else
#define RB_HD
#endif

namespace rb {

// Minimum-norm least squares for a rank-deficient system (fewer inliers than unknowns, or all directions of the set
// parallel): c = sum over the eigenpairs of G with lambda > 1e-12 lambda_max of (v.b / lambda) v - what LAPACK's gelsd
// returns when the discarded singular values are exactly zero.  Cyclic Jacobi on the symmetric 3x3 (D == 2: padded).
RB_HD inline void solve_pinv(const double (&g)[6], const double (&b)[3], double (&c)[3]) {
    double a[3][3] = {{g[0], g[1], g[3]}, {g[1], g[2], g[4]}, {g[3], g[4], g[5]}};
    double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 12; ++sweep) {
#pragma unroll
        for (int pair = 0; pair < 3; ++pair) {
            const int p = pair == 2 ? 1 : 0, q = pair == 0 ? 1 : 2;        // (0,1), (0,2), (1,2)
            const double apq = a[p][q];
            if (fabs(apq) <= 1e-300) continue;
            const double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
            const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
#pragma unroll
            for (int k = 0; k < 3; ++k) {                                   // A <- A J, V <- V J
                const double akp = a[k][p], akq = a[k][q];
                a[k][p] = cs * akp - sn * akq;
                a[k][q] = sn * akp + cs * akq;
                const double vkp = v[k][p], vkq = v[k][q];
                v[k][p] = cs * vkp - sn * vkq;
                v[k][q] = sn * vkp + cs * vkq;
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {                                   // A <- J^T A
                const double apk = a[p][k], aqk = a[q][k];
                a[p][k] = cs * apk - sn * aqk;
                a[q][k] = sn * apk + cs * aqk;
            }
        }
    }
    const double lmax = fmax(a[0][0], fmax(a[1][1], a[2][2]));
    c[0] = c[1] = c[2] = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double lam = a[i][i];
        if (lam > 1e-12 * lmax && lam > 0.0) {
            const double w = (v[0][i] * b[0] + v[1][i] * b[1] + v[2][i] * b[2]) / lam;
            c[0] += w * v[0][i]; c[1] += w * v[1][i]; c[2] += w * v[2][i];
        }
    }
}

// Normal equations G c = b (D x D, symmetric) by Cholesky; rank-deficient G goes to solve_pinv.
template <int D>
RB_HD inline void solve_normal(const double (&g)[6], const double (&b)[3], double (&c)[3]) {
    // g: [00, 01, 11, 02, 12, 22]
    c[0] = c[1] = c[2] = 0.0;
    const double tr = g[0] + g[2] + (D == 3 ? g[5] : 0.0);
    const double tiny = 1e-10 * tr;
    const double l00 = sqrt(g[0]);
    bool ok = g[0] > tiny;
    const double l10 = g[1] / l00;
    const double d1 = g[2] - l10 * l10;
    ok = ok && d1 > tiny;
    const double l11 = sqrt(d1);
    if (D == 2) {
        if (ok) {
            const double z0 = b[0] / l00;
            const double z1 = (b[1] - l10 * z0) / l11;
            c[1] = z1 / l11;
            c[0] = (z0 - l10 * c[1]) / l00;
            return;
        }
    } else {
        const double l20 = g[3] / l00;
        const double l21 = (g[4] - l20 * l10) / l11;
        const double d2 = g[5] - l20 * l20 - l21 * l21;
        ok = ok && d2 > tiny;
        if (ok) {
            const double l22 = sqrt(d2);
            const double z0 = b[0] / l00;
            const double z1 = (b[1] - l10 * z0) / l11;
            const double z2 = (b[2] - l20 * z0 - l21 * z1) / l22;
            c[2] = z2 / l22;
            c[1] = (z1 - l21 * c[2]) / l11;
            c[0] = (z0 - l10 * c[1] - l20 * c[2]) / l00;
            return;
        }
    }
    if (tr > 0.0) solve_pinv(g, b, c);
}

}  // namespace rb
