// CPU check of the ego-velocity solvers (csrc/vel_solve.cuh): full-rank Cholesky path against a residual check, and the
// minimum-norm path (rank 1 and rank 2 systems) against the two properties that define LAPACK gelsd's answer:
// the normal equations hold and the solution lies in the row space.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include "vel_solve.cuh"

static double urand() { return rand() / (double)RAND_MAX * 2.0 - 1.0; }

template <int D>
static int run(int n_rows, int trial) {
    double h[8][3], y[8], g[6] = {0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0}, c[3];
    for (int i = 0; i < n_rows; ++i) {
        double x[3] = {urand(), urand(), D == 3 ? urand() : 0.0};
        const double nrm = std::sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
        for (int k = 0; k < 3; ++k) h[i][k] = x[k] / nrm;
        y[i] = urand();
        g[0] += h[i][0] * h[i][0]; g[1] += h[i][0] * h[i][1]; g[2] += h[i][1] * h[i][1];
        g[3] += h[i][0] * h[i][2]; g[4] += h[i][1] * h[i][2]; g[5] += h[i][2] * h[i][2];
        for (int k = 0; k < 3; ++k) b[k] += h[i][k] * y[i];
    }
    rb::solve_normal<D>(g, b, c);
    // normal equations G c = b
    const double r0 = g[0] * c[0] + g[1] * c[1] + g[3] * c[2] - b[0];
    const double r1 = g[1] * c[0] + g[2] * c[1] + g[4] * c[2] - b[1];
    const double r2 = g[3] * c[0] + g[4] * c[1] + g[5] * c[2] - b[2];
    double err = std::fabs(r0) + std::fabs(r1) + std::fabs(r2);
    // minimum norm: c must be a combination of the rows (orthogonal to their null space)
    if (n_rows < D) {
        double nul[3];
        if (n_rows == 1) {                       // any vector orthogonal to h0: check the component of c off h0 is 0
            const double dot = c[0] * h[0][0] + c[1] * h[0][1] + c[2] * h[0][2];
            for (int k = 0; k < 3; ++k) err += std::fabs(c[k] - dot * h[0][k]);
        } else {                                 // D == 3, two rows: null space = h0 x h1
            nul[0] = h[0][1] * h[1][2] - h[0][2] * h[1][1];
            nul[1] = h[0][2] * h[1][0] - h[0][0] * h[1][2];
            nul[2] = h[0][0] * h[1][1] - h[0][1] * h[1][0];
            err += std::fabs(c[0] * nul[0] + c[1] * nul[1] + c[2] * nul[2]);
        }
    }
    if (!(err < 1e-11)) { std::printf("FAIL D=%d rows=%d trial=%d err=%.3e c=(%g,%g,%g)\n", D, n_rows, trial, err, c[0], c[1], c[2]); return 1; }
    return 0;
}

int main() {
    srand(5);
    int bad = 0;
    for (int t = 0; t < 200; ++t) {
        bad += run<2>(1, t) + run<2>(2, t) + run<2>(8, t);
        bad += run<3>(1, t) + run<3>(2, t) + run<3>(3, t) + run<3>(8, t);
    }
    std::printf(bad ? "host_vel_solve_check: %d failures\n" : "host_vel_solve_check: ok\n", bad);
    return bad ? 1 : 0;
}
