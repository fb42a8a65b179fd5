// SAR self-calibration of the synthetic-array geometry (reference
// processors/simple_synthetic_array_beamformer_processor_multiFrame.py:673-816): three strong scatterers are picked from
// the mean range profile and the beamformed image, the element-to-element phase steps they show after steering are
// regressed on their directions, and the accumulated corrections move the element positions.  O(S*S*N) arithmetic on
// an aperture of a few hundred KB: latency-bound, done in fp64 so every wrap / peak decision matches the fp64 reference.
//   phase 1  sar_cal_profile_kernel : one CTA per range bin k: mean over elements of 20 log10 |FFT_S(w[n] X[:, n])[k]|
//   phase 2  sar_cal_targets_kernel : one CTA: scipy find_peaks(height=0) + top-3 ranges, strongest azimuth peak of each
//   phase 3  sar_cal_correct_kernel : one CTA: steered phases, wrapped steps, per-step least squares, running sum
#include "radar_internal.h"
#include "vel_solve.cuh"

namespace rb {
namespace {

constexpr int kCalThreads = 256;
constexpr int kCalTargets = 3;

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    __syncthreads();
    return t;
}

// X[k, n] of the windowed aperture: sum_s w[n] x[s, n] e^{-2 pi i k s / S}; tw[j] = e^{-2 pi i j / S} in shared memory.
__device__ __forceinline__ double2 dft_bin(const cf* __restrict__ x, int S, int N, int n, int k, double w, const double2* tw) {
    double re = 0.0, im = 0.0;
    int j = 0;
    for (int s = 0; s < S; ++s) {
        const cf v = x[(size_t)s * N + n];
        const double2 t = tw[j];
        re += (double)v.x * t.x - (double)v.y * t.y;
        im += (double)v.x * t.y + (double)v.y * t.x;
        j += k; if (j >= S) j -= S;
    }
    return make_double2(w * re, w * im);
}

__device__ __forceinline__ void load_twiddles(double2* tw, int S) {
    for (int j = threadIdx.x; j < S; j += blockDim.x) {
        double sn, cs;
        sincospi(-2.0 * (double)j / (double)S, &sn, &cs);
        tw[j] = make_double2(cs, sn);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kCalThreads) sar_cal_profile_kernel(const cf* __restrict__ x, int S, int N,
                                                                      const double* __restrict__ win, double* __restrict__ avg_db) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2* tw = reinterpret_cast<double2*>(smem_raw);
    __shared__ double red[kCalThreads / 32];
    load_twiddles(tw, S);
    const int k = blockIdx.x;
    double acc = 0.0;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const double2 f = dft_bin(x, S, N, n, k, win[n], tw);
        acc += 20.0 * log10(hypot(f.x, f.y));
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) avg_db[k] = acc / (double)N;
}

// scipy.signal._peak_finding_utils._local_maxima_1d (plateau midpoints, never the end samples) restricted to values >= 0
// (find_peaks(height=0)); serial - the rows are at most a few hundred samples.  Calls visit(index) for every peak.
template <typename F>
__device__ void for_each_peak(const double* v, int n, F visit) {
    int i = 1;
    const int i_max = n - 1;
    while (i < i_max) {
        if (v[i - 1] < v[i]) {
            int ahead = i + 1;
            while (ahead < i_max && v[ahead] == v[i]) ++ahead;
            if (v[ahead] < v[i]) {
                const int mid = (i + ahead - 1) / 2;
                if (v[mid] >= 0.0) visit(mid);
                i = ahead;
            }
        }
        ++i;
    }
}

// targets_out: [rng0, az0, rng1, az1, rng2, az2, status]; status = number of targets found (3 = usable), -1 when the range
// profile has fewer than three peaks (the reference indexes past its peak list there).
__global__ void __launch_bounds__(kCalThreads) sar_cal_targets_kernel(const double* __restrict__ avg_db, int S,
                                                                      const cf* __restrict__ image, int G, int A, int E,
                                                                      int32_t* __restrict__ targets_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* row = reinterpret_cast<double*>(smem_raw);            // [A]
    __shared__ int rng[kCalTargets];
    __shared__ int n_rng;
    if (threadIdx.x == 0) {
        int best[kCalTargets] = {-1, -1, -1};
        int count = 0;
        for_each_peak(avg_db, S, [&](int p) {
            ++count;
            int q = p;
            for (int t = 0; t < kCalTargets; ++t) {               // insertion into the descending top three;
                if (q < 0) break;                                 // on equal values the later peak ranks first (argsort()[::-1])
                if (best[t] < 0 || avg_db[q] >= avg_db[best[t]]) { const int o = best[t]; best[t] = q; q = o; }
            }
        });
        n_rng = count;
        for (int t = 0; t < kCalTargets; ++t) rng[t] = best[t];
    }
    __syncthreads();
    if (n_rng < kCalTargets) {
        if (threadIdx.x == 0) {
            for (int t = 0; t < 2 * kCalTargets; ++t) targets_out[t] = -1;
            targets_out[2 * kCalTargets] = -1;
        }
        return;
    }
    int found = 0;
    for (int t = 0; t < kCalTargets; ++t) {
        for (int a = threadIdx.x; a < A; a += blockDim.x) {
            const cf v = image[(size_t)rng[t] * G + (size_t)a * E];
            row[a] = fabs(10.0 * log10(hypot((double)v.x, (double)v.y)));
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int best = -1;
            for_each_peak(row, A, [&](int p) { if (best < 0 || row[p] > row[best]) best = p; });   // first maximum (np.argmax)
            if (best >= 0) {
                targets_out[2 * found] = rng[t];
                targets_out[2 * found + 1] = best;
                ++found;
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        for (int t = found; t < kCalTargets; ++t) targets_out[2 * t] = targets_out[2 * t + 1] = -1;
        targets_out[2 * kCalTargets] = found;
    }
}

struct CalArgs {
    const cf* x; int S, N;
    const double* win;
    const double* pos;          // [3][N]
    const double* dirs;         // [3][G]
    int G, E;
    double inv_lambda;
    const int32_t* targets;
    double* phase;              // scratch [3][N]
    double* step;               // scratch [2][N-1]
    double* pos_out;            // [3][N]
};

__global__ void __launch_bounds__(kCalThreads) sar_cal_correct_kernel(const __grid_constant__ CalArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2* tw = reinterpret_cast<double2*>(smem_raw);
    const int N = a.N;
    for (int i = threadIdx.x; i < 3 * N; i += blockDim.x) a.pos_out[i] = a.pos[i];
    if (a.targets[2 * kCalTargets] != kCalTargets) return;        // uniform: geometry returned unchanged
    load_twiddles(tw, a.S);
    double d[kCalTargets][3];
    for (int t = 0; t < kCalTargets; ++t) {
        const size_t g = (size_t)a.targets[2 * t + 1] * a.E;
        for (int c = 0; c < 3; ++c) d[t][c] = a.dirs[(size_t)c * a.G + g];
    }
    // steered phase of every element at each target's range bin: angle(X[rng, n] e^{+2 pi i d.p_n / lambda})
    for (int i = threadIdx.x; i < kCalTargets * N; i += blockDim.x) {
        const int t = i / N, n = i - t * N;
        const double2 f = dft_bin(a.x, a.S, N, n, a.targets[2 * t], a.win[n], tw);
        const double turns = (d[t][0] * a.pos[n] + d[t][1] * a.pos[N + n] + d[t][2] * a.pos[2 * N + n]) * a.inv_lambda;
        double sn, cs;
        sincospi(2.0 * (turns - rint(turns)), &sn, &cs);
        a.phase[i] = atan2(f.x * sn + f.y * cs, f.x * cs - f.y * sn);
    }
    __syncthreads();
    // diff(np.unwrap(phase)): the step between neighbours, wrapped into [-pi, pi] when it is pi or more in size;
    // per step the 3x2 least squares Phi = (2 pi / lambda) D[:, 0:2] delta (minimum norm when the directions coincide)
    const double k2pi = 6.283185307179586476925286766559, kpi = 3.141592653589793238462643383279;
    double gm[6] = {0, 0, 0, 0, 0, 0};
    for (int t = 0; t < kCalTargets; ++t) {
        const double u = k2pi * a.inv_lambda * d[t][0], v = k2pi * a.inv_lambda * d[t][1];
        gm[0] += u * u; gm[1] += u * v; gm[2] += v * v;
    }
    for (int i = threadIdx.x; i < N - 1; i += blockDim.x) {
        double b[3] = {0, 0, 0}, c[3];
        for (int t = 0; t < kCalTargets; ++t) {
            const double dd = a.phase[t * N + i + 1] - a.phase[t * N + i];
            double step = dd;
            if (fabs(dd) >= kpi) {
                double m = fmod(dd + kpi, k2pi);
                if (m < 0.0) m += k2pi;                           // numpy's mod takes the divisor's sign
                m -= kpi;
                if (m == -kpi && dd > 0.0) m = kpi;
                step = m;
            }
            b[0] += k2pi * a.inv_lambda * d[t][0] * step;
            b[1] += k2pi * a.inv_lambda * d[t][1] * step;
        }
        solve_pinv(gm, b, c);
        a.step[i] = c[0];
        a.step[(N - 1) + i] = c[1];
    }
    __syncthreads();
    if (threadIdx.x < 2) {                                        // np.cumsum order: one serial running sum per axis
        const int ax = threadIdx.x;
        double run = 0.0;
        for (int i = 0; i < N - 1; ++i) {
            run += a.step[ax * (N - 1) + i];
            a.pos_out[ax * N + i + 1] = a.pos[ax * N + i + 1] - run;
        }
    }
}

}  // namespace

int launch_sar_calibrate(radar_ctx* ctx, const cf* aperture, int S, int N, const double* window, const double* positions,
                         const cf* image, int n_dirs, int n_elev, const double* dirs, double lambda_m, double* positions_out,
                         int32_t* targets_out, cudaStream_t st) {
    if (S < 3 || N < 2) return set_error(ctx, RADAR_E_INVALID, "calibration needs S >= 3 and N >= 2 (S %d, N %d)", S, N);
    if (n_elev <= 0 || n_dirs <= 0 || n_dirs % n_elev != 0)
        return set_error(ctx, RADAR_E_INVALID, "n_dirs %d is not a multiple of n_elev %d", n_dirs, n_elev);
    if (!(lambda_m > 0.0)) return set_error(ctx, RADAR_E_INVALID, "lambda must be positive");
    const int A = n_dirs / n_elev;
    const size_t tw_bytes = (size_t)S * sizeof(double2), row_bytes = (size_t)A * sizeof(double);
    if (tw_bytes > 200 * 1024 || row_bytes > 200 * 1024)
        return set_error(ctx, RADAR_E_INVALID, "calibration tables do not fit shared memory (S %d, azimuth bins %d)", S, A);
    const size_t n_doubles = (size_t)S + 3 * (size_t)N + 2 * (size_t)(N - 1);
    if (int rc = ensure_scratch(ctx, n_doubles * sizeof(double))) return rc;
    double* avg = reinterpret_cast<double*>(ctx->scratch);
    double* phase = avg + S;
    double* step = phase + 3 * (size_t)N;
    if (tw_bytes > 48 * 1024) {
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(sar_cal_profile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tw_bytes));
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(sar_cal_correct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tw_bytes));
    }
    if (row_bytes > 48 * 1024)
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(sar_cal_targets_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)row_bytes));
    sar_cal_profile_kernel<<<(unsigned)S, kCalThreads, tw_bytes, st>>>(aperture, S, N, window, avg);
    RB_CUDA_OK(ctx, cudaGetLastError());
    sar_cal_targets_kernel<<<1, kCalThreads, row_bytes, st>>>(avg, S, image, n_dirs, A, n_elev, targets_out);
    RB_CUDA_OK(ctx, cudaGetLastError());
    CalArgs a{aperture, S, N, window, positions, dirs, n_dirs, n_elev, 1.0 / lambda_m, targets_out, phase, step, positions_out};
    sar_cal_correct_kernel<<<1, kCalThreads, tw_bytes, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb

// ------------------------------------------------------------------------------------------------------------------
// Polar -> cartesian resampling of the SAR image at elevation index 0 (reference ...multiFrame.py:587-613:
// scipy.interpolate.griddata(method="linear", fill_value=0) = Delaunay triangulation + barycentric weights).  On a polar
// grid every cell is a cyclic trapezoid, so both diagonals are Delaunay and Qhull's pick is arbitrary per cell; the
// kernel splits every cell along `diagonal` (0: (i,j)-(i+1,j+1), 1: (i,j+1)-(i+1,j)).  One pixel per thread: sector by
// binary search on the angle, ring by binary search on the bisector coordinate (rings are bounded by CHORDS in the
// plane, not arcs), barycentric weights in fp64, values complex64.  Writes 2 floats per pixel: HBM-trivial.
// ------------------------------------------------------------------------------------------------------------------
namespace rb {
namespace {

struct P2cArgs {
    const cf* image; int S, G, E, A;
    const double* r; const double* az; const int32_t* az_order;
    double x0, dx, y0, dy; int nx, ny, diagonal;
    cf* out; uint8_t* inside;
};

__device__ __forceinline__ int upper_cell(const double* v, int n, double q) {      // searchsorted(v, q, "right") - 1, clipped
    int lo = 0, hi = n;
    while (lo < hi) { const int m = (lo + hi) >> 1; if (v[m] <= q) lo = m + 1; else hi = m; }
    return min(max(lo - 1, 0), n - 2);
}

__global__ void __launch_bounds__(256) polar_to_cart_kernel(const __grid_constant__ P2cArgs a) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)a.nx * a.ny) return;
    const int ix = (int)(idx / a.ny), iy = (int)(idx - (long long)ix * a.ny);
    const double x = a.x0 + ix * a.dx, y = a.y0 + iy * a.dy;
    const double th = atan2(y, x);
    const int j = upper_cell(a.az, a.A, th);
    const double az0 = a.az[j], az1 = a.az[j + 1];
    double s0, c0, s1, c1, sm, cm;
    sincos(az0, &s0, &c0);
    sincos(az1, &s1, &c1);
    sincos(0.5 * (az0 + az1), &sm, &cm);
    const double u = (x * cm + y * sm) / cos(0.5 * (az1 - az0));
    const int i = upper_cell(a.r, a.S, u);
    cf val = cmake(0.f, 0.f);
    bool hit = false;
    if (th >= a.az[0] && th <= a.az[a.A - 1] && u >= a.r[0] && u <= a.r[a.S - 1]) {
        const double r0 = a.r[i], r1 = a.r[i + 1];
        const int g0 = (a.az_order ? a.az_order[j] : j) * a.E, g1 = (a.az_order ? a.az_order[j + 1] : j + 1) * a.E;
        // corners: 0 = (i,j), 1 = (i,j+1), 2 = (i+1,j), 3 = (i+1,j+1)
        const double px[4] = {r0 * c0, r0 * c1, r1 * c0, r1 * c1}, py[4] = {r0 * s0, r0 * s1, r1 * s0, r1 * s1};
        // at r == 0 every azimuth sample sits on the origin; Qhull keeps the first duplicate (image column 0)
        const size_t at[4] = {(size_t)i * a.G + (r0 == 0.0 ? 0 : g0), (size_t)i * a.G + (r0 == 0.0 ? 0 : g1),
                              (size_t)(i + 1) * a.G + g0, (size_t)(i + 1) * a.G + g1};
        const int tri[2][2][3] = {{{0, 3, 2}, {0, 1, 3}}, {{0, 1, 2}, {1, 3, 2}}};
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            const int ia = tri[a.diagonal][t][0], ib = tri[a.diagonal][t][1], ic = tri[a.diagonal][t][2];
            const double det = (py[ib] - py[ic]) * (px[ia] - px[ic]) + (px[ic] - px[ib]) * (py[ia] - py[ic]);
            if (hit || det == 0.0) continue;
            const double l0 = ((py[ib] - py[ic]) * (x - px[ic]) + (px[ic] - px[ib]) * (y - py[ic])) / det;
            const double l1 = ((py[ic] - py[ia]) * (x - px[ic]) + (px[ia] - px[ic]) * (y - py[ic])) / det;
            const double l2 = 1.0 - l0 - l1;
            if (l0 >= -1e-12 && l1 >= -1e-12 && l2 >= -1e-12) {
                const cf va = a.image[at[ia]], vb = a.image[at[ib]], vc = a.image[at[ic]];
                val = cmake((float)(l0 * va.x + l1 * vb.x + l2 * vc.x), (float)(l0 * va.y + l1 * vb.y + l2 * vc.y));
                hit = true;
            }
        }
    }
    a.out[idx] = val;
    if (a.inside) a.inside[idx] = hit ? 1 : 0;
}

}  // namespace

int launch_polar_to_cart(radar_ctx* ctx, const cf* image, int S, int n_dirs, int n_elev, const double* range_bins,
                         const double* az_bins, const int32_t* az_order, int n_az, double x0, double dx, int nx, double y0,
                         double dy, int ny, int diagonal, cf* out, uint8_t* inside_out, cudaStream_t st) {
    if (S < 2 || n_az < 2) return set_error(ctx, RADAR_E_INVALID, "resampling needs at least 2 range and 2 azimuth bins");
    if (n_elev <= 0 || n_dirs != n_az * n_elev)
        return set_error(ctx, RADAR_E_INVALID, "n_dirs %d != n_az %d * n_elev %d", n_dirs, n_az, n_elev);
    if (diagonal != 0 && diagonal != 1) return set_error(ctx, RADAR_E_INVALID, "diagonal must be 0 or 1");
    if (nx < 0 || ny < 0) return set_error(ctx, RADAR_E_INVALID, "negative grid size");
    const long long total = (long long)nx * ny;
    if (total == 0) return RADAR_OK;
    P2cArgs a{image, S, n_dirs, n_elev, n_az, range_bins, az_bins, az_order, x0, dx, y0, dy, nx, ny, diagonal, out, inside_out};
    polar_to_cart_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb
