// Point cloud -> ego velocity: least squares inside RANSAC, one warp per frame.
//
// Reference arithmetic reproduced (paths relative to mmwave_radar_processing/ in the reference):
//   point_cloud_processing/vel_estimator.py:76-170  y = -v, H = xyz[:D]/|xyz[:D]|,
//       RANSACRegressor(LinearRegression(fit_intercept=False), residual_threshold=0.15, random_state=42,
//                       max_trials=20, min_samples=10)
// and, behind it, scikit-learn's fit loop (linear_model/_ransac.py): one random subset per trial, least squares on the
// subset, inliers |y - H c| <= threshold, the trial with more inliers wins (ties: better R^2 on its inliers), max_trials
// shrinks with the inlier ratio (stop_probability 0.99), final least squares on the winning inlier set.
//
// random_state=42 makes sklearn build a FRESH np.random.RandomState(42) on every fit, so the subset of trial t depends
// only on (number of points, t).  The host layer tabulates those subsets with numpy's own RandomState (the generator the
// reference uses) and uploads the table once; the kernel is deterministic given the table.
// Everything runs in fp64: H is formed with the same IEEE operations numpy uses (mul, add, sqrt, div - no contraction),
// so the inputs of the fit are bit-identical; the 2x2 / 3x3 normal equations are solved by Cholesky (sklearn: LAPACK
// gelsd), which moves coefficients by ~1e-15 relative - far below what can flip a 0.15 m/s inlier decision.
#include "radar_internal.h"
#include "vel_solve.cuh"

namespace rb {

constexpr int kVelWarps = 4;

struct EgoArgs {
    const double* points;         // [rows][4] (x, y, z, v)
    const int32_t* offsets;       // [n_frames + 1]
    int n_frames, dims;
    const int32_t* subsets;       // [n_blocks][max_trials][min_samples]
    const int32_t* subset_index;  // [table_max_n + 1]: block of a cloud of n points, -1 = not tabulated
    int table_max_n, max_trials, min_samples;
    double thr, stop_prob;
    double* vel_out;              // [n_frames][3]
    double* r2_out;               // [n_frames]
    double* frac_out;             // [n_frames]
    int32_t* status_out;          // [n_frames] number of inliers of the final model, 0 = no estimate
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) { return __reduce_add_sync(0xffffffffu, v); }

struct Row { double h[3]; double y; };

// H row and target of one point, operation for operation what numpy does (np.linalg.norm: sqrt(sum(x*x)), then divide).
// RAW: the row already holds (h0, h1, h2, y) - the 1-D regressions of the ADC-domain estimator
// (processors/velocity_estimator.py:409-506) hand their own design matrix.
template <int D, bool RAW = false>
__device__ __forceinline__ Row load_row(const double* __restrict__ p) {
    Row r;
    if (RAW) {
        r.h[0] = p[0]; r.h[1] = D >= 2 ? p[1] : 0.0; r.h[2] = D == 3 ? p[2] : 0.0; r.y = p[3];
        return r;
    }
    double s = __dmul_rn(p[0], p[0]);
    s = __dadd_rn(s, __dmul_rn(p[1], p[1]));
    if (D == 3) s = __dadd_rn(s, __dmul_rn(p[2], p[2]));
    const double nrm = sqrt(s);
    r.h[0] = p[0] / nrm;
    r.h[1] = p[1] / nrm;
    r.h[2] = (D == 3) ? p[2] / nrm : 0.0;
    r.y = -p[3];
    return r;
}

template <int D>
__device__ __forceinline__ void accumulate(const Row& r, double (&g)[6], double (&b)[3]) {
    g[0] += r.h[0] * r.h[0]; g[1] += r.h[0] * r.h[1]; g[2] += r.h[1] * r.h[1];
    b[0] += r.h[0] * r.y;    b[1] += r.h[1] * r.y;
    if (D == 3) {
        g[3] += r.h[0] * r.h[2]; g[4] += r.h[1] * r.h[2]; g[5] += r.h[2] * r.h[2];
        b[2] += r.h[2] * r.y;
    }
}

template <int D>
__device__ __forceinline__ double residual(const Row& r, const double (&c)[3]) {
    double pred = r.h[0] * c[0] + r.h[1] * c[1];
    if (D == 3) pred += r.h[2] * c[2];
    return fabs(r.y - pred);
}

// Inlier count, R^2 of model c on its inliers (sklearn r2_score, force_finite): two passes like the reference
// (mean first, then the sums of squares).  All lanes return the same values.
template <int D, bool RAW>
__device__ __forceinline__ void score_model(const double* __restrict__ pts, int n, int lane, const double (&c)[3], double thr,
                                            int* n_in_out, double* r2_out) {
    int cnt = 0;
    double sy = 0.0;
    for (int i = lane; i < n; i += 32) {
        const Row r = load_row<D, RAW>(pts + 4 * (size_t)i);
        if (residual<D>(r, c) <= thr) { ++cnt; sy += r.y; }
    }
    cnt = warp_sum(cnt);
    sy = warp_sum(sy);
    *n_in_out = cnt;
    if (cnt < 2) { *r2_out = __longlong_as_double(0x7ff8000000000000LL); return; }
    const double mean = sy / (double)cnt;
    double ss_res = 0.0, ss_tot = 0.0;
    for (int i = lane; i < n; i += 32) {
        const Row r = load_row<D, RAW>(pts + 4 * (size_t)i);
        const double e = residual<D>(r, c);
        if (e <= thr) { ss_res += e * e; ss_tot += (r.y - mean) * (r.y - mean); }
    }
    ss_res = warp_sum(ss_res);
    ss_tot = warp_sum(ss_tot);
    *r2_out = (ss_tot == 0.0) ? (ss_res == 0.0 ? 1.0 : 0.0) : 1.0 - ss_res / ss_tot;
}

__device__ __forceinline__ double dynamic_max_trials(int n_inliers, int n_samples, int min_samples, double prob) {
    const double eps = 2.220446049250313e-16;
    const double ratio = (double)n_inliers / (double)n_samples;
    const double nom = fmax(eps, 1.0 - prob);
    const double denom = fmax(eps, 1.0 - pow(ratio, (double)min_samples));
    if (nom == 1.0) return 0.0;
    if (denom == 1.0) return __longlong_as_double(0x7ff0000000000000LL);
    return fabs(ceil(log(nom) / log(denom)));
}

template <int D, bool RAW>
__global__ void __launch_bounds__(kVelWarps * 32) ego_velocity_kernel(const __grid_constant__ EgoArgs a) {
    const int lane = threadIdx.x & 31;
    const int frame = blockIdx.x * kVelWarps + (threadIdx.x >> 5);
    if (frame >= a.n_frames) return;
    const int begin = a.offsets[frame], n = a.offsets[frame + 1] - begin;
    const double* pts = a.points + 4 * (size_t)begin;
    double vel[3] = {0.0, 0.0, 0.0}, r2 = 0.0, frac = 0.0;
    int block = -1;
    if (n >= a.min_samples && n <= a.table_max_n) block = a.subset_index[n];
    int status = (n >= a.min_samples && block < 0) ? -1 : 0;     // -1: no subset block tabulated for this point count
    bool ok = block >= 0;
    if (ok) {                                  // sklearn's input validation: any nan / inf -> ValueError -> zeros
        bool finite = true;
        for (int i = lane; i < n; i += 32) {
            const Row r = load_row<D, RAW>(pts + 4 * (size_t)i);
            finite = finite && isfinite(r.h[0]) && isfinite(r.h[1]) && isfinite(r.h[2]) && isfinite(r.y);
        }
        ok = __all_sync(0xffffffffu, finite);
    }
    if (ok) {
        const int32_t* table = a.subsets + (size_t)block * a.max_trials * a.min_samples;
        int n_best = 1;
        double score_best = -__longlong_as_double(0x7ff0000000000000LL);
        double best[3] = {0.0, 0.0, 0.0};
        bool have = false;
        double limit = (double)a.max_trials;
        for (int t = 0; (double)t < limit; ++t) {
            double g[6] = {0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0}, c[3];
            for (int j = 0; j < a.min_samples; ++j)           // every lane forms the same small system (broadcast loads)
                accumulate<D>(load_row<D, RAW>(pts + 4 * (size_t)table[t * a.min_samples + j]), g, b);
            solve_normal<D>(g, b, c);
            int n_in;
            double score;
            score_model<D, RAW>(pts, n, lane, c, a.thr, &n_in, &score);
            if (n_in < n_best) continue;
            if (n_in == n_best && score < score_best) continue;
            n_best = n_in; score_best = score; have = true;
            best[0] = c[0]; best[1] = c[1]; best[2] = c[2];
            limit = fmin(limit, dynamic_max_trials(n_best, n, a.min_samples, a.stop_prob));
        }
        if (have) {                            // final model: least squares over the inliers of the best trial
            double g[6] = {0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0}, c[3];
            for (int i = lane; i < n; i += 32) {
                const Row r = load_row<D, RAW>(pts + 4 * (size_t)i);
                if (residual<D>(r, best) <= a.thr) accumulate<D>(r, g, b);
            }
#pragma unroll
            for (int k = 0; k < 6; ++k) g[k] = warp_sum(g[k]);
#pragma unroll
            for (int k = 0; k < 3; ++k) b[k] = warp_sum(b[k]);
            solve_normal<D>(g, b, c);
            // R^2 of the final model on the SAME inlier set (mask of the best trial, not re-classified)
            int cnt = 0;
            double sy = 0.0;
            for (int i = lane; i < n; i += 32) {
                const Row r = load_row<D, RAW>(pts + 4 * (size_t)i);
                if (residual<D>(r, best) <= a.thr) { ++cnt; sy += r.y; }
            }
            cnt = warp_sum(cnt);
            sy = warp_sum(sy);
            const double mean = sy / (double)cnt;
            double ss_res = 0.0, ss_tot = 0.0;
            for (int i = lane; i < n; i += 32) {
                const Row r = load_row<D, RAW>(pts + 4 * (size_t)i);
                if (residual<D>(r, best) <= a.thr) {
                    const double e = residual<D>(r, c);
                    ss_res += e * e; ss_tot += (r.y - mean) * (r.y - mean);
                }
            }
            ss_res = warp_sum(ss_res);
            ss_tot = warp_sum(ss_tot);
            vel[0] = c[0]; vel[1] = c[1]; vel[2] = (D == 3) ? c[2] : 0.0;
            r2 = (cnt > 3) ? ((ss_tot == 0.0) ? (ss_res == 0.0 ? 1.0 : 0.0) : 1.0 - ss_res / ss_tot) : 0.0;
            frac = (double)cnt / (double)n;
            status = cnt;
        }
    }
    if (lane == 0) {
        a.vel_out[3 * (size_t)frame + 0] = vel[0];
        a.vel_out[3 * (size_t)frame + 1] = vel[1];
        a.vel_out[3 * (size_t)frame + 2] = vel[2];
        if (a.r2_out) a.r2_out[frame] = r2;
        if (a.frac_out) a.frac_out[frame] = frac;
        if (a.status_out) a.status_out[frame] = status;
    }
}

int launch_ego_velocity(radar_ctx* ctx, const double* points, const int32_t* offsets, int n_frames, int dims,
                        const int32_t* subsets, const int32_t* subset_index, int table_max_n, int max_trials,
                        int min_samples, double residual_threshold, double stop_probability, double* vel_out,
                        double* r2_out, double* frac_out, int32_t* status_out, cudaStream_t st) {
    if (dims != 2 && dims != 3 && dims != -1 && dims != -2 && dims != -3)
        return set_error(ctx, RADAR_E_INVALID, "dims must be 2 or 3 (point rows) or -1..-3 (raw design rows), got %d", dims);
    if (max_trials <= 0 || min_samples <= 0 || table_max_n < min_samples)
        return set_error(ctx, RADAR_E_INVALID, "bad RANSAC table (max_trials %d, min_samples %d, table_max_n %d)", max_trials, min_samples, table_max_n);
    if (n_frames <= 0) return RADAR_OK;
    EgoArgs a{points, offsets, n_frames, dims, subsets, subset_index, table_max_n, max_trials, min_samples, residual_threshold,
              stop_probability, vel_out, r2_out, frac_out, status_out};
    const unsigned grid = (unsigned)((n_frames + kVelWarps - 1) / kVelWarps);
    if (dims == 2) ego_velocity_kernel<2, false><<<grid, kVelWarps * 32, 0, st>>>(a);
    else if (dims == 3) ego_velocity_kernel<3, false><<<grid, kVelWarps * 32, 0, st>>>(a);
    else if (dims == -3) ego_velocity_kernel<3, true><<<grid, kVelWarps * 32, 0, st>>>(a);
    else ego_velocity_kernel<2, true><<<grid, kVelWarps * 32, 0, st>>>(a);      // one or two raw columns (a zero column is dropped by the minimum-norm solve)
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb
