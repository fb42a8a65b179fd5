// Altimeter support: zoom range profile and prominence peak picking.
//
// Reference arithmetic reproduced (paths relative to mmwave_radar_processing/ in the reference):
//   processors/range_resp.py:59-103   RangeProcessor.zoom_fft   (scipy.signal.ZoomFFT, m = n points, endpoint=False)
//   processors/range_resp.py:105-150  RangeProcessor.find_peaks (scipy.signal.find_peaks(prominence=6) + 20 dB gate + top-N)
//
// ZoomFFT is a chirp-Z transform; what it evaluates is X[k] = sum_s x[s] exp(-2 pi i s r_k / range_max),
// r_k = start + k (stop - start)/n.  With n <= 512 samples the direct sum is a few hundred thousand complex MACs per
// frame, so it is evaluated as such: the phase s*r_k/range_max is reduced modulo 1 in fp64 (it reaches hundreds of turns)
// before the fp32 sincospi, the sums run in fp32 per antenna.
#include "radar_internal.h"
#include "rd_core.cuh"

namespace rb {

// defined in kernels_rd.cu
struct AdcViewAlt {
    const void* base;
    int in_kind;
    int R, S, cpl, L, V, C;
};
__device__ __forceinline__ cf adc_fetch_alt(const AdcViewAlt& a, long long frame, int v, int s, int l) {
    if (a.in_kind == RADAR_IN_IQ16) {
        const int tx = v / a.R, rx = v - tx * a.R;
        const long long row = ((frame * a.C + (long long)l * a.cpl + tx) * a.R + rx) * a.S + s;
        const unsigned w = __ldg(reinterpret_cast<const unsigned*>(a.base) + row);
        return cmake((float)(short)(w & 0xffffu), (float)(short)(w >> 16));
    }
    const long long idx = ((frame * a.V + v) * a.S + s) * (long long)a.L + l;
    return __ldg(reinterpret_cast<const cf*>(a.base) + idx);
}

constexpr int kZoomThreads = 128;

struct ZoomArgs {
    AdcViewAlt adc;
    int loop_idx;
    const float* winS;
    const double* start_m; const double* stop_m;     // device [n_frames]
    double inv_range_max;
    float* out;                                      // [frame][S]
};

__global__ void __launch_bounds__(kZoomThreads) range_zoom_kernel(const __grid_constant__ ZoomArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cf* x = reinterpret_cast<cf*>(smem_raw);         // [V][S] windowed samples of the chirp
    const int S = a.adc.S, V = a.adc.V;
    const long long frame = blockIdx.x;
    for (int i = threadIdx.x; i < V * S; i += blockDim.x) {
        const int v = i / S, s = i - v * S;
        x[i] = cscale(adc_fetch_alt(a.adc, frame, v, s, a.loop_idx), a.winS[s]);
    }
    __syncthreads();
    const double start = a.start_m[frame], stop = a.stop_m[frame];
    for (int k = threadIdx.x; k < S; k += blockDim.x) {
        const double turns_per_sample = (start + (double)k * (stop - start) / (double)S) * a.inv_range_max;
        float sum_mag = 0.0f;
        for (int v0 = 0; v0 < V; v0 += 4) {          // four antennas share one twiddle evaluation
            cf acc[4] = {cmake(0.f, 0.f), cmake(0.f, 0.f), cmake(0.f, 0.f), cmake(0.f, 0.f)};
            const int nv = min(4, V - v0);
            for (int s = 0; s < S; ++s) {
                const double t = turns_per_sample * (double)s;
                const float frac = (float)(t - floor(t));
                float sn, cs;
                sincospif(-2.0f * frac, &sn, &cs);
                for (int j = 0; j < nv; ++j) {
                    const cf xv = x[(v0 + j) * S + s];
                    acc[j].x = fmaf(xv.x, cs, fmaf(-xv.y, sn, acc[j].x));
                    acc[j].y = fmaf(xv.x, sn, fmaf(xv.y, cs, acc[j].y));
                }
            }
            for (int j = 0; j < nv; ++j) sum_mag += sqrtf(acc[j].x * acc[j].x + acc[j].y * acc[j].y);
        }
        a.out[frame * S + k] = sum_mag / (float)V;
    }
}

int launch_range_zoom(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx, const double* start_m,
                      const double* stop_m, double range_max_m, float* out, cudaStream_t st) {
    if (in_kind != RADAR_IN_IQ16 && in_kind != RADAR_IN_CUBE64) return set_error(ctx, RADAR_E_INVALID, "unknown in_kind %d", in_kind);
    if (loop_idx < 0 || loop_idx >= ctx->shape.num_loops) return set_error(ctx, RADAR_E_INVALID, "loop_idx %d out of range", loop_idx);
    if (!(range_max_m > 0.0)) return set_error(ctx, RADAR_E_INVALID, "range_max_m must be positive");
    if (n_frames <= 0) return RADAR_OK;
    ZoomArgs a;
    a.adc.base = adc; a.adc.in_kind = in_kind;
    a.adc.R = ctx->shape.num_rx; a.adc.S = ctx->shape.num_samples; a.adc.cpl = ctx->shape.chirps_per_loop;
    a.adc.L = ctx->shape.num_loops; a.adc.V = ctx->V; a.adc.C = ctx->C;
    a.loop_idx = loop_idx;
    a.winS = ctx->fftS.hann;
    a.start_m = start_m; a.stop_m = stop_m;
    a.inv_range_max = 1.0 / range_max_m;
    a.out = out;
    const size_t smem = (size_t)ctx->V * ctx->shape.num_samples * sizeof(cf);
    if (smem > (size_t)ctx->max_smem_optin) return set_error(ctx, RADAR_E_UNSUPPORTED, "zoom profile: chirp exceeds shared memory");
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(range_zoom_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    range_zoom_kernel<<<(unsigned)n_frames, kZoomThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ------------------------------------------------------------------------------------------------
// scipy.signal.find_peaks(x, prominence=P): local maxima (plateau -> midpoint, never the end samples), prominence with
// wlen=None (walk left/right until a HIGHER sample, keep the minimum), then the reference's gate and ordering.
// One CTA per profile; n <= 4096.
// ------------------------------------------------------------------------------------------------
constexpr int kPeakThreads = 128;
constexpr int kPeakMaxN = 4096;

struct PeakArgs {
    const float* profiles; int n; int in_db;
    double prominence, threshold_db; int max_peaks;
    int32_t* idx_out; double* db_out; int32_t* count_out;
};

__global__ void __launch_bounds__(kPeakThreads) find_peaks_kernel(const __grid_constant__ PeakArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* x = reinterpret_cast<double*>(smem_raw);                 // [n] response in dB
    int* cand = reinterpret_cast<int*>(x + a.n);                     // [n/2 + 1] peaks that passed the prominence test
    __shared__ int n_cand;
    const long long prof = blockIdx.x;
    const float* src = a.profiles + prof * a.n;
    if (threadIdx.x == 0) n_cand = 0;
    for (int i = threadIdx.x; i < a.n; i += blockDim.x)
        x[i] = a.in_db ? (double)src[i] : 20.0 * log10((double)src[i]);
    __syncthreads();
    const int i_max = a.n - 1;
    for (int i = 1 + threadIdx.x; i < i_max; i += blockDim.x) {
        if (!(x[i - 1] < x[i])) continue;                            // i must be the left edge of a (possibly flat) top
        int ahead = i + 1;
        while (ahead < i_max && x[ahead] == x[i]) ++ahead;
        if (!(x[ahead] < x[i])) continue;
        const int p = (i + ahead - 1) / 2;
        const double xp = x[p];
        double left_min = xp, right_min = xp;
        for (int j = p; j >= 0 && x[j] <= xp; --j) left_min = fmin(left_min, x[j]);
        for (int j = p; j <= i_max && x[j] <= xp; ++j) right_min = fmin(right_min, x[j]);
        if (xp - fmax(left_min, right_min) >= a.prominence) cand[atomicAdd(&n_cand, 1)] = p;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    int32_t* idx = a.idx_out + prof * a.max_peaks;
    double* dbo = a.db_out ? a.db_out + prof * a.max_peaks : nullptr;
    double best = -INFINITY;
    for (int c = 0; c < n_cand; ++c) best = fmax(best, x[cand[c]]);
    const double gate = best - a.threshold_db;
    int n_out = 0;
    // np.argsort(vals)[::-1][:max_peaks]: strongest first; equal values -> the later peak first
    double prev_val = INFINITY; int prev_idx = 0x7fffffff;
    for (int k = 0; k < a.max_peaks; ++k) {
        int sel = -1; double sel_val = -INFINITY;
        for (int c = 0; c < n_cand; ++c) {
            const int p = cand[c]; const double v = x[p];
            if (!(v >= gate)) continue;
            const bool after_prev = v < prev_val || (v == prev_val && p < prev_idx);   // not yet emitted
            if (!after_prev) continue;
            if (v > sel_val || (v == sel_val && p > sel)) { sel = p; sel_val = v; }
        }
        idx[k] = sel;
        if (dbo) dbo[k] = sel >= 0 ? sel_val : 0.0;
        if (sel < 0) { for (int r = k + 1; r < a.max_peaks; ++r) { idx[r] = -1; if (dbo) dbo[r] = 0.0; } break; }
        ++n_out; prev_val = sel_val; prev_idx = sel;
    }
    if (a.count_out) a.count_out[prof] = n_out;
}

int launch_find_peaks(radar_ctx* ctx, const float* profiles, int n_profiles, int n, int in_db, double prominence,
                      double threshold_db, int max_peaks, int32_t* idx_out, double* db_out, int32_t* count_out,
                      cudaStream_t st) {
    if (n < 1 || n > kPeakMaxN) return set_error(ctx, RADAR_E_INVALID, "profile length %d outside 1..%d", n, kPeakMaxN);
    if (max_peaks < 1) return set_error(ctx, RADAR_E_INVALID, "max_peaks must be positive");
    if (!idx_out) return set_error(ctx, RADAR_E_INVALID, "peak_idx_out is null");
    if (n_profiles <= 0) return RADAR_OK;
    PeakArgs a;
    a.profiles = profiles; a.n = n; a.in_db = in_db;
    a.prominence = prominence; a.threshold_db = threshold_db; a.max_peaks = max_peaks;
    a.idx_out = idx_out; a.db_out = db_out; a.count_out = count_out;
    const size_t smem = (size_t)n * sizeof(double) + ((size_t)n / 2 + 2) * sizeof(int);
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(find_peaks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    find_peaks_kernel<<<(unsigned)n_profiles, kPeakThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb
