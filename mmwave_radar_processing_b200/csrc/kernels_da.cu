// Doppler-azimuth family: angle DFT over every (range, Doppler) cell of a cube with the
// range average fused in, the zoom (chirp-Z) Doppler transform of the "precise" variant, and the row reduction of the
// micro-Doppler spectrogram.
//
// Reference arithmetic reproduced (paths relative to mmwave_radar_processing/ in the reference):
//   processors/doppler_azimuth_resp.py:84-100   Hann(S) x Hann(L) (x Hann(antennas) for standard arrays with TDM)
//   processors/doppler_azimuth_resp.py:102-128  range FFT + range-window mask
//   processors/doppler_azimuth_resp.py:296-334  coarse: zero-padded fft2 over (chirp, antenna), fftshift, abs
//   processors/doppler_azimuth_resp.py:130-162, 207-294  precise: scipy.signal.ZoomFFT over chirps, angle FFT, abs
//   processors/doppler_azimuth_resp.py:483-491  valid-angle columns, mean over range
//   processors/micro_doppler_resp.py:91-114     un-windowed |RD| of one antenna, max over the range rows kept
//
// The coarse response is  out[d][a] = mean_k | sum_v w[v] RD[v][k][d] e^{-2 pi i v a'/A} |  with RD the windowed,
// Doppler-shifted range-Doppler cube that radar_range_doppler already produces, so the kernel starts from that cube.
// The inner dimension (antennas, 4..12) is far too short for a tensor-core GEMM in a precision that keeps 1e-4
// (TF32 does not); the zero-padded angle DFT is Horner's rule on the FP32 pipe with the antenna samples in registers.
#include <mutex>
#include "radar_internal.h"
#include "rd_core.cuh"

namespace rb {

constexpr int kDaThreads = 256;
constexpr int kDaBins = 8;        // angle bins a thread carries per pass

struct DaArgs {
    const cf* cube;               // [frame][n_ant][rows][cols]
    int n_ant, rows, cols, k0, k1, A, shift;
    const float* weight;          // [n_ant] or nullptr
    float* out;                   // [frame][cols][A]
};

template <int NA>
__global__ void __launch_bounds__(kDaThreads) doppler_azimuth_kernel(const __grid_constant__ DaArgs a) {
    __shared__ float wgt[NA];
    for (int v = threadIdx.x; v < NA; v += blockDim.x) wgt[v] = (v < a.n_ant) ? (a.weight ? a.weight[v] : 1.0f) : 0.0f;
    __syncthreads();
    const long long frame = blockIdx.x;
    const int dper = min(a.cols, (int)blockDim.x);
    const int groups = blockDim.x / dper;
    const int g = threadIdx.x / dper, dl = threadIdx.x - g * dper;
    if (g >= groups) return;
    const size_t ant_stride = (size_t)a.rows * a.cols;
    const cf* base = a.cube + (size_t)frame * a.n_ant * ant_stride;
    const float inv_k = 1.0f / (float)(a.k1 - a.k0);      // empty range window: 0/0 = nan, like np.mean of nothing
    for (int d = dl; d < a.cols; d += dper) {
        for (int a0 = g * kDaBins; a0 < a.A; a0 += groups * kDaBins) {
            cf w[kDaBins];
            float acc[kDaBins];
#pragma unroll
            for (int i = 0; i < kDaBins; ++i) {
                const int bin = a0 + i;
                const int src = a.shift ? (bin + a.A - a.A / 2) % a.A : bin;   // fftshift: out[i] = in[(i - A//2) mod A]
                double s, c;
                sincospi(-2.0 * (double)src / (double)a.A, &s, &c);
                w[i] = cmake((float)c, (float)s);
                acc[i] = 0.0f;
            }
            for (int k = a.k0; k < a.k1; ++k) {
                cf z[NA];
                const cf* p = base + (size_t)k * a.cols + d;
#pragma unroll
                for (int v = 0; v < NA; ++v) {
                    z[v] = cmake(0.f, 0.f);
                    if (v < a.n_ant) z[v] = cscale(__ldg(p + v * ant_stride), wgt[v]);
                }
#pragma unroll
                for (int i = 0; i < kDaBins; ++i) {
                    cf h = cmake(0.f, 0.f);
#pragma unroll
                    for (int v = NA - 1; v >= 0; --v) {
                        if (v < a.n_ant) {          // uniform branch: Horner over the antennas present
                            const float hx = fmaf(h.x, w[i].x, fmaf(-h.y, w[i].y, z[v].x));
                            const float hy = fmaf(h.x, w[i].y, fmaf(h.y, w[i].x, z[v].y));
                            h.x = hx; h.y = hy;
                        }
                    }
                    acc[i] += sqrtf(fmaf(h.x, h.x, h.y * h.y));
                }
            }
#pragma unroll
            for (int i = 0; i < kDaBins; ++i)
                if (a0 + i < a.A) a.out[((size_t)frame * a.cols + d) * a.A + a0 + i] = acc[i] * inv_k;
        }
    }
}

int launch_doppler_azimuth(radar_ctx* ctx, const cf* cube, int n_frames, int n_ant, int rows, int cols, int k_begin,
                           int k_end, const float* weight, int num_angle_bins, int shift_angle, float* out,
                           cudaStream_t st) {
    if (n_ant <= 0 || n_ant > 16) return set_error(ctx, RADAR_E_UNSUPPORTED, "doppler-azimuth handles 1..16 antennas, got %d", n_ant);
    if (rows <= 0 || cols <= 0 || num_angle_bins <= 0) return set_error(ctx, RADAR_E_INVALID, "empty cube or angle axis");
    if (k_begin < 0 || k_end > rows || k_begin > k_end) return set_error(ctx, RADAR_E_INVALID, "range rows [%d, %d) outside [0, %d)", k_begin, k_end, rows);
    if (n_frames <= 0) return RADAR_OK;
    DaArgs a{cube, n_ant, rows, cols, k_begin, k_end, num_angle_bins, shift_angle ? 1 : 0, weight, out};
    if (n_ant <= 4) doppler_azimuth_kernel<4><<<(unsigned)n_frames, kDaThreads, 0, st>>>(a);
    else if (n_ant <= 8) doppler_azimuth_kernel<8><<<(unsigned)n_frames, kDaThreads, 0, st>>>(a);
    else if (n_ant <= 12) doppler_azimuth_kernel<12><<<(unsigned)n_frames, kDaThreads, 0, st>>>(a);
    else doppler_azimuth_kernel<16><<<(unsigned)n_frames, kDaThreads, 0, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ---- zoom Doppler transform -------------------------------------------------------------------------------------
// scipy.signal.ZoomFFT(n, [f0, f1], m = n, fs, endpoint=False) evaluates X[j] = sum_l x[l] exp(-2 pi i l nu_j) with
// nu_j = (f0 + j (f1 - f0)/m)/fs turns per chirp.  The host layer forms nu_j in fp64 with the reference's expressions;
// the twiddle table exp(-2 pi i frac(l nu_j)) is built once per call (phase reduced modulo 1 in fp64), then every
// (frame, antenna) tile is a small complex matrix product [rows x L] . [L x J].
__global__ void zoom_twiddle_kernel(const double* __restrict__ turns, int L, int J, cf* __restrict__ tw) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L * J) return;
    const int l = i / J, j = i - l * J;
    double ph = (double)l * turns[j];
    ph -= rint(ph);
    double s, c;
    sincospi(-2.0 * ph, &s, &c);
    tw[i] = cmake((float)c, (float)s);
}

struct CzArgs {
    const cf* cube;               // [frame][n_ant][S][L]  range transform done, chirp axis in loop order
    int S, L, k0, k1, J;
    const cf* tw;                 // [L][J]
    cf* out;                      // [frame][n_ant][k1-k0][J]
};

__global__ void __launch_bounds__(kDaThreads) chirp_zoom_kernel(const __grid_constant__ CzArgs a) {
    const size_t tile = blockIdx.x;                       // frame * n_ant + antenna
    const cf* src = a.cube + tile * (size_t)a.S * a.L;
    const int K = a.k1 - a.k0;
    cf* dst = a.out + tile * (size_t)K * a.J;
    for (int idx = threadIdx.x; idx < K * a.J; idx += blockDim.x) {
        const int kk = idx / a.J, j = idx - kk * a.J;
        const cf* row = src + (size_t)(a.k0 + kk) * a.L;
        cf acc = cmake(0.f, 0.f);
        for (int l = 0; l < a.L; ++l) {
            const cf y = __ldg(row + l), t = __ldg(a.tw + (size_t)l * a.J + j);
            acc.x = fmaf(y.x, t.x, fmaf(-y.y, t.y, acc.x));
            acc.y = fmaf(y.x, t.y, fmaf(y.y, t.x, acc.y));
        }
        dst[idx] = acc;
    }
}

int launch_chirp_zoom(radar_ctx* ctx, const cf* range_cube, int n_frames, int n_ant, int k_begin, int k_end,
                      const double* turns, int n_out, cf* out, cudaStream_t st) {
    const int S = ctx->shape.num_samples, L = ctx->shape.num_loops;
    if (n_ant <= 0 || n_out <= 0) return set_error(ctx, RADAR_E_INVALID, "empty antenna list or frequency list");
    if (k_begin < 0 || k_end > S || k_begin > k_end) return set_error(ctx, RADAR_E_INVALID, "range rows [%d, %d) outside [0, %d)", k_begin, k_end, S);
    if (n_frames <= 0 || k_begin == k_end) return RADAR_OK;
    if (int rc = ensure_scratch(ctx, (size_t)L * n_out * sizeof(cf))) return rc;
    cf* tw = reinterpret_cast<cf*>(ctx->scratch);
    const int n_tw = L * n_out;
    zoom_twiddle_kernel<<<(n_tw + 255) / 256, 256, 0, st>>>(turns, L, n_out, tw);
    RB_CUDA_OK(ctx, cudaGetLastError());
    CzArgs a{range_cube, S, L, k_begin, k_end, n_out, tw, out};
    const long long grid = (long long)n_frames * n_ant;
    if (grid > 0x7fffffffLL) return set_error(ctx, RADAR_E_INVALID, "too many frames in one call");
    chirp_zoom_kernel<<<(unsigned)grid, kDaThreads, 0, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ---- zoom Doppler transform without any allocation (radar_chirp_zoom_static) ------------------------------------------
// Same two kernels as launch_chirp_zoom, but the per-call twiddle table lives in a file-scope device array instead of the
// context's cudaMalloc'ed scratch, so the call allocates nothing (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  Every GPU that loads the library holds the array (kZoomStaticEntries * 8 bytes = 512 KiB) until exit.
// All calls on a device share it: the entries [0, L*n_out) are rewritten by zoom_twiddle_static_kernel at the start of every
// call (nothing has to start at zero, nothing of the previous call is read), and two calls that overlap on the GPU
// (different streams, no dependency) race on it.
constexpr int kZoomStaticEntries = 1 << 16;            // twiddles: num_loops * n_out <= 65536
constexpr int kZoomStaticDevices = 64;
static __device__ __align__(256) cf g_zoom_tw_static[kZoomStaticEntries];
// In host code `g_zoom_tw_static` is only a handle; the device address is looked up once per device, with that device
// current, and cached here.
static cf* g_zoom_tw_addr[kZoomStaticDevices] = {nullptr};
static std::mutex g_zoom_tw_mutex;

// device code may name the array directly: the table builder does, the transform kernel takes the looked-up address
__global__ void zoom_twiddle_static_kernel(const double* __restrict__ turns, int L, int J) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L * J) return;
    const int l = i / J, j = i - l * J;
    double ph = (double)l * turns[j];
    ph -= rint(ph);
    double s, c;
    sincospi(-2.0 * ph, &s, &c);
    g_zoom_tw_static[i] = cmake((float)c, (float)s);
}

static int zoom_static_table(radar_ctx* ctx, cf** out) {
    int dev = -1;
    RB_CUDA_OK(ctx, cudaGetDevice(&dev));
    if (dev != ctx->device) return set_error(ctx, RADAR_E_INVALID, "device %d is current, the context lives on %d", dev, ctx->device);
    if (dev < 0 || dev >= kZoomStaticDevices) return set_error(ctx, RADAR_E_UNSUPPORTED, "device index %d beyond %d", dev, kZoomStaticDevices - 1);
    std::lock_guard<std::mutex> lock(g_zoom_tw_mutex);
    if (!g_zoom_tw_addr[dev]) {
        cf* p = nullptr;
        RB_CUDA_OK(ctx, cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_zoom_tw_static));
        if (!p) return set_error(ctx, RADAR_E_CUDA, "cudaGetSymbolAddress returned a null address");
        g_zoom_tw_addr[dev] = p;
    }
    *out = g_zoom_tw_addr[dev];
    return RADAR_OK;
}

int chirp_zoom_static_capacity() { return kZoomStaticEntries; }

int launch_chirp_zoom_static(radar_ctx* ctx, const cf* range_cube, int n_frames, int n_ant, int k_begin, int k_end,
                             const double* turns, int n_out, cf* out, cudaStream_t st) {
    const int S = ctx->shape.num_samples, L = ctx->shape.num_loops;
    if (n_ant <= 0 || n_out <= 0) return set_error(ctx, RADAR_E_INVALID, "empty antenna list or frequency list");
    if (k_begin < 0 || k_end > S || k_begin > k_end) return set_error(ctx, RADAR_E_INVALID, "range rows [%d, %d) outside [0, %d)", k_begin, k_end, S);
    if ((long long)L * n_out > kZoomStaticEntries)
        return set_error(ctx, RADAR_E_UNSUPPORTED, "radar_chirp_zoom_static: num_loops * n_out = %d * %d exceeds the static twiddle table (%d entries); use radar_chirp_zoom",
                         L, n_out, kZoomStaticEntries);
    cf* tw = nullptr;
    if (int rc = zoom_static_table(ctx, &tw)) return rc;          // set-up: also what a call with n_frames == 0 does
    if (n_frames <= 0 || k_begin == k_end) return RADAR_OK;
    const long long grid = (long long)n_frames * n_ant;
    if (grid > 0x7fffffffLL) return set_error(ctx, RADAR_E_INVALID, "too many frames in one call");
    const int n_tw = L * n_out;
    zoom_twiddle_static_kernel<<<(n_tw + 255) / 256, 256, 0, st>>>(turns, L, n_out);
    RB_CUDA_OK(ctx, cudaGetLastError());
    CzArgs a{range_cube, S, L, k_begin, k_end, n_out, tw, out};
    chirp_zoom_kernel<<<(unsigned)grid, kDaThreads, 0, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ---- row reduction of a batch of maps (micro-Doppler: max over the range rows kept) -------------------------------
struct RowReduceArgs {
    const float* maps; int rows, cols; const uint8_t* row_select; int op; float* out;
};

__global__ void __launch_bounds__(256) row_reduce_kernel(const __grid_constant__ RowReduceArgs a) {
    const size_t m = blockIdx.x;
    const float* src = a.maps + m * (size_t)a.rows * a.cols;
    for (int c = threadIdx.x; c < a.cols; c += blockDim.x) {
        float best = -__int_as_float(0x7f800000);
        double sum = 0.0;
        int n = 0;
        bool any_nan = false;
        for (int r = 0; r < a.rows; ++r) {
            if (a.row_select && !a.row_select[r]) continue;
            const float x = src[(size_t)r * a.cols + c];
            any_nan |= (x != x);
            best = fmaxf(best, x);
            sum += (double)x;
            ++n;
        }
        float res;
        if (a.op == RADAR_REDUCE_MAX) res = any_nan ? __int_as_float(0x7fc00000) : best;     // np.max propagates nan
        else res = (float)(sum / (double)n);
        a.out[m * a.cols + c] = res;
    }
}

int launch_row_reduce(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const uint8_t* row_select,
                      int op, float* out, cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return set_error(ctx, RADAR_E_INVALID, "empty map");
    if (op != RADAR_REDUCE_MAX && op != RADAR_REDUCE_MEAN) return set_error(ctx, RADAR_E_INVALID, "unknown reduction %d", op);
    if (n_maps <= 0) return RADAR_OK;
    RowReduceArgs a{maps, rows, cols, row_select, op, out};
    row_reduce_kernel<<<(unsigned)n_maps, 256, 0, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb
