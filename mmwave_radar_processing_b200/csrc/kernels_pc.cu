// K3b: per-detection zero-padded angle DFT + first-max argmax (azimuth and elevation antenna sets) and the
// spherical -> cartesian conversion; one warp per detection, one CTA per frame.
//
// Reference arithmetic reproduced: processors/point_cloud_generator.py:143-248 (gather raw[ant][r][d],
// zero-pad to num_angle_bins, FFT, optional fftshift, abs, np.argmax (first maximum), angle_bins[idx],
// x = r cos(el) cos(az), y = r cos(el) sin(az), z = r sin(el)).
//
// The DFT, the magnitude (hypot) and the trigonometry run in fp64 on the fp32 RD samples: the work is a few
// hundred FMAs per detection and the argmax index has to be bit-exact against the fp64 reference.
#include "radar_internal.h"

namespace rb {

constexpr int kPcThreads = 128;
constexpr int kPcWarps = kPcThreads / 32;
constexpr int kPcMaxAnt = 32;
constexpr int kPcGather = 2048;      // RD samples staged per chunk of detections (16 KB)
constexpr int kPcTwMax = 256;        // angle DFT tables up to this length are staged in shared memory

struct PcArgs {
    const cf* rd;
    int V, S, L;
    const int32_t* dets; const int32_t* counts; int max_dets;
    const int32_t* az_idx; int n_az;
    const int32_t* el_idx; int n_el;
    int shift_az, shift_el, A;
    const double2* tw64;
    const double* range_bins; const double* vel_bins;
    const double* angle_cos; const double* angle_sin;     // cos / sin of angle_bins (host fp64, uploaded with the bins)
    double* points; int32_t* az_bin; int32_t* el_bin;
    const int32_t* offsets;      // optional exclusive scan of the clamped counts: packed output rows
};

// argmax_a | sum_j z[j] w^j |, w = W_A^{a'}, a' = (a - A/2) mod A when shifted; first maximum wins.  z: this detection's n
// gathered RD samples (shared memory), tw: W_A^e in fp64.  The sum is evaluated by Horner's rule in fp64 (one twiddle per
// bin) and |.|^2 is compared (sqrt is monotone, so the ordering is the one of |.|).  NC > 0: antenna count compiled in, the
// samples are converted to fp64 once and stay in registers.
template <int NC>
__device__ __forceinline__ int angle_argmax(const cf* z, int n_rt, int A, bool shift, int lane, const double2* tw) {
    const int n = NC > 0 ? NC : n_rt;
    double zr[NC > 0 ? NC : 1], zi[NC > 0 ? NC : 1];
    if (NC > 0) {
#pragma unroll
        for (int j = 0; j < NC; ++j) { const cf v = z[j]; zr[j] = (double)v.x; zi[j] = (double)v.y; }
    }
    double best = -1.0;
    int best_a = 0x7fffffff;
    for (int ao = lane; ao < A; ao += 32) {
        int ap = ao;
        if (shift) { ap = ao - A / 2; if (ap < 0) ap += A; }
        const double2 w = tw[ap];
        double sr, si;
        if (NC > 0) {
            sr = zr[NC - 1]; si = zi[NC - 1];
#pragma unroll
            for (int j = NC - 2; j >= 0; --j) {
                const double nr = fma(sr, w.x, fma(-si, w.y, zr[j]));
                si = fma(sr, w.y, fma(si, w.x, zi[j]));
                sr = nr;
            }
        } else {
            sr = (double)z[n - 1].x; si = (double)z[n - 1].y;
            for (int j = n - 2; j >= 0; --j) {
                const cf v = z[j];
                const double nr = fma(sr, w.x, fma(-si, w.y, (double)v.x));
                si = fma(sr, w.y, fma(si, w.x, (double)v.y));
                sr = nr;
            }
        }
        const double mag = sr * sr + si * si;
        if (mag > best) { best = mag; best_a = ao; }      // ascending ao per lane: strict > keeps the first
    }
    // warp argmax, ties -> smaller index: non-negative doubles order like their bit patterns, so three warp reductions
    // (high word, low word, index) replace a shuffle tree on 96-bit keys.  Idle lanes (A < 32) carry -1 -> key 0.
    const unsigned long long bits = best >= 0.0 ? (unsigned long long)__double_as_longlong(best) : 0ull;
    const unsigned hi = (unsigned)(bits >> 32), lo = (unsigned)bits;
    const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    return (int)__reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? (unsigned)best_a : 0x7fffffffu);
}

__device__ __forceinline__ int angle_argmax_any(const cf* z, int n, int A, bool shift, int lane, const double2* tw) {
    switch (n) {
        case 1: return angle_argmax<1>(z, n, A, shift, lane, tw);
        case 2: return angle_argmax<2>(z, n, A, shift, lane, tw);
        case 3: return angle_argmax<3>(z, n, A, shift, lane, tw);
        case 4: return angle_argmax<4>(z, n, A, shift, lane, tw);
        case 8: return angle_argmax<8>(z, n, A, shift, lane, tw);
        default: return angle_argmax<0>(z, n, A, shift, lane, tw);
    }
}

// One CTA per frame (grid-stride).  The scattered RD samples of a whole chunk of detections are gathered into shared
// memory first - every thread issues independent loads, so the chunk pays ONE memory latency instead of one per
// detection - then the warps walk the chunk, one detection each.
__global__ void __launch_bounds__(kPcThreads) point_cloud_kernel(const __grid_constant__ PcArgs a, int n_frames) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ cf zs[kPcGather];
    __shared__ double2 tw_s[kPcTwMax];
    const bool tw_in_smem = a.A <= kPcTwMax;
    if (tw_in_smem)
        for (int i = threadIdx.x; i < a.A; i += kPcThreads) tw_s[i] = a.tw64[i];
    const double2* tw = tw_in_smem ? tw_s : a.tw64;
    const int nant = a.n_az + a.n_el;
    const int chunk = nant > 0 ? kPcGather / nant : a.max_dets;
    const long long plane = (long long)a.S * a.L;
    for (int frame = blockIdx.x; frame < n_frames; frame += gridDim.x) {
        const int n = min(a.counts[frame], a.max_dets);
        const long long slot0 = (long long)frame * a.max_dets;
        const long long row0 = a.offsets ? (long long)a.offsets[frame] : slot0;
        const cf* cube = a.rd + (long long)frame * a.V * plane;
        for (int c0 = 0; c0 < n; c0 += chunk) {
            const int m = min(chunk, n - c0);
            __syncthreads();                                  // previous chunk consumed (and tw_s written, first time)
            for (int i = threadIdx.x; i < m * nant; i += kPcThreads) {
                const int slot = i / nant, j = i - slot * nant;
                const int2 rdx = reinterpret_cast<const int2*>(a.dets)[slot0 + c0 + slot];
                const int ant = j < a.n_az ? a.az_idx[j] : a.el_idx[j - a.n_az];
                zs[i] = __ldg(cube + ant * plane + (long long)rdx.x * a.L + rdx.y);
            }
            __syncthreads();
            for (int slot = warp; slot < m; slot += kPcWarps) {
                const cf* z = zs + slot * nant;
                int az_i = 0, el_i = 0;
                if (a.n_az > 0) az_i = angle_argmax_any(z, a.n_az, a.A, a.shift_az != 0, lane, tw);
                if (a.n_el > 0) el_i = angle_argmax_any(z + a.n_az, a.n_el, a.A, a.shift_el != 0, lane, tw);
                if (lane == 0) {
                    const int2 rdx = reinterpret_cast<const int2*>(a.dets)[slot0 + c0 + slot];
                    const double ca = a.n_az > 0 ? a.angle_cos[az_i] : 1.0, sa = a.n_az > 0 ? a.angle_sin[az_i] : 0.0;
                    const double ce = a.n_el > 0 ? a.angle_cos[el_i] : 1.0, se = a.n_el > 0 ? a.angle_sin[el_i] : 0.0;
                    const double rng = a.range_bins[rdx.x], vel = a.vel_bins[rdx.y];
                    double4 p;
                    p.x = rng * ce * ca;
                    p.y = rng * ce * sa;
                    p.z = rng * se;
                    p.w = vel;
                    const long long row = row0 + c0 + slot;
                    reinterpret_cast<double4*>(a.points)[row] = p;
                    if (a.az_bin) a.az_bin[row] = az_i;
                    if (a.el_bin) a.el_bin[row] = el_i;
                }
            }
        }
    }
}

int launch_point_cloud(radar_ctx* ctx, const cf* rd, int n_frames, const int32_t* dets, const int32_t* counts,
                       int max_dets, const int32_t* az_idx, int n_az, const int32_t* el_idx, int n_el, int shift_az,
                       int shift_el, int num_angle_bins, const int32_t* offsets, double* points_out, int32_t* az_bin_out,
                       int32_t* el_bin_out, cudaStream_t st) {
    if (n_az < 0 || n_el < 0 || n_az > kPcMaxAnt || n_el > kPcMaxAnt)
        return set_error(ctx, RADAR_E_INVALID, "antenna sets must hold 0..%d indices", kPcMaxAnt);
    if (num_angle_bins < 1 || num_angle_bins > 8192 || num_angle_bins < n_az || num_angle_bins < n_el)
        return set_error(ctx, RADAR_E_INVALID, "num_angle_bins %d invalid", num_angle_bins);
    if (!ctx->range_bins || !ctx->vel_bins) return set_error(ctx, RADAR_E_INVALID, "radar_ctx_set_axes not called");
    if ((n_az > 0 || n_el > 0) && (!ctx->angle_bins || ctx->n_angle_bins < num_angle_bins))
        return set_error(ctx, RADAR_E_INVALID, "radar_ctx_set_angle_bins not called with >= %d bins", num_angle_bins);
    if (max_dets <= 0) return set_error(ctx, RADAR_E_INVALID, "max_dets must be positive");
    if (n_frames <= 0) return RADAR_OK;
    AngleTables at;
    if (int rc = get_angle_tables(ctx, num_angle_bins, &at)) return rc;
    PcArgs a;
    a.rd = rd; a.V = ctx->V; a.S = ctx->shape.num_samples; a.L = ctx->shape.num_loops;
    a.dets = dets; a.counts = counts; a.max_dets = max_dets;
    a.az_idx = az_idx; a.n_az = n_az; a.el_idx = el_idx; a.n_el = n_el;
    a.shift_az = shift_az; a.shift_el = shift_el; a.A = num_angle_bins;
    a.tw64 = at.tw64;
    a.range_bins = ctx->range_bins; a.vel_bins = ctx->vel_bins;
    a.angle_cos = ctx->angle_cos; a.angle_sin = ctx->angle_sin;
    a.points = points_out; a.az_bin = az_bin_out; a.el_bin = el_bin_out;
    a.offsets = offsets;
    const int grid = n_frames < ctx->sm_count * 64 ? n_frames : ctx->sm_count * 64;
    point_cloud_kernel<<<(unsigned)grid, kPcThreads, 0, st>>>(a, n_frames);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ------------------------------------------------------------------------------------------------
// Exclusive scan of the per-frame detection counts (clamped to max_dets): offsets[f] is the first row of
// frame f in the packed point list, offsets[n_frames] the total.  One CTA; every thread scans a contiguous
// chunk of frames, the chunk totals go through a shared-memory block scan.
// ------------------------------------------------------------------------------------------------
constexpr int kScanThreads = 1024;

__global__ void __launch_bounds__(kScanThreads) det_offsets_kernel(const int32_t* __restrict__ counts, int n_frames,
                                                                   int max_dets, int32_t* __restrict__ offsets) {
    __shared__ int warp_tot[kScanThreads / 32];
    const int chunk = (n_frames + kScanThreads - 1) / kScanThreads;
    const int begin = min(threadIdx.x * chunk, n_frames), end = min(begin + chunk, n_frames);
    int sum = 0;
    for (int f = begin; f < end; ++f) sum += min(max(counts[f], 0), max_dets);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = sum;
    for (int s = 1; s < 32; s <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, s);
        if (lane >= s) incl += o;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = warp_tot[lane];
        for (int s = 1; s < 32; s <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, w, s);
            if (lane >= s) w += o;
        }
        warp_tot[lane] = w;
    }
    __syncthreads();
    int run = incl - sum + (warp > 0 ? warp_tot[warp - 1] : 0);
    for (int f = begin; f < end; ++f) {
        offsets[f] = run;
        run += min(max(counts[f], 0), max_dets);
    }
    if (threadIdx.x == kScanThreads - 1) offsets[n_frames] = warp_tot[kScanThreads / 32 - 1];
}

int launch_det_offsets(radar_ctx* ctx, const int32_t* counts, int n_frames, int max_dets, int32_t* offsets_out,
                       cudaStream_t st) {
    if (max_dets <= 0) return set_error(ctx, RADAR_E_INVALID, "max_dets must be positive");
    if ((long long)n_frames * max_dets > 0x7fffffffLL)
        return set_error(ctx, RADAR_E_INVALID, "n_frames*max_dets overflows the int32 offsets");
    det_offsets_kernel<<<1, kScanThreads, 0, st>>>(counts, n_frames, max_dets, offsets_out);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb
