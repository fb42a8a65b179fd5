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
    double* mirrors[16];         // radar_ctx_set_point_mirrors: every packed row below mirror_cap is stored there too
    int n_mirrors, mirror_cap;   // (peer-mapped gathered buffers: the all-gather of the points, done by these stores)
};

__device__ __forceinline__ void store_point(const PcArgs& a, long long row, double4 p) {
    reinterpret_cast<double4*>(a.points)[row] = p;
    if (a.n_mirrors > 0 && row < a.mirror_cap) {
        for (int m = 0; m < a.n_mirrors; ++m) reinterpret_cast<double4*>(a.mirrors[m])[row] = p;
    }
}

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
    const unsigned arg = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? (unsigned)best_a : 0x7fffffffu);
    return arg < (unsigned)A ? (int)arg : 0;       // every |DFT| NaN (NaN / inf in the cube): no lane ever updated - stay in range
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
                    store_point(a, row, p);
                    if (a.az_bin) a.az_bin[row] = az_i;
                    if (a.el_bin) a.el_bin[row] = el_i;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Thread-per-detection variant (antenna sets of at most 8, power-of-two bin count): the warp-per-detection kernel above
// spends ~230 warp instructions per detection, most of them on the warp reductions, the lane-0 tail and index arithmetic
// (profiles/r1f_ncu_point_cloud_kernel_1843_RadVel.txt).  Here one thread owns a detection end to end.  The zero-padded
// A-point transform of NP inputs (set padded with zeros to a power of two) splits by residue: bin a' = r + (A/NP) q is
// sum_j (z_j W_A^(j r)) W_NP^(j q), so each residue costs NP-1 twiddles and one NP-point butterfly, all in fp64 registers;
// the running first-maximum replaces the reductions, and every lane converts and stores its own point.
// ------------------------------------------------------------------------------------------------
struct D2 { double x, y; };
__device__ __forceinline__ D2 dadd(D2 a, D2 b) { return {a.x + b.x, a.y + b.y}; }
__device__ __forceinline__ D2 dsub(D2 a, D2 b) { return {a.x - b.x, a.y - b.y}; }
__device__ __forceinline__ D2 dmul(D2 a, double2 w) { return {fma(a.x, w.x, -a.y * w.y), fma(a.x, w.y, a.y * w.x)}; }
__device__ __forceinline__ D2 mul_mi(D2 a) { return {a.y, -a.x}; }      // times -i

template <int NP> __device__ __forceinline__ void dft_regs(D2 (&v)[NP]);
template <> __device__ __forceinline__ void dft_regs<1>(D2 (&)[1]) {}
template <> __device__ __forceinline__ void dft_regs<2>(D2 (&v)[2]) {
    const D2 a = v[0], b = v[1];
    v[0] = dadd(a, b); v[1] = dsub(a, b);
}
template <> __device__ __forceinline__ void dft_regs<4>(D2 (&v)[4]) {
    const D2 s02 = dadd(v[0], v[2]), d02 = dsub(v[0], v[2]), s13 = dadd(v[1], v[3]), d13 = mul_mi(dsub(v[1], v[3]));
    v[0] = dadd(s02, s13); v[1] = dadd(d02, d13); v[2] = dsub(s02, s13); v[3] = dsub(d02, d13);
}
template <> __device__ __forceinline__ void dft_regs<8>(D2 (&v)[8]) {
    D2 e[4] = {v[0], v[2], v[4], v[6]}, o[4] = {v[1], v[3], v[5], v[7]};
    dft_regs<4>(e);
    dft_regs<4>(o);
    const double q = 0.70710678118654752440;
    const D2 o1 = {q * (o[1].x + o[1].y), q * (o[1].y - o[1].x)};      // o[1] W8^1
    const D2 o2 = mul_mi(o[2]);                                          // o[2] W8^2
    const D2 o3 = {q * (o[3].y - o[3].x), -q * (o[3].x + o[3].y)};     // o[3] W8^3
    v[0] = dadd(e[0], o[0]); v[4] = dsub(e[0], o[0]);
    v[1] = dadd(e[1], o1);   v[5] = dsub(e[1], o1);
    v[2] = dadd(e[2], o2);   v[6] = dsub(e[2], o2);
    v[3] = dadd(e[3], o3);   v[7] = dsub(e[3], o3);
}

// first argmax over the (optionally shifted) output index of | zero-padded A-point DFT of z[0..n) |, n <= NP
template <int NP>
__device__ __forceinline__ int angle_argmax_thread(const cf* __restrict__ cube, long long plane, long long cell,
                                                   const int32_t* __restrict__ ant, int n, int A, bool shift,
                                                   const double2* tw) {
    D2 z[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        z[j] = {0.0, 0.0};
        if (j < n) { const cf v = __ldg(cube + ant[j] * plane + cell); z[j] = {(double)v.x, (double)v.y}; }
    }
    if (NP == 1) return 0;                                  // every bin has the magnitude of the one sample
    const int M = A / NP, half = shift ? A / 2 : 0;
    double best = -1.0;
    int best_a = 0;
    for (int r = 0; r < M; ++r) {
        D2 v[NP];
        v[0] = z[0];
#pragma unroll
        for (int j = 1; j < NP; ++j) v[j] = dmul(z[j], tw[(j * r) & (A - 1)]);
        dft_regs<NP>(v);
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const double mag = fma(v[q].x, v[q].x, v[q].y * v[q].y);
            const int ao = (r + M * q + half) & (A - 1);
            if (mag > best || (mag == best && ao < best_a)) { best = mag; best_a = ao; }
        }
    }
    return best_a;
}

__device__ __forceinline__ int angle_argmax_thread_any(const cf* cube, long long plane, long long cell, const int32_t* ant,
                                                       int n, int A, bool shift, const double2* tw) {
    if (n <= 1) return angle_argmax_thread<1>(cube, plane, cell, ant, n, A, shift, tw);
    if (n == 2) return angle_argmax_thread<2>(cube, plane, cell, ant, n, A, shift, tw);
    if (n <= 4) return angle_argmax_thread<4>(cube, plane, cell, ant, n, A, shift, tw);
    return angle_argmax_thread<8>(cube, plane, cell, ant, n, A, shift, tw);
}

constexpr int kPcTThreads = 128;

__global__ void __launch_bounds__(kPcTThreads, 8) point_cloud_thread_kernel(const __grid_constant__ PcArgs a, int n_frames) {
    __shared__ double2 tw_s[kPcTwMax];
    const bool tw_in_smem = a.A <= kPcTwMax;
    if (tw_in_smem)
        for (int i = threadIdx.x; i < a.A; i += kPcTThreads) tw_s[i] = a.tw64[i];
    __syncthreads();
    const double2* tw = tw_in_smem ? tw_s : a.tw64;
    const long long plane = (long long)a.S * a.L;
    for (int frame = blockIdx.x; frame < n_frames; frame += gridDim.x) {
        const int n = min(a.counts[frame], a.max_dets);
        const long long slot0 = (long long)frame * a.max_dets;
        const long long row0 = a.offsets ? (long long)a.offsets[frame] : slot0;
        const cf* cube = a.rd + (long long)frame * a.V * plane;
        for (int slot = threadIdx.x; slot < n; slot += kPcTThreads) {
            const int2 rdx = reinterpret_cast<const int2*>(a.dets)[slot0 + slot];
            const long long cell = (long long)rdx.x * a.L + rdx.y;
            int az_i = 0, el_i = 0;
            if (a.n_az > 0) az_i = angle_argmax_thread_any(cube, plane, cell, a.az_idx, a.n_az, a.A, a.shift_az != 0, tw);
            if (a.n_el > 0) el_i = angle_argmax_thread_any(cube, plane, cell, a.el_idx, a.n_el, a.A, a.shift_el != 0, tw);
            const double ca = a.n_az > 0 ? a.angle_cos[az_i] : 1.0, sa = a.n_az > 0 ? a.angle_sin[az_i] : 0.0;
            const double ce = a.n_el > 0 ? a.angle_cos[el_i] : 1.0, se = a.n_el > 0 ? a.angle_sin[el_i] : 0.0;
            const double rng = a.range_bins[rdx.x], vel = a.vel_bins[rdx.y];
            double4 p;
            p.x = rng * ce * ca;
            p.y = rng * ce * sa;
            p.z = rng * se;
            p.w = vel;
            const long long row = row0 + slot;
            store_point(a, row, p);
            if (a.az_bin) a.az_bin[row] = az_i;
            if (a.el_bin) a.el_bin[row] = el_i;
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
    a.n_mirrors = offsets ? ctx->n_point_mirrors : 0;       // mirrors follow the packed layout only
    a.mirror_cap = ctx->point_mirror_cap;
    for (int i = 0; i < 16; ++i) a.mirrors[i] = ctx->point_mirrors[i];
    const int grid = n_frames < ctx->sm_count * 64 ? n_frames : ctx->sm_count * 64;
    const bool pow2_bins = (num_angle_bins & (num_angle_bins - 1)) == 0 && num_angle_bins >= 8;
    if (pow2_bins && n_az <= 8 && n_el <= 8 && ctx->pc_mode != 1)
        point_cloud_thread_kernel<<<(unsigned)grid, kPcTThreads, 0, st>>>(a, n_frames);
    else
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
