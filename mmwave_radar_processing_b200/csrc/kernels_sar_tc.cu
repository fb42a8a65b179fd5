// K5a on the 5th-generation tensor cores: the synthetic-array beamformer as a split-precision FP16 GEMM (tcgen05 + TMEM).
//
// Reference arithmetic (processors/simple_synthetic_array_beamformer_processor_multiFrame.py:499-585):
//   out[:, g] = FFT_S( hanning(S) * sum_n hamming(N)[n] * x[s, n] * exp(+2 pi i (d_g . p_n) / lambda) )
//
// The complex product  beams[g][s] = sum_n W[n][g] X[s][n]  is written as ONE real GEMM  D[M x Nb] = A[M x K] B[K x Nb]:
//   M  = directions (128 per accumulator tile),          K = 2 N  (k = 2n: Re W, k = 2n+1: Im W),
//   Nb = 2 SP columns (SP = S rounded up to 8): column s holds Re beams, column SP + s holds Im beams,
//   B[2n][s] = Xr, B[2n+1][s] = -Xi, B[2n][SP+s] = Xi, B[2n+1][SP+s] = Xr   (X pre-multiplied by hamming(N)[n]).
// * A is never stored in HBM: every thread owns one direction and generates its row of the current K chunk - the phase
//   (d.p)/lambda reaches 1e2..1e4 turns, so it is reduced modulo 1 in fp64 before the fp32 sincospi - and writes it
//   straight into shared memory in the UMMA canonical K-major layout (no swizzle: 8-row x 16-byte core matrices).
// * B is prepared once per call (sar_tc_prep_kernel) in global memory in exactly that layout, chunk by chunk, so a CTA
//   moves a chunk (hi and lo, 48 KB for S = 90, 32 aperture elements) with ONE bulk copy of the TMA engine (cp.async.bulk, completion counted
//   on an mbarrier) - no thread touches it.
// * One 16-bit operand keeps 11 significant bits, far from the 1e-4 bar, so both operands are split  v = hi + lo  (hi = v
//   rounded to FP16, lo = v - hi rounded to FP16) and three MMAs accumulate  hi.hi + hi.lo + lo.hi  in the fp32 TMEM
//   accumulator: relative error ~2^-21 per product, the dropped lo.lo term is ~2^-22.  FP16 rather than TF32 because
//   kind::f16 runs at twice the TF32 rate (K = 16 per instruction instead of 8) for the same 11 + 11 bits.  FP16's narrow
//   exponent is handled by exact power-of-two scales: the steering coefficients (|w| <= 1) are scaled by 2^13, the aperture
//   samples by the power of two that brings the largest |x| just below 2^15 (sar_tc_max_kernel finds it), so hi and lo
//   of everything that matters are normal numbers; the range-FFT kernel multiplies the beams by the inverse power of two
//   (exact), so the scales leave no trace in the result.
// * One thread issues tcgen05.mma; tcgen05.commit arrives on an mbarrier per shared-memory stage, so generating chunk
//   c+1 overlaps the MMAs of chunk c (two stages).  The accumulators (128 lanes x Nb columns per tile, MT tiles per CTA so
//   a B chunk is reused for 128 MT directions) stay in TMEM for the whole K loop and are read once with tcgen05.ld.
// * The tensor core adds into the accumulator with truncation; the K loop is cut into segments that are drained from
//   TMEM and summed in ordinary fp32 (see the drain in sar_tc_kernel).
// * The beams go to a scratch array [S][width] (coalesced: lanes = directions); sar_range_fft_kernel applies Hann and the
//   range FFT per direction with the shared-memory FFT of the range-Doppler family.
#include <cuda_fp16.h>
#include "radar_internal.h"
#include "rd_core.cuh"
#include "fft_radix.cuh"

namespace rb {

constexpr int kTcKC = 64;                 // real k-values (halves) per chunk row = 128 bytes; 4 MMA k-steps of 16
constexpr int kTcElems = kTcKC / 2;       // aperture elements per chunk: one 32-bit word (Re, Im as two halves) each
constexpr int kTcATile = 128 * kTcElems;  // 32-bit words of one A tile (128 rows)
constexpr int kTcLin = 16;                // elements of one "equispaced" unit (two per chunk)
constexpr int kTcFftDirs = 32;
constexpr float kTcAScale = 8192.0f;      // 2^13: steering coefficients as FP16 hi + lo without subnormals

// word index of element n (0..31) of row `row` inside a K-major no-swizzle tile: core matrices of 8 rows x 16 bytes
__host__ __device__ inline int tc_canon(int row, int n) {
    return ((row >> 3) * (kTcElems / 4) + (n >> 2)) * 32 + (row & 7) * 4 + (n & 3);
}

// FP16 hi / lo words of the complex value (re, im): low half = re (k = 2n), high half = im (k = 2n + 1)
__device__ __forceinline__ void split_h2(float re, float im, uint32_t* hi, uint32_t* lo) {
    const __half2 h = __floats2half2_rn(re, im);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(re - hf.x, im - hf.y);
    *hi = *reinterpret_cast<const uint32_t*>(&h);
    *lo = *reinterpret_cast<const uint32_t*>(&l);
}

// Power-of-two scale of the aperture samples from the largest |x| of the call: max * scale in [2^14, 2^15).
__device__ __forceinline__ int sar_scale_exp(float mx) {
    if (!(mx > 0.0f) || mx > 3.0e38f) return 0;             // empty / all-zero / non-finite input: no scaling
    int e;
    frexpf(mx, &e);                                          // mx = m 2^e, m in [1/2, 1)
    e = e < -110 ? -110 : e;                                 // keeps both 2^(15 - e) and its inverse 2^(e - 28) exact floats (e <= 128)
    return 15 - e;
}

// One thread per aperture element n (coalesced rows of the [S][N] cube), blockIdx.y cuts the range rows: the Hamming weight
// of n is formed once per thread (fp64 cospi) and stored as a float for sar_tc_prep_kernel, the largest |x hamming| of
// the call goes to max_bits.
constexpr int kTcMaxThreads = 128;
__global__ void __launch_bounds__(kTcMaxThreads) sar_tc_max_kernel(const cf* __restrict__ aperture, int S, int N, float* __restrict__ hamf,
                                                                   uint32_t* __restrict__ max_bits) {
    __shared__ float red[kTcMaxThreads / 32];
    const int n = blockIdx.x * kTcMaxThreads + threadIdx.x;
    float m = 0.0f;
    if (n < N) {
        const float ham = (float)(N > 1 ? 0.54 - 0.46 * cospi(2.0 * (double)n / (double)(N - 1)) : 1.0);
        if (blockIdx.y == 0) hamf[n] = ham;
        const int rows = (S + gridDim.y - 1) / gridDim.y;
        const int s0 = blockIdx.y * rows, s1 = min(S, s0 + rows);
#pragma unroll 4
        for (int r = s0; r < s1; ++r) {
            const cf x = cscale(__ldg(aperture + (size_t)r * N + n), ham);
            m = fmaxf(m, fmaxf(fabsf(x.x), fabsf(x.y)));     // fmaxf drops NaNs: they reach the image through the MMAs
        }
    }
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kTcMaxThreads / 32; ++w) m = fmaxf(m, red[w]);
        atomicMax(max_bits, __float_as_uint(m));             // non-negative floats order like their bit patterns
    }
}

// ---- operand preparation -----------------------------------------------------------------------------------------
struct SarPrepArgs {
    const cf* aperture;   // [S][N]
    int S, N, SP, NB, n_chunks;
    const double* pos;    // [3][N]
    double inv_lambda;
    uint32_t* bblk;       // [n_chunks][2][NB * kTcElems] words (hi block, lo block)
    double* posl;         // [3][n_chunks * kTcElems]  positions / lambda, zero padded
    float* posf;          // [6][n_chunks * kTcElems]  the same as float pairs: rows 2a = hi, 2a+1 = lo of axis a
    float* stepf;         // [6][n_units]  per-axis element spacing / lambda (float hi / lo pairs) of an equispaced unit
    uint8_t* lin;         // [n_units]     1 = the unit (kTcLin elements) is an arithmetic progression to 1e-9 wavelengths (no padding), 2 = and continues the unit one chunk earlier
    const uint32_t* max_bits;   // largest |x hamming| of the call (float bits), from sar_tc_max_kernel
    const float* hamf;    // [N] np.hamming(N) as floats, from sar_tc_max_kernel
};

// Block (c, y): chunk c of the aperture, quarter y of the B columns; the y = 0 blocks also write the chunk's geometry.
constexpr int kTcPrepSplit = 4;
__global__ void __launch_bounds__(256) sar_tc_prep_kernel(const __grid_constant__ SarPrepArgs a) {
    const int c = blockIdx.x;
    const int npad = a.n_chunks * kTcElems;
    constexpr int kUnits = kTcElems / kTcLin;
    const int n_units = a.n_chunks * kUnits;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (blockIdx.y == 0) {
        for (int i = threadIdx.x; i < 3 * kTcElems; i += blockDim.x) {
            const int axis = i / kTcElems, n = c * kTcElems + (i - axis * kTcElems);
            const double v = n < a.N ? a.pos[(size_t)axis * a.N + n] * a.inv_lambda : 0.0;
            a.posl[(size_t)axis * npad + n] = v;
            const float vh = (float)v;
            a.posf[(size_t)(2 * axis) * npad + n] = vh;
            a.posf[(size_t)(2 * axis + 1) * npad + n] = (float)(v - (double)vh);
        }
        if (warp < kUnits) {
            // One warp per unit u of kTcLin elements, one lane per element.  Flag 1: the unit lies inside the aperture and is
            // an arithmetic progression to 1e-9 wavelengths (constant platform velocity inside a frame): the generators may
            // then step the phase by a constant.  Flag 2: it also continues the progression of the unit one chunk earlier
            // (same step, kTcElems elements further on), so its coefficients are the earlier unit's times
            // exp(2 pi i kTcElems delta).
            const int u = c * kUnits + warp;
            const bool inside = (u + 1) * kTcLin <= a.N, has_prev = u >= kUnits;
            double step[3] = {0.0, 0.0, 0.0};
            bool ok = inside, okp = inside && has_prev, okc = inside && has_prev;
            if (inside) {
                const int e = lane & (kTcLin - 1);
#pragma unroll
                for (int axis = 0; axis < 3; ++axis) {
                    const double* v = a.pos + (size_t)axis * a.N + (size_t)u * kTcLin;
                    const double v0 = v[0] * a.inv_lambda;
                    step[axis] = (v[kTcLin - 1] * a.inv_lambda - v0) / (double)(kTcLin - 1);
                    ok = ok && fabs(v[e] * a.inv_lambda - (v0 + (double)e * step[axis])) <= 1e-9;
                    if (has_prev) {
                        const double* w = v - kTcElems;
                        const double w0 = w[0] * a.inv_lambda;
                        const double prev = (w[kTcLin - 1] * a.inv_lambda - w0) / (double)(kTcLin - 1);
                        okp = okp && fabs(w[e] * a.inv_lambda - (w0 + (double)e * prev)) <= 1e-9;
                        okc = okc && fabs((v[e] - w[e]) * a.inv_lambda - (double)kTcElems * prev) <= 1e-9;
                    }
                }
            }
            const bool lin = __all_sync(0xffffffffu, ok);
            const bool cont = lin && __all_sync(0xffffffffu, okp) && __all_sync(0xffffffffu, okc);
            if (lane == 0) {
                a.lin[u] = lin ? (cont ? 2 : 1) : 0;
                for (int axis = 0; axis < 3; ++axis) {
                    const float sh = (float)step[axis];
                    a.stepf[(size_t)(2 * axis) * n_units + u] = sh;
                    a.stepf[(size_t)(2 * axis + 1) * n_units + u] = (float)(step[axis] - (double)sh);
                }
            }
        }
    }
    const float bscale = ldexpf(1.0f, sar_scale_exp(__uint_as_float(__ldg(a.max_bits))));
    uint32_t* hi = a.bblk + (size_t)c * 2 * a.NB * kTcElems;
    uint32_t* lo = hi + a.NB * kTcElems;
    const int cols = (a.NB + kTcPrepSplit - 1) / kTcPrepSplit;
    const int col0 = blockIdx.y * cols, col1 = min(a.NB, col0 + cols);
    for (int i = col0 * kTcElems + threadIdx.x; i < col1 * kTcElems; i += blockDim.x) {
        const int col = i / kTcElems, nl = i - col * kTcElems;      // consecutive threads: consecutive n (coalesced reads)
        const int n = c * kTcElems + nl;
        const int s = col >= a.SP ? col - a.SP : col;
        cf x = cmake(0.f, 0.f);
        if (s < a.S && n < a.N) x = cscale(__ldg(a.aperture + (size_t)s * a.N + n), __ldg(a.hamf + n));
        const float v0 = col >= a.SP ? x.y : x.x;       // multiplies Re W
        const float v1 = col >= a.SP ? x.x : -x.y;      // multiplies Im W
        const int idx = tc_canon(col, nl);
        split_h2(v0 * bscale, v1 * bscale, hi + idx, lo + idx);
    }
}

// ---- tcgen05 / mbarrier helpers ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t ok = 0, spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
        if (!ok && ++spins > (1u << 27)) __trap();       // a lost arrive must not hang the device
    } while (!ok);
}
// K-major, no swizzle: core matrices of 8 rows x 16 bytes; LBO = 128 B between the two 16-byte K chunks of one MMA,
// SBO = (row bytes / 16) * 128 B between 8-row groups; descriptor version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)((kTcElems / 4) * 128 >> 4) << 32) |
           (1ull << 46);
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
}

// Fractional part of (d . p) in turns as a float pair, WITHOUT the fp64 pipe: d = dh + dl and p = ph + pl per axis as float
// pairs.  dh*ph splits exactly into a rounded product and its FMA remainder; the rounded product (up to 1e4 turns) loses its
// integer part exactly (x - rint(x) is exact below 2^22, rint by the magic-number add); the three fractions are summed with
// error-free TwoSums, everything below goes into the low word.  th + tl is good to ~1e-9 turns; |th| <= 1/2 + ulp.
__device__ __forceinline__ void phase_pair(const float (&dh)[3], const float (&dl)[3], const float (&ph)[3], const float (&pl)[3],
                                           float* th, float* tl) {
    float fr[3], small = 0.0f;
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        const float hi = dh[ax] * ph[ax];
        const float lo = fmaf(dh[ax], ph[ax], -hi);
        fr[ax] = hi - ((hi + 12582912.0f) - 12582912.0f);
        small += fmaf(dh[ax], pl[ax], fmaf(dl[ax], ph[ax], lo));
    }
    const float s1 = fr[0] + fr[1], b1 = s1 - fr[0], e1 = (fr[0] - (s1 - b1)) + (fr[1] - b1);
    const float s2 = s1 + fr[2], b2 = s2 - s1, e2 = (s1 - (s2 - b2)) + (fr[2] - b2);
    const float h = s2 - ((s2 + 12582912.0f) - 12582912.0f);
    const float l = (e1 + e2) + small;
    const float t = h + l;
    *th = t;
    *tl = (h - t) + l;
}

// exp(2 pi i t) for t in [-1/2, 1/2] turns: quadrant by the magic-number round (no conversion-unit instructions), Taylor
// polynomials on |2 pi r| <= pi/4 (sin to x^9, cos to x^8: truncation error < 3e-8), quadrant rotation by sign-bit XORs.
__device__ __forceinline__ void sincos_turns(float t, float* sn, float* cs) {
    const float m = fmaf(4.0f, t, 12582912.0f);                 // 1.5 * 2^23: the low mantissa bits hold rint(4 t)
    const int q = __float_as_int(m);
    const float r = fmaf(-0.25f, m - 12582912.0f, t);           // |r| <= 1/8 turn
    const float r2 = r * r;
    float s = fmaf(r2, 42.0586939448976f, -76.7058597530614f);          // (2pi)^9/9! = 42.0586939, -(2pi)^7/7!
    s = fmaf(r2, s, 81.6052492760750f);                                                // (2pi)^5/5!
    s = fmaf(r2, s, -41.3417022403998f);                                               // -(2pi)^3/3!
    s = fmaf(r2, s, 6.28318530717959f);
    s *= r;
    float c = fmaf(r2, 60.2446413718763f, -85.4568172066937f);                         // (2pi)^8/8!, -(2pi)^6/6!
    c = fmaf(r2, c, 64.9393940226683f);                                                // (2pi)^4/4!
    c = fmaf(r2, c, -19.7392088021787f);                                               // -(2pi)^2/2!
    c = fmaf(r2, c, 1.0f);
    const bool swap = q & 1;
    const float c0 = swap ? s : c, s0 = swap ? c : s;
    *cs = __int_as_float(__float_as_int(c0) ^ (((q + 1) & 2) << 30));
    *sn = __int_as_float(__float_as_int(s0) ^ ((q & 2) << 30));
}

// ---- the GEMM ---------------------------------------------------------------------------------------------------
struct SarTcArgs {
    const uint32_t* bblk; // prepared B chunks (FP16 pairs)
    const double* posl;   // [3][npad] positions / lambda
    const float* posf;    // [6][npad] hi / lo float pairs of the same
    const float* stepf;   // [6][n_units] element spacing / lambda of the equispaced units, float hi / lo pairs
    const uint8_t* lin;   // [n_units] unit of kTcLin elements is equispaced (null: never use the stepping path)
    const double* dirs;   // [3][G]
    int npad, n_chunks, seg_chunks, G, S, SP, NB;
    int g_begin, g_end;
    cf* beams;            // [S][g_end - g_begin]; split-K: one such array per accumulation segment
    // split-K over gridDim.y: CTA row y owns the accumulation segments [y*segs_per_split, (y+1)*segs_per_split).  With
    // ksplit > 1 every segment is drained into ITS OWN array and sar_range_fft_kernel adds the arrays in segment order - the
    // same fp32 additions in the same order as the in-place accumulation of ksplit == 1, so the image does not depend on
    // how the aperture is cut over CTAs (slices of the grid, which pick another ksplit, stay bit-identical to the full grid).
    int ksplit, segs_per_split;
    int parts;            // 1: one array per segment (always with ksplit > 1; RADAR_B200_SAR_PARTS=1 forces it for ksplit == 1)
    size_t part_stride;   // cells between two per-segment arrays (S of the whole call x width; S here is one column tile of it)
    int skip_drain_stores;   // RADAR_B200_SAR_DBG & 64: timing experiment, the drains read TMEM but do not touch `beams` (image invalid)
};

// MT accumulator tiles (128 directions each) per CTA, TPR threads per direction row (they split the elements of a chunk and
// the columns of a drain).  Warp roles: GW = 4 MT TPR generator warps (steering rows -> shared memory, drains), one MMA
// warp (a single lane issues tcgen05.mma / commit) and one TMA warp (a single lane issues the bulk copies of B).  The roles
// meet only through mbarriers - no CTA-wide barrier inside the K loop:
//   empty[s]   tcgen05.commit: the MMAs that read stage s are complete            -> generators, TMA warp
//   full[s]    one arrive per generator warp: the A rows of stage s are written    -> MMA warp
//   bfull[s]   TMA transaction bytes: the B chunk of stage s has landed            -> MMA warp
//   done       tcgen05.commit at the end of an accumulation segment               -> generators (drain)
//   drained    one arrive per generator warp: the accumulators have been read      -> MMA warp (next segment may overwrite)
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int MT, int TPR, int UNR, bool CS>
__global__ void __launch_bounds__(128 * MT * TPR + 64, 1) sar_tc_kernel(const __grid_constant__ SarTcArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int kGenThreads = 128 * MT * TPR, kGenWarps = kGenThreads / 32;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int b_words = a.NB * kTcElems;                       // one of B_hi / B_lo
    const int stage_words = MT * 2 * kTcATile + 2 * b_words;
    uint32_t* stage0 = reinterpret_cast<uint32_t*>(smem_raw);
    uint64_t* bars = reinterpret_cast<uint64_t*>(stage0 + 2 * stage_words);
    uint64_t *empty = bars, *full = bars + 2, *bfull = bars + 4, *done = bars + 6, *drained = bars + 7;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

    if (tid == 0) {
        mbar_init(&empty[0], 1); mbar_init(&empty[1], 1);
        mbar_init(&full[0], kGenWarps); mbar_init(&full[1], kGenWarps);
        mbar_init(&bfull[0], 1); mbar_init(&bfull[1], 1);
        mbar_init(done, 1); mbar_init(drained, kGenWarps);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {                                           // one warp allocates (and later frees) the accumulators
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(256 * MT) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    const int width = a.g_end - a.g_begin;
    // this CTA row's share of the aperture: whole accumulation segments
    const int seg0 = blockIdx.y * a.segs_per_split;
    const int c0 = seg0 * a.seg_chunks;
    const int n_local = min(a.n_chunks, (seg0 + a.segs_per_split) * a.seg_chunks) - c0;

    if (warp == kGenWarps) {
        // ===== MMA issuer =====
        if (lane == 0) {
            // M = 128, N = NB, fp32 accumulate, FP16 x FP16, both operands K-major
            const uint32_t idesc = (1u << 4) | ((uint32_t)(a.NB >> 3) << 17) | ((128u >> 4) << 24);
            int n_seg = 0;
            for (int c = 0; c < n_local; ++c) {
                const int s = c & 1, c_seg = c % a.seg_chunks;
                const bool seg_end = (c_seg == a.seg_chunks - 1) || (c == n_local - 1);
                uint32_t* st = stage0 + s * stage_words;
                mbar_wait(&full[s], (c >> 1) & 1);
                mbar_wait(&bfull[s], (c >> 1) & 1);
                if (c_seg == 0 && c > 0) mbar_wait(drained, (n_seg - 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t b_hi = smem_u32(st + MT * 2 * kTcATile), b_lo = b_hi + b_words * 4;
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    const uint32_t ah = smem_u32(st + m * kTcATile), al = smem_u32(st + MT * kTcATile + m * kTcATile);
                    const uint32_t d = tmem_base + m * 256;
#pragma unroll
                    for (int j = 0; j < kTcKC / 16; ++j) {                 // 16 k-values = two 16-byte chunks = 256 B further on
                        const uint32_t off = j * 256;
                        umma_f16(d, umma_desc(al + off), umma_desc(b_hi + off), idesc, (c_seg | j) != 0);
                        umma_f16(d, umma_desc(ah + off), umma_desc(b_lo + off), idesc, 1);
                        umma_f16(d, umma_desc(ah + off), umma_desc(b_hi + off), idesc, 1);
                    }
                }
                umma_commit(&empty[s]);
                if (seg_end) { umma_commit(done); ++n_seg; }
            }
        }
    } else if (warp == kGenWarps + 1) {
        // ===== B loader: one bulk copy (hi and lo, canonical layout) per chunk =====
        // Lanes 8..20 pull the NEXT chunk's geometry (positions, steps, flags: what every generator thread is about to
        // load, the same addresses for the whole CTA) into L1 one chunk ahead of the generators.
        const int n_units = a.n_chunks * (kTcElems / kTcLin);
        const int row6 = (lane - 8) % 6, which = lane < 8 ? -1 : (lane - 8) / 6;
        const uint32_t bytes = 2u * (uint32_t)b_words * 4u;
        for (int c = 0; c < n_local; ++c) {
            const int s = c & 1;
            if (c >= 2) mbar_wait(&empty[s], ((c >> 1) - 1) & 1);               // warp uniform
            if (lane == 0) {
                const uint32_t bar = smem_u32(&bfull[s]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(stage0 + s * stage_words + MT * 2 * kTcATile)), "l"(a.bblk + (size_t)(c0 + c) * 2 * b_words),
                               "r"(bytes), "r"(bar) : "memory");
            } else if (which >= 0 && c + 1 < n_local) {
                const int cc = c0 + c + 1;
                const void* ptr = nullptr;
                if (which == 0) ptr = a.posf + (size_t)row6 * a.npad + (size_t)cc * kTcElems;        // 128 bytes = one line
                else if (which == 1) ptr = a.stepf + (size_t)row6 * n_units + (size_t)cc * (kTcElems / kTcLin);
                else if (lane == 20 && a.lin) ptr = a.lin + (size_t)cc * (kTcElems / kTcLin);
                if (ptr) asm volatile("prefetch.global.L1 [%0];" ::"l"(ptr));
            }
            __syncwarp();
        }
    } else {
        // ===== generators =====
        const int part = tid / (128 * MT), rt = tid - part * (128 * MT);
        const int mt = rt >> 7, row = rt & 127;
        // this thread's direction (clamped: surplus rows of the last tile compute a valid direction and are not written)
        const int g_out = blockIdx.x * (128 * MT) + mt * 128 + row;            // column of `beams`
        const int g = a.g_begin + min(g_out, width - 1);
        const double dx = a.dirs[g], dy = a.dirs[(size_t)a.G + g], dz = a.dirs[2 * (size_t)a.G + g];
        // float pairs of the direction for the fp32 phase (CS): d = dh + dl to 2^-48
        const float dh[3] = {(float)dx, (float)dy, (float)dz};
        const float dl[3] = {(float)(dx - (double)dh[0]), (float)(dy - (double)dh[1]), (float)(dz - (double)dh[2])};
        constexpr int W = kTcElems / TPR;                          // elements of a chunk this thread generates
        static_assert(W % kTcLin == 0, "a thread generates whole units of kTcLin elements");
        const int n_units = a.n_chunks * (kTcElems / kTcLin);
        const int row_word = ((row >> 3) * (kTcElems / 4)) * 32 + (row & 7) * 4;     // word of element 0 of this row
        int n_drained = 0;
        // Equispaced units: phase(e) = t0 + e * delta.  Per unit the thread keeps the first coefficient w0 = exp(2 pi i t0)
        // (scaled by 2^13) and the powers exp(2 pi i delta 2^k), k = 0..3 and 5, in registers from one chunk to the next.
        // A unit that continues the progression of the unit one chunk earlier (flag 2) only multiplies w0 by the k = 5
        // power (32 elements further on) - no phase, no sincos.  The chain restarts from the phases (flag 1 path) at the
        // start of every accumulation segment and then every 8 chunks, so at most 7 such multiplies (~1e-7 each) stack
        // up, the same ones whatever the tiling or the split of the aperture over CTAs.  The 16 coefficients of the unit
        // are w0 times at most four of the powers (a ladder of 15 complex multiplies): ~7 instructions per element
        // instead of ~46 for the general path.
        constexpr int U = W / kTcLin;
        float w0r[U], w0i[U], pwr[U][5], pwi[U][5];
#pragma unroll
        for (int i = 0; i < U; ++i) {
            w0r[i] = 0.0f; w0i[i] = 0.0f;
#pragma unroll
            for (int k = 0; k < 5; ++k) { pwr[i][k] = 1.0f; pwi[i][k] = 0.0f; }
        }
        for (int cl = 0; cl < n_local; ++cl) {
            const int c = c0 + cl;                                 // chunk of the aperture; cl counts this CTA's stages
            const int s = cl & 1, c_seg = cl % a.seg_chunks;
            const bool seg_end = (c_seg == a.seg_chunks - 1) || (cl == n_local - 1);
            uint32_t* st = stage0 + s * stage_words;
            uint32_t* a_hi = st + mt * kTcATile + row_word;
            uint32_t* a_lo = st + MT * kTcATile + mt * kTcATile + row_word;
            if (cl >= 2) mbar_wait(&empty[s], ((cl >> 1) - 1) & 1);          // the MMAs of stage cl-2 have finished reading this stage
            const double* px = a.posl + (size_t)c * kTcElems;
            const float* pf = a.posf + (size_t)c * kTcElems;
#pragma unroll
            for (int ul = 0; ul < U; ++ul) {
                const int nl0 = part * W + ul * kTcLin;            // first element (inside the chunk) of this unit
                const int unit = (c * kTcElems + nl0) / kTcLin;
                float wr[kTcLin], wi[kTcLin];                      // the coefficients, scaled by 2^13
                const int flag = (CS && a.lin) ? a.lin[unit] : 0;
                if (flag) {
                    if (flag == 2 && (c_seg & 7) != 0) {
                        const float xr = w0r[ul], xi = w0i[ul];
                        w0r[ul] = fmaf(xr, pwr[ul][4], -xi * pwi[ul][4]);
                        w0i[ul] = fmaf(xr, pwi[ul][4], xi * pwr[ul][4]);
                    } else {
                        // the unit's first phase and delta as float pairs (phase_pair), exp(2 pi i delta 2^k) by sincos_turns
                        // of the exactly doubled and re-reduced pair (no squaring error); the six sincos chains are independent
                        float sh[3], sl[3], ph[3], pl[3];
#pragma unroll
                        for (int ax = 0; ax < 3; ++ax) {
                            sh[ax] = __ldg(a.stepf + (size_t)(2 * ax) * n_units + unit);
                            sl[ax] = __ldg(a.stepf + (size_t)(2 * ax + 1) * n_units + unit);
                            ph[ax] = __ldg(pf + (size_t)(2 * ax) * a.npad + nl0);
                            pl[ax] = __ldg(pf + (size_t)(2 * ax + 1) * a.npad + nl0);
                        }
                        float uh, ul2, t0h, t0l;
                        phase_pair(dh, dl, sh, sl, &uh, &ul2);
                        phase_pair(dh, dl, ph, pl, &t0h, &t0l);
#pragma unroll
                        for (int k = 0; k < 6; ++k) {
                            if (k != 4) sincos_turns(uh + ul2, &pwi[ul][k < 4 ? k : 4], &pwr[ul][k < 4 ? k : 4]);
                            uh *= 2.0f; ul2 *= 2.0f;
                            uh -= (uh + 12582912.0f) - 12582912.0f;
                        }
                        sincos_turns(t0h + t0l, &w0i[ul], &w0r[ul]);
                        w0r[ul] *= kTcAScale; w0i[ul] *= kTcAScale;      // exact; every product inherits the scale
                    }
                    wr[0] = w0r[ul]; wi[0] = w0i[ul];
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
#pragma unroll
                        for (int e = 0; e < (1 << k); ++e) {
                            wr[(1 << k) + e] = fmaf(wr[e], pwr[ul][k], -wi[e] * pwi[ul][k]);
                            wi[(1 << k) + e] = fmaf(wr[e], pwi[ul][k], wi[e] * pwr[ul][k]);
                        }
                    }
                } else {
#pragma unroll UNR
                    for (int e = 0; e < kTcLin; ++e) {
                        const int nl = nl0 + e;
                        float sn, cs;
                        if (CS) {
                            // Phase (d.p)/lambda modulo 1 WITHOUT the fp64 pipe (ncu: a third of all stall samples were
                            // math-pipe throttle on DFMA / DMUL).  d = dh + dl, p = ph + pl as float pairs; dh*ph is split
                            // exactly into a rounded product and its FMA remainder; the rounded product (up to 1e4 turns)
                            // loses its integer part exactly (x - rint(x) is exact below 2^22, rint by the magic-number
                            // add); what is left are terms below one turn, summed in fp32: ~1e-7 turns of error.
                            float acc = 0.0f, small = 0.0f;
#pragma unroll
                            for (int ax = 0; ax < 3; ++ax) {
                                const float ph = __ldg(pf + (size_t)(2 * ax) * a.npad + nl), pl = __ldg(pf + (size_t)(2 * ax + 1) * a.npad + nl);
                                const float hi = dh[ax] * ph;
                                const float lo = fmaf(dh[ax], ph, -hi);
                                const float ri = (hi + 12582912.0f) - 12582912.0f;
                                acc += hi - ri;
                                small += fmaf(dh[ax], pl, fmaf(dl[ax], ph, lo));
                            }
                            float t = acc + small;
                            t -= (t + 12582912.0f) - 12582912.0f;
                            sincos_turns(t, &sn, &cs);
                        } else {
                            const double turns = fma(dx, __ldg(px + nl), fma(dy, __ldg(px + a.npad + nl), dz * __ldg(px + 2 * (size_t)a.npad + nl)));
                            const float frac = (float)(turns - rint(turns));
                            sincospif(2.0f * frac, &sn, &cs);
                        }
                        wr[e] = cs * kTcAScale; wi[e] = sn * kTcAScale;
                    }
                }
                // FP16 hi / lo words of the coefficients: 16-byte core-matrix rows of four elements
#pragma unroll
                for (int q = 0; q < kTcLin; q += 4) {
                    uint4 h, l;
                    split_h2(wr[q], wi[q], &h.x, &l.x);
                    split_h2(wr[q + 1], wi[q + 1], &h.y, &l.y);
                    split_h2(wr[q + 2], wi[q + 2], &h.z, &l.z);
                    split_h2(wr[q + 3], wi[q + 3], &h.w, &l.w);
                    const int idx = ((nl0 + q) >> 2) * 32;
                    *reinterpret_cast<uint4*>(a_hi + idx) = h;
                    *reinterpret_cast<uint4*>(a_lo + idx) = l;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[s]);
            if (seg_end) {
                // The tensor core adds into the fp32 accumulator with truncation, a bias that grows with the length of the
                // accumulation chain (measured 1.3e-8 N relative).  The chain is therefore cut into segments: each one is
                // drained from TMEM and summed into `beams` with ordinary round-to-nearest fp32 adds.
                mbar_wait(done, n_drained & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                // TMEM lane = direction row, column s = Re, column SP + s = Im
                const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + mt * 256;
                for (int s0 = part * 16; s0 < a.SP; s0 += 16 * TPR) {
                    uint32_t re[16], im[16];
                    tmem_ld16(taddr + s0, re);
                    tmem_ld16(taddr + a.SP + s0, im);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (g_out < width && !a.skip_drain_stores) {
                        // ksplit > 1: this segment's own array, plain stores; ksplit == 1: accumulate in place - the first
                        // segment stores, the later ones add with fire-and-forget reductions (red.global.add, round to
                        // nearest in the L2): nothing is loaded, and the additions to one address come from this one
                        // thread in program order, so the sum is the same sequence of fp32 adds as the per-segment arrays.
                        cf* dst = a.beams + (a.parts ? (size_t)(seg0 + n_drained) * a.part_stride : (size_t)0) + (size_t)s0 * width + g_out;
                        if (n_drained > 0 && !a.parts) {
#pragma unroll
                            for (int i = 0; i < 16; ++i)
                                if (s0 + i < a.S)
                                    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst + (size_t)i * width), "f"(__uint_as_float(re[i])),
                                                 "f"(__uint_as_float(im[i])) : "memory");
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i)
                                if (s0 + i < a.S) __stcg(dst + (size_t)i * width, cmake(__uint_as_float(re[i]), __uint_as_float(im[i])));
                        }
                    }
                }
                ++n_drained;
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(drained);
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256 * MT) : "memory");
}

// ---- Hann + range FFT of every beam -----------------------------------------------------------------------------
struct SarFftArgs {
    const cf* beams;      // [n_part][S][width]: per-segment arrays (split-K) or one accumulated array
    int n_part;
    int S, width;
    const float* winS; const cf* twS; const uint16_t* posS;
    cf* out;              // [S][width]
    FftPlan planS;
    FastDiv divDirs;
    const uint32_t* max_bits;   // the operand scales of the GEMM are undone here (exact powers of two)
    // radar_ctx_set_image_mirrors: every cell is also stored at mirrors[i][k * mstride + mcol0 + column] - the full images of
    // all ranks over NVLink peer memory (the all-gather of the tiles, done by the stores of this kernel)
    cf* mirrors[16];
    int n_mirrors;
    long long mstride, mcol0;
};

__global__ void __launch_bounds__(256) sar_range_fft_kernel(const __grid_constant__ SarFftArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = a.S, Gp = kTcFftDirs | 1;
    SmemArena ar(smem_raw);
    cf* tile = ar.take<cf>(S * Gp);
    cf* tw = ar.take<cf>(S);
    float* win = ar.take<float>(S);
    uint16_t* pos = ar.take<uint16_t>(S);
    load_table(tw, a.twS, S);
    load_table(win, a.winS, S);
    load_table(pos, a.posS, S);
    __syncthreads();
    const float descale = ldexpf(1.0f, -13 - sar_scale_exp(__uint_as_float(__ldg(a.max_bits))));
    const int g0 = blockIdx.x * kTcFftDirs;
    const int cells = S * kTcFftDirs;
    for (int i = threadIdx.x; i < cells; i += blockDim.x) {
        const int s = i / kTcFftDirs, gl = i - s * kTcFftDirs;
        cf v = cmake(0.f, 0.f);
        if (g0 + gl < a.width) {
            const cf* src = a.beams + (size_t)s * a.width + g0 + gl;
            v = src[0];
            for (int k = 1; k < a.n_part; ++k) {               // segment order, the additions of the in-place accumulation
                const cf u = src[(size_t)k * a.S * a.width];
                v = cmake(u.x + v.x, u.y + v.y);
            }
        }
        tile[s * Gp + gl] = cscale(cscale(v, descale), win[s]);
    }
    __syncthreads();
    fft_axis(tile, Gp, 1, kTcFftDirs, a.divDirs, a.planS, tw);
    for (int i = threadIdx.x; i < cells; i += blockDim.x) {
        const int k = i / kTcFftDirs, gl = i - k * kTcFftDirs;
        if (g0 + gl < a.width) {
            const cf v = tile[pos[k] * Gp + gl];
            a.out[(size_t)k * a.width + g0 + gl] = v;
            for (int m = 0; m < a.n_mirrors; ++m) a.mirrors[m][(size_t)k * a.mstride + a.mcol0 + g0 + gl] = v;
        }
    }
}

// The same for S = 2 * H with H odd and a register transform Dft<H> (S = 90: the RadSAR cfgs): no shared-memory tile at
// all.  Lanes are directions (every load and store is a coalesced row segment of [S][width]); lanes l and l ^ 16 work on
// the same direction and are the two output classes k = k1 (mod 2) of the prime-factor split 2 x H:
//   X[(H k1 + (H + 1) k2) mod S] = sum_n2 ( v[2 n2 mod S] + (-1)^k1 v[(H + 2 n2) mod S] ) W_H^(n2 k2)
// Lane k1 loads the H samples v[(H k1 + 2 n2) mod S] (every sample is read once), gets the partner's by shuffle, forms
// the H sums / differences and runs one H-point transform in registers.
constexpr int kTcFftRegDirs = 64;
template <int H>
__global__ void __launch_bounds__(2 * kTcFftRegDirs) sar_range_fft_reg_kernel(const __grid_constant__ SarFftArgs a) {
    constexpr int S = 2 * H;
    __shared__ float win[S];
    for (int i = threadIdx.x; i < S; i += blockDim.x) win[i] = a.winS[i];
    __syncthreads();
    const float descale = ldexpf(1.0f, -13 - sar_scale_exp(__uint_as_float(__ldg(a.max_bits))));
    const int lane = threadIdx.x & 31, k1 = lane >> 4;
    const int g_raw = blockIdx.x * kTcFftRegDirs + (threadIdx.x >> 5) * 16 + (lane & 15);
    const int g = min(g_raw, a.width - 1);                                // surplus lanes repeat the last direction, no store
    const cf* src = a.beams + g;
    const size_t part_stride = (size_t)S * a.width;
    cf y[H];
#pragma unroll
    for (int n2 = 0; n2 < H; ++n2) {
        const int s_mine = k1 ? (H + 2 * n2) % S : (2 * n2) % S;
        cf v = __ldcg(src + (size_t)s_mine * a.width);
        if (a.n_part > 1)
            for (int k = 1; k < a.n_part; ++k) {                          // segment order, the additions of the in-place accumulation
                const cf u = __ldcg(src + (size_t)k * part_stride + (size_t)s_mine * a.width);
                v = cmake(u.x + v.x, u.y + v.y);
            }
        v = cscale(cscale(v, descale), win[s_mine]);
        const cf o = cmake(__shfl_xor_sync(0xffffffffu, v.x, 16), __shfl_xor_sync(0xffffffffu, v.y, 16));
        y[n2] = k1 ? cmake(o.x - v.x, o.y - v.y) : cmake(v.x + o.x, v.y + o.y);
    }
    Dft<H>::run(y);
    if (g_raw >= a.width) return;
    cf* dst = a.out + g;
#pragma unroll
    for (int k2 = 0; k2 < H; ++k2) {
        const int k = (H * k1 + (H + 1) * k2) % S;
        dst[(size_t)k * a.width] = y[k2];
    }
    for (int m = 0; m < a.n_mirrors; ++m) {
        cf* md = a.mirrors[m] + a.mcol0 + g;
#pragma unroll
        for (int k2 = 0; k2 < H; ++k2) {
            const int k = (H * k1 + (H + 1) * k2) % S;
            md[(size_t)k * a.mstride] = y[k2];
        }
    }
}

// 1 = shape not handled here (caller falls back to the FP32-pipe kernel), 0 = launched, < 0 = error
int launch_sar_image_tc(radar_ctx* ctx, const cf* aperture, int S, int N, const double* positions, const double* dirs,
                        int G, double lambda_m, int g_begin, int g_end, cf* out, cudaStream_t st) {
    // tcgen05.mma with M = 128 takes N in 16..256 (multiples of 16) and N = 2 * samples here, so more than 128 range samples
    // are cut into column tiles of 96 samples (192 columns: two accumulator tiles of 128 directions still fit the shared
    // memory stages and TMEM, as for S = 90); every tile is one pass of prep + GEMM over the whole aperture into its rows
    // of the beams, the range FFT then runs once over all S.
    const int ST = S <= 128 ? S : 96;
    const int n_tiles = (S + ST - 1) / ST;
    const int SP = (ST + 7) & ~7, NB = 2 * SP;                  // the widest tile: sizes the stages and the B blocks
    if (NB > 256 || (NB & 15) != 0) return 1;
    const int width = g_end - g_begin;
    const int n_chunks = (N + kTcElems - 1) / kTcElems, npad = n_chunks * kTcElems;
    // segment length: the truncation bias grows linearly with it (measured on S = 90, N = 2000 against fp64: 2.1e-6 relative
    // at 288 elements, 3.5e-6 at 512, 6.9e-6 at 1024, 1.3e-5 undivided; tools/sar_seg_error.py); keep it near 4e-6
    int seg_chunks = (int)(1.05e6 / (double)N) / kTcElems;
    seg_chunks = seg_chunks < 2 ? 2 : (seg_chunks > n_chunks ? n_chunks : seg_chunks);
    if (ctx->sar_seg_chunks > 0) seg_chunks = ctx->sar_seg_chunks;
    const int n_seg = (n_chunks + seg_chunks - 1) / seg_chunks;
    auto smem_for = [&](int tiles) {
        const size_t stage_words = (size_t)tiles * 2 * kTcATile + 2 * (size_t)NB * kTcElems;
        return 2 * stage_words * sizeof(uint32_t) + 8 * sizeof(uint64_t) + 16;
    };
    // two accumulator tiles per CTA (a B chunk serves 256 directions) once every SM has work anyway and the stages fit
    int mt = (width > 128 * ctx->sm_count && smem_for(2) <= (size_t)ctx->max_smem_optin) ? 2 : 1;
    if (ctx->sar_mode == 2) mt = 1;
    if (ctx->sar_mode == 3 && smem_for(2) <= (size_t)ctx->max_smem_optin) mt = 2;
    const size_t smem = smem_for(mt);
    if (smem > (size_t)ctx->max_smem_optin) return 1;
    const int per_cta = 128 * mt;
    const unsigned grid = (unsigned)((width + per_cta - 1) / per_cta);
    // split-K: every CTA walks its share of the aperture one chunk at a time (~1.2 us per chunk, whatever the number of
    // directions), so a grid that leaves SMs empty - a voxel slice of a multi-GPU run - is cut along the aperture too, in
    // whole accumulation segments (see SarTcArgs: the image does not depend on the cut).  RADAR_B200_SAR_KSPLIT forces it.
    int want_split = (int)(ctx->sm_count / (grid > 0 ? grid : 1));
    if (ctx->sar_ksplit > 0) want_split = ctx->sar_ksplit;
    if (want_split > n_seg) want_split = n_seg;
    if (want_split < 1) want_split = 1;
    const int segs_per_split = (n_seg + want_split - 1) / want_split;
    const int ksplit = (n_seg + segs_per_split - 1) / segs_per_split;
    const int n_part = (ksplit > 1 || ctx->sar_parts) ? n_seg : 1;

    const size_t bblk_bytes = (size_t)n_chunks * 2 * NB * kTcElems * sizeof(uint32_t);
    const int n_units = n_chunks * (kTcElems / kTcLin);
    const size_t posl_bytes = (size_t)3 * npad * sizeof(double);
    const size_t posf_bytes = (size_t)6 * npad * sizeof(float);
    const size_t beams_bytes = (size_t)n_part * S * width * sizeof(cf);
    const size_t step_bytes = (size_t)6 * n_units * sizeof(float), lin_bytes = (size_t)n_units;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t ham_bytes = (size_t)N * sizeof(float);
    if (int rc = ensure_scratch(ctx, up(bblk_bytes) + up(posl_bytes) + up(posf_bytes) + up(beams_bytes) + up(step_bytes) + up(lin_bytes) + 256 + up(ham_bytes))) return rc;
    unsigned char* base = reinterpret_cast<unsigned char*>(ctx->scratch);
    uint32_t* bblk = reinterpret_cast<uint32_t*>(base);
    double* posl = reinterpret_cast<double*>(base + up(bblk_bytes));
    float* posf = reinterpret_cast<float*>(base + up(bblk_bytes) + up(posl_bytes));
    cf* beams = reinterpret_cast<cf*>(base + up(bblk_bytes) + up(posl_bytes) + up(posf_bytes));
    float* stepf = reinterpret_cast<float*>(base + up(bblk_bytes) + up(posl_bytes) + up(posf_bytes) + up(beams_bytes));
    uint8_t* lin = reinterpret_cast<uint8_t*>(stepf) + up(step_bytes);
    uint32_t* max_bits = reinterpret_cast<uint32_t*>(lin + up(lin_bytes));
    float* hamf = reinterpret_cast<float*>(lin + up(lin_bytes) + 256);

    // largest |x hamming| of the call -> power-of-two scale of the FP16 B operand
    RB_CUDA_OK(ctx, cudaMemsetAsync(max_bits, 0, sizeof(uint32_t), st));
    {
        const int slices = S >= 64 ? 8 : (S >= 8 ? 4 : 1);
        sar_tc_max_kernel<<<dim3((unsigned)((N + kTcMaxThreads - 1) / kTcMaxThreads), (unsigned)slices), kTcMaxThreads, 0, st>>>(aperture, S, N, hamf, max_bits);
        RB_CUDA_OK(ctx, cudaGetLastError());
    }
    const size_t fsmem = arena_bytes((size_t)S * (kTcFftDirs | 1), sizeof(cf)) + arena_bytes(S, sizeof(cf)) + arena_bytes(S, 4) + arena_bytes(S, 2);
    if (fsmem > (size_t)ctx->max_smem_optin) return 1;          // the range FFT of the beams keeps S x 33 cells in shared memory
    for (int t = 0; t < n_tiles; ++t) {
        const int s_base = t * ST, S_t = (S - s_base < ST) ? S - s_base : ST, SP_t = (S_t + 7) & ~7, NB_t = 2 * SP_t;
        SarPrepArgs p{aperture + (size_t)s_base * N, S_t, N, SP_t, NB_t, n_chunks, positions, 1.0 / lambda_m, bblk, posl, posf, stepf, lin, max_bits, hamf};
        sar_tc_prep_kernel<<<dim3((unsigned)n_chunks, kTcPrepSplit), 256, 0, st>>>(p);
        RB_CUDA_OK(ctx, cudaGetLastError());

        SarTcArgs a{bblk, posl, posf, stepf, (ctx->sar_dbg & 32) ? nullptr : lin, dirs, npad, n_chunks, seg_chunks, G, S_t, SP_t, NB_t, g_begin, g_end,
                    beams + (size_t)s_base * width, ksplit, segs_per_split, n_part > 1 ? 1 : 0, (size_t)S * width, (ctx->sar_dbg & 64) ? 1 : 0};
        const int tpr = ctx->sar_tpr > 0 ? ctx->sar_tpr : (mt == 2 ? 1 : 2);      // measured: 0.443 / 0.450 ms (2 tiles), 0.550 / 0.503 ms (1 tile)
        const int unr = ctx->sar_dbg & 15, cs = !(ctx->sar_dbg & 16);      // RADAR_B200_SAR_DBG: unroll of the element loop (0 = default), +16 = sincospif
#define RB_SAR_TC_LAUNCH(MT_, TPR_, UNR_, CS_)                                                                               \
    do {                                                                                                                     \
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(sar_tc_kernel<MT_, TPR_, UNR_, CS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        sar_tc_kernel<MT_, TPR_, UNR_, CS_><<<dim3(grid, (unsigned)ksplit), 128 * MT_ * TPR_ + 64, smem, st>>>(a);                                   \
    } while (0)
#define RB_SAR_TC_VARIANTS(MT_, TPR_)                                                                                        \
    do {                                                                                                                     \
        if (cs) { if (unr == 1) RB_SAR_TC_LAUNCH(MT_, TPR_, 1, true); else if (unr == 2) RB_SAR_TC_LAUNCH(MT_, TPR_, 2, true); else RB_SAR_TC_LAUNCH(MT_, TPR_, 16, true); } \
        else { if (unr == 1) RB_SAR_TC_LAUNCH(MT_, TPR_, 1, false); else if (unr == 2) RB_SAR_TC_LAUNCH(MT_, TPR_, 2, false); else RB_SAR_TC_LAUNCH(MT_, TPR_, 16, false); } \
    } while (0)
        if (mt == 2 && tpr == 2) RB_SAR_TC_VARIANTS(2, 2);
        else if (mt == 2) RB_SAR_TC_VARIANTS(2, 1);
        else if (tpr == 2) RB_SAR_TC_VARIANTS(1, 2);
        else RB_SAR_TC_VARIANTS(1, 1);
#undef RB_SAR_TC_VARIANTS
#undef RB_SAR_TC_LAUNCH
        RB_CUDA_OK(ctx, cudaGetLastError());
    }

    SarFftArgs f{beams, n_part, S, width, ctx->fftS.hann, ctx->fftS.tw, ctx->fftS.pos, out, ctx->fftS.plan, make_fastdiv(kTcFftDirs), max_bits, {nullptr}, ctx->n_image_mirrors, ctx->image_row_stride, ctx->image_col0};
    for (int m = 0; m < ctx->n_image_mirrors; ++m) f.mirrors[m] = ctx->image_mirrors[m];
    if (S == 90 && !(ctx->sar_dbg & 128)) {            // RADAR_B200_SAR_DBG & 128: the shared-memory transform for every S (comparison)
        sar_range_fft_reg_kernel<45><<<(unsigned)((width + kTcFftRegDirs - 1) / kTcFftRegDirs), 2 * kTcFftRegDirs, 0, st>>>(f);
        RB_CUDA_OK(ctx, cudaGetLastError());
        return RADAR_OK;
    }
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(sar_range_fft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
    sar_range_fft_kernel<<<(unsigned)((width + kTcFftDirs - 1) / kTcFftDirs), 256, fsmem, st>>>(f);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb
