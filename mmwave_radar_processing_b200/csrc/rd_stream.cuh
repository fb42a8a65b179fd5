// Streaming range-Doppler kernel (K1+K2), third generation: Doppler transform FIRST, range transform LAST.
//
// Reference arithmetic: processors/virtual_array_reformater.py:44-65 (TDM regroup), processors/range_doppler_resp.py:49-110
// (Hann(S) x Hann(L), fft2, Doppler fftshift), range_doppler_detection/range_doppler_detector.py:62-80 (|RD[0]|).
//
// Why this order.  ncu on the previous kernel (profiles/r1c_ncu_rd_spec_1843_RadVel.txt) showed the L1/shared-memory data
// pipe 93.6 % busy (one 128-byte wavefront per clock per SM) with the FMA pipe half idle: the kernel is bound by
// shared-memory WAVEFRONTS, then by issue slots, not by HBM.  So the layout work here removes wavefronts:
//   * the int16 words of a chirp row are contiguous along the sample axis s, and so are the rows of neighbouring RX
//     antennas.  A thread that owns one (antenna, s) column reads its L (or L/DL2) words straight from global memory
//     with lanes along s - coalesced, no shared-memory staging of the raw words at all;
//   * the Doppler transform of that column runs in registers (whole length when L <= 32: prime-factor butterflies,
//     no twiddles; otherwise two radix stages), writing the tile as tile[s][c] at an odd pitch (conflict free);
//   * the range transform then runs with lanes along the Doppler bin c, which is the contiguous axis of the tile
//     AND of the output cube: both shared-memory passes are conflict free and the last radix stage stores whole
//     256-byte runs of rd_out[k][d] (the fftshift is a rotation of the lane -> d map, two contiguous pieces).
//   * twiddles and the Doppler window come from the kernel parameter block (constant bank, LDC), not shared memory.
// Per 100x30 tile this is ~1 200 wavefronts (stores included) against ~3 200 before.
//
// CTAs are persistent: gridDim = SMs x resident CTAs, each loops over (frame, antenna group) items, so index
// arithmetic that depends only on the thread is hoisted out of the item loop by the compiler.
//
// Everything that touches data is __host__ __device__ so tests/native/host_rd_stream_check.cu can run the very same
// phases on the CPU (threads emulated by a loop, __syncthreads by phase boundaries).
#pragma once
#include <cstdint>
#include "fft_radix.cuh"

namespace rb {

template <int S_, int L_, int CR_, int RS1_, int RS2_, int DL1_, int DL2_, int G_, int NT_, int MINB_, int DF_ = 1>
struct RdStreamShape {
    static constexpr int S = S_, L = L_, CR = CR_, RS1 = RS1_, RS2 = RS2_, DL1 = DL1_, DL2 = DL2_, G = G_, NT = NT_, MINB = MINB_, DF = DF_;
    static constexpr int Lp = L | 1;             // odd pitch (complex elements) of a tile row
    static constexpr int TILE = S * Lp;          // elements per antenna tile
    static constexpr int COLS = G * S;           // (antenna, sample) columns handled by one CTA
    static constexpr int ROW = CR * S;           // words between consecutive loops of one TX slot (CR = chirps_per_loop * num_rx)
    static constexpr int NW = NT / 32;
    // phase A register file of raw words: whole column (DL2 == 1) or one radix-DL1 group per round of stage 0
    static constexpr int A_TASKS = DL2 * COLS;
    static constexpr int A_ROUNDS = (A_TASKS + NT - 1) / NT;
    static constexpr int RAW_WORDS = A_ROUNDS * DL1;
    // phase B: when a Doppler row fits one warp (L <= 32) every warp task is one (antenna, n2) row - no bank conflicts at row
    // boundaries and warp-uniform twiddles; longer rows are flattened over the threads
    static constexpr bool ROW_PER_WARP = L <= 32 && L >= 24;
    // tile double buffer (phase A of item i+1 overlaps the stores of item i) when it does not cost residency
    static constexpr int NBUF = (size_t)2 * G * TILE * sizeof(cf) * MINB <= 220 * 1024 ? 2 : 1;
    // DF > 1: the Doppler length is DL1*DF with DL1 and DF coprime (prime-factor map, no twiddles).  Phase A runs the DL1-point
    // transforms of the DF residue classes l = (DF*n1 + DL1*h) mod L in registers (DL1 words in flight per thread instead of
    // L: half the registers, twice the resident warps) and the radix-DF combine is fused into the first range stage, which
    // reads the DF partial columns it needs anyway.  Tile position p = h*DL1 + k1 holds Doppler bin (E1*k1 + E2*h) mod L.
    static constexpr int PFA_E1 = DF * pfa_inv(DF % DL1, DL1), PFA_E2 = DL1 * pfa_inv(DL1 % (DF > 1 ? DF : 2), DF > 1 ? DF : 2);
    static_assert(RS1 * RS2 == S && DL1 * DL2 * DF == L, "radix split must multiply to the length");
    static_assert(DF == 1 || (DL2 == 1 && NT >= COLS && 32 / DL1 >= 1), "fused Doppler combine: one thread per column");
    static_assert(NT % 32 == 0, "whole warps");
    static constexpr size_t smem_bytes() { return (size_t)NBUF * G * TILE * sizeof(cf) + (size_t)S * sizeof(float); }
};

template <class C>
struct RdStreamParams {
    const uint32_t* iq;        // int16 pairs as 32-bit words: [frame][chirp][rx][sample]
    cf* rd_out;                // [frame][V][S][L] or null
    float* mag_out;            // see radar_range_doppler
    long long n_items;         // frames * items_per_frame
    int items_per_frame;       // V / G
    int R, cpl, V;
    int mag_antenna;
    float winL[C::L];
    float winS[C::S];
    cf twS[C::RS2 * C::RS1];   // [n2][k1] = W_S^(n2*k1)
    cf twD[C::DL2 * C::DL1];   // [n2][k1] = W_L^(n2*k1) (two-stage Doppler only)
};

// int16 pair -> complex float, exactly, without the conversion unit: 0x4B000000 | u is the float 2^23 + u for u < 2^16, and
// u = x + 32768 for the sign-flipped half word, so subtracting 2^23 + 32768 leaves x (every step is exact).
RB_HD cf iq_word(uint32_t w) {
#ifdef __CUDA_ARCH__
    // (w & 0xffff) ^ 0x4B008000 flips the sign bit of the low half and sets the exponent in one LOP3 (the fields do not overlap)
    const float lo = __uint_as_float((w & 0xffffu) ^ 0x4B008000u);
    const float hi = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7632) ^ 0x8000u);
    return cadd(cmake(lo, hi), cmake(-8421376.0f, -8421376.0f));
#else
    return cmake((float)(short)(w & 0xffffu), (float)(short)(w >> 16));
#endif
}

#ifdef __CUDA_ARCH__
#define RB_LDG32(p) __ldg(p)
#else
#define RB_LDG32(p) (*(p))
#endif

// ---- phase A loads: the raw words of this thread's columns, issued one item ahead of their use --------------------------
// stage-0 task u = n2*COLS + t (t = g*S + s) reads words (n2 + DL2*i, t), i < DL1, at src[l*ROW + t].
template <class C>
RB_HD void rd_stream_load_raw(int tid, const uint32_t* src, uint32_t (&w)[C::RAW_WORDS]) {
    constexpr int DL1 = C::DL1, DL2 = C::DL2;
#pragma unroll
    for (int j = 0; j < C::A_ROUNDS; ++j) {
        const int u = tid + j * C::NT;
        if (u < C::A_TASKS) {
            const int n2 = u / C::COLS, t = u - n2 * C::COLS;
#pragma unroll
            for (int i = 0; i < DL1; ++i) w[j * DL1 + i] = RB_LDG32(src + (n2 + DL2 * i) * C::ROW + t);
        }
    }
}

// ---- phase A stage 0: window, radix DL1 over l = n2 + DL2*i, twiddle, store column k1 + DL1*n2 of row t ------------------
// (DL2 == 1: the whole Doppler transform; the range window is applied here because no second stage follows)
template <class C>
RB_HD void rd_stream_doppler_s0(const RdStreamParams<C>& a, int tid, const uint32_t (&w)[C::RAW_WORDS], cf* tile,
                                const float* winS_s) {
    constexpr int DL1 = C::DL1, DL2 = C::DL2;
#pragma unroll
    for (int j = 0; j < C::A_ROUNDS; ++j) {
        const int u = tid + j * C::NT;
        if (u < C::A_TASKS) {
            const int n2 = u / C::COLS, t = u - n2 * C::COLS;
            cf v[DL1];
#pragma unroll
            for (int i = 0; i < DL1; ++i) v[i] = cscale(iq_word(w[j * DL1 + i]), a.winL[n2 + DL2 * i]);
            Dft<DL1>::run(v);
            cf* dst = tile + t * C::Lp + DL1 * n2;
            if (DL2 > 1) {
#pragma unroll
                for (int k = 1; k < DL1; ++k) v[k] = cmul(v[k], a.twD[n2 * DL1 + k]);
#pragma unroll
                for (int k = 0; k < DL1; ++k) dst[k] = v[k];
            } else {
                const float ws = winS_s[t % C::S];
#pragma unroll
                for (int k = 0; k < DL1; ++k) dst[k] = cscale(v[k], ws);
            }
        }
    }
}
// ---- phase A stage 1 (two-stage Doppler only): radix DL2 over n2, range window ---------------------------------------------
template <class C>
RB_HD void rd_stream_doppler_s1(int tid, cf* tile, const float* winS_s) {
    constexpr int DL1 = C::DL1, DL2 = C::DL2;
#pragma unroll
    for (int u = tid; u < DL1 * C::COLS; u += C::NT) {
        const int k1 = u / C::COLS, t = u - k1 * C::COLS;
        cf* p = tile + t * C::Lp + k1;
        cf v[DL2];
#pragma unroll
        for (int i = 0; i < DL2; ++i) v[i] = p[DL1 * i];
        Dft<DL2>::run(v);
        const float ws = winS_s[t % C::S];
#pragma unroll
        for (int k = 0; k < DL2; ++k) p[DL1 * k] = cscale(v[k], ws);
    }
}

// ---- fused mode (DF > 1): phase A works on the residue class H of the Doppler axis ------------------------------------
template <class C, int H>
RB_HD void rd_stream_load_half(int tid, const uint32_t* src, uint32_t (&w)[C::DL1]) {
    if (tid < C::COLS) {
#pragma unroll
        for (int n1 = 0; n1 < C::DL1; ++n1) w[n1] = RB_LDG32(src + ((C::DF * n1 + C::DL1 * H) % C::L) * C::ROW + tid);
    }
}
template <class C, int H>
RB_HD void rd_stream_prefetch_half(int tid, const uint32_t* src) {
#ifdef __CUDA_ARCH__
    if (tid < C::COLS && (tid & 7) == 0) {          // one request per 32-byte sector
#pragma unroll
        for (int n1 = 0; n1 < C::DL1; ++n1)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(src + ((C::DF * n1 + C::DL1 * H) % C::L) * C::ROW + tid));
    }
#endif
}
template <class C, int H>
RB_HD void rd_stream_convert_half(const RdStreamParams<C>& a, int tid, const uint32_t (&w)[C::DL1], cf (&v)[C::DL1]) {
    if (tid < C::COLS) {
#pragma unroll
        for (int n1 = 0; n1 < C::DL1; ++n1) v[n1] = cscale(iq_word(w[n1]), a.winL[(C::DF * n1 + C::DL1 * H) % C::L]);
    }
}
template <class C, int H>
RB_HD void rd_stream_transform_half(int tid, cf (&v)[C::DL1], cf* tile, const float* winS_s) {
    constexpr int DL1 = C::DL1;
    if (tid < C::COLS) {
        Dft<DL1>::run(v);
        const float ws = winS_s[tid % C::S];
        cf* dst = tile + tid * C::Lp + H * DL1;
#pragma unroll
        for (int k = 0; k < DL1; ++k) dst[k] = cscale(v[k], ws);
    }
}
template <class C, int H>
RB_HD void rd_stream_doppler_half(const RdStreamParams<C>& a, int tid, const uint32_t (&w)[C::DL1], cf* tile, const float* winS_s) {
    cf v[C::DL1];
    rd_stream_convert_half<C, H>(a, tid, w, v);
    rd_stream_transform_half<C, H>(tid, v, tile, winS_s);
}
// first range stage with the radix-2 Doppler combine in front: a thread owns partial columns k1 and DL1 + k1 of the rows
// n2 + RS2*i; DL1 lanes per (antenna, n2) row group, 32/DL1 row groups per warp.
template <class C>
RB_HD void rd_stream_range_s0_fused(const RdStreamParams<C>& a, int tid, cf* tile) {
    static_assert(C::DF <= 2, "radix-2 combine");
    constexpr int RS1 = C::RS1, RS2 = C::RS2, Lp = C::Lp, DL1 = C::DL1, RPW = 32 / DL1;
    const int warp = tid >> 5, lane = tid & 31;
    const int sub = lane / DL1, k1 = lane - sub * DL1;
    if (sub >= RPW) return;
    for (int q = warp * RPW + sub; q < C::G * RS2; q += C::NW * RPW) {
        const int g = q / RS2, n2 = q - g * RS2;
        cf* p = tile + g * C::TILE + n2 * Lp + k1;
        cf x[RS1], y[RS1];
#pragma unroll
        for (int i = 0; i < RS1; ++i) {
            const cf e = p[i * RS2 * Lp], o = p[i * RS2 * Lp + DL1];
            x[i] = cadd(e, o);
            y[i] = csub(e, o);
        }
        Dft<RS1>::run(x);
#pragma unroll
        for (int k = 1; k < RS1; ++k) x[k] = cmul(x[k], a.twS[n2 * RS1 + k]);
#pragma unroll
        for (int k = 0; k < RS1; ++k) p[k * RS2 * Lp] = x[k];
        Dft<RS1>::run(y);
#pragma unroll
        for (int k = 1; k < RS1; ++k) y[k] = cmul(y[k], a.twS[n2 * RS1 + k]);
#pragma unroll
        for (int k = 0; k < RS1; ++k) p[k * RS2 * Lp + DL1] = y[k];
    }
}

// Phase B task enumeration: calls f(q, c) for every (row group q, Doppler column c) this thread owns, q < n_q.
template <class C, class F>
RB_HD void rd_stream_for_rows(int tid, int n_q, F f) {
    if (C::ROW_PER_WARP) {
        const int warp = tid >> 5, lane = tid & 31;
        if (lane < C::L)
            for (int q = warp; q < n_q; q += C::NW) f(q, lane);
    } else {
        // flattened over PADDED rows (pitch Lp, one idle lane per row): consecutive lanes touch consecutive tile elements
        // even across a row boundary, so no shared-memory access of phase B has a bank conflict
        for (int u = tid; u < n_q * C::Lp; u += C::NT) {
            const int q = u / C::Lp, c = u - q * C::Lp;
            if (c < C::L) f(q, c);
        }
    }
}

// ---- phase B stage 0: radix RS1 over s = n2 + RS2*i, lanes along the Doppler column c --------------------------------
template <class C>
RB_HD void rd_stream_range_s0(const RdStreamParams<C>& a, int tid, cf* tile) {
    constexpr int RS1 = C::RS1, RS2 = C::RS2, Lp = C::Lp;
    rd_stream_for_rows<C>(tid, C::G * RS2, [&](int q, int c) {
        const int g = q / RS2, n2 = q - g * RS2;
        cf* p = tile + g * C::TILE + n2 * Lp + c;
        cf v[RS1];
#pragma unroll
        for (int i = 0; i < RS1; ++i) v[i] = p[i * RS2 * Lp];
        Dft<RS1>::run(v);
        if (RS2 > 1) {
#pragma unroll
            for (int k = 1; k < RS1; ++k) v[k] = cmul(v[k], a.twS[n2 * RS1 + k]);
        }
#pragma unroll
        for (int k = 0; k < RS1; ++k) p[k * RS2 * Lp] = v[k];
    });
}

// ---- phase B stage 1: radix RS2 over n2, results leave the registers straight to rd_out[k][d] -------------------------
// out_base: element offset of antenna 0 of this item in the [.., S, L] cubes; mag_base likewise for the magnitude map;
// mag_g: antenna of this item whose magnitude is wanted (-1: all of them, layout like rd_out).
template <class C>
RB_HD void rd_stream_range_s1(int tid, const cf* tile, cf* rd_out, long long out_base, float* mag_out,
                              long long mag_base, int mag_g) {
    constexpr int RS1 = C::RS1, RS2 = C::RS2, L = C::L, Lp = C::Lp, S = C::S;
    rd_stream_for_rows<C>(tid, C::G * RS1, [&](int q, int c) {
        const int g = q / RS1, k1 = q - g * RS1;
        const cf* p = tile + g * C::TILE + (RS2 * k1) * Lp + c;
        cf v[RS2];
#pragma unroll
        for (int i = 0; i < RS2; ++i) v[i] = p[i * Lp];
        Dft<RS2>::run(v);
        int kd = c;                              // Doppler frequency bin held at tile position c
        if (C::DF > 1) { const int h = c / C::DL1; kd = (C::PFA_E1 * (c - h * C::DL1) + C::PFA_E2 * h) % L; }
        int d = kd + L / 2;                      // fftshift: frequency bin kd lands at d = (kd + L//2) mod L
        if (d >= L) d -= L;
        const int o = (g * S + k1) * L + d;      // range bin k = k1 + RS1*k2
        if (rd_out) {
#pragma unroll
            for (int k = 0; k < RS2; ++k) rd_out[out_base + o + k * (RS1 * L)] = v[k];
        }
        if (mag_out && (mag_g < 0 || mag_g == g)) {
            const int om = mag_g < 0 ? o : o - g * S * L;
#pragma unroll
            for (int k = 0; k < RS2; ++k)
                mag_out[mag_base + om + k * (RS1 * L)] = sqrtf(v[k].x * v[k].x + v[k].y * v[k].y);
        }
    });
}

}  // namespace rb
