// Register-resident range-Doppler phases (K1+K2), fourth generation: every 1-D transform lives in ONE thread's registers.
//
// Reference arithmetic: processors/virtual_array_reformater.py:44-65 (TDM regroup), processors/range_doppler_resp.py:49-110
// (Hann(S) x Hann(L), fft2, Doppler fftshift), range_doppler_detection/range_doppler_detector.py:62-80 (|RD[0]|).
//
// Why.  ncu on the multi-stage shared-memory kernel (profiles/r1m_ncu_rd_stream_6843_RadVel_ods_20Hz.txt) shows it bound by
// the L1/shared data pipe (63 %: three round trips of the tile through shared memory plus the global sectors) and by issue
// slots (46 % with 24 warps per SM, 2.6 of 11 stall cycles on CTA barriers), at 37 % of DRAM.  This version
//   * moves the raw int16 rows with the TMA engine (cp.async.bulk -> shared-memory ring, completion on mbarriers): no LDG
//     instruction, no load latency in any warp, three items of prefetch;
//   * phase A: one thread per (antenna, sample) column reads its L raw words from the ring (lanes along the sample axis,
//     conflict free), converts, windows and runs the WHOLE L-point Doppler transform in registers (prime-factor map: no
//     twiddles, every index a compile-time constant), then writes the row tile[col][d] in fftshifted order with 128-bit
//     stores at a pitch that is conflict free for them;
//   * phase B: one thread per (antenna, Doppler bin) column reads its S values (lanes along d: conflict free), runs the
//     WHOLE S-point range transform in registers and stores straight to rd_out[k][d] (lanes along the contiguous axis).
// The tile makes ONE trip through shared memory (552 wavefronts per 63x70 tile against 2 176 before), the thread-instruction
// count per point drops from 62 to about 35, and the only synchronisation is producer/consumer mbarriers between the warp
// roles (A warps, B warps, one TMA warp) - no CTA-wide barrier in the steady state.
//
// Everything that touches data is __host__ __device__ so tests/native/host_rd_reg_check.cu runs the same phases on the CPU.
#pragma once
#include <cstdint>
#include "fft_radix.cuh"
#include "rd_stream.cuh"     // iq_word

namespace rb {

// ---- whole-length transforms in registers -------------------------------------------------------------------------
template <> struct Dft<63> : DftPfa<7, 9> {};
template <> struct Dft<70> : DftPfa<7, 10> {};
template <> struct Dft<40> : DftPfa<5, 8> {};
template <> struct Dft<35> : DftPfa<5, 7> {};
template <> struct Dft<45> : DftPfa<5, 9> {};

// DF = 2: the Doppler column is split over two threads by the prime-factor map L = 2 * (L/2), L/2 odd: thread h transforms
// the L/2 inputs l = (h*L/2 + 2m) mod L and X[k] = Y0[k mod L/2] + (-1)^k Y1[k mod L/2] is formed by phase B as it loads
// (twice the A warps with half the registers and half the straight-line code each; the code of all roles then fits the
// 32 KB instruction cache level, which the undivided version overflowed: 44 % of its phase-A cycles were no-instruction stalls).
template <int S_, int L_, int CR_, int G_, int DF_ = 1, int NSTAGE_ = 3, int NTILE_ = 1>
struct RdRegShape {
    static constexpr int S = S_, L = L_, CR = CR_, G = G_, DF = DF_, NTILE = NTILE_;
    static constexpr int LH = L / DF;                        // points per phase-A thread
    static constexpr int HP = (LH + 1) / 2 * 2;              // tile offset between the halves (16-byte aligned)
    static constexpr int ROW = CR * S;                       // words between consecutive loops of one TX slot
    static constexpr int COLS_A = G * S, COLS_B = G * L;
    static constexpr int NA1 = (COLS_A + 31) / 32;           // phase-A warps per residue class
    static constexpr int NA = NA1 * DF, NB = (COLS_B + 31) / 32;             // warps per role
    static constexpr int NT = 32 * (NA + NB + 1);            // + one TMA warp
    // raw ring: one stage = L rows of the item's G*S words, fetched from the 16-byte aligned address below the first word
    static constexpr int ROW_COPY = ((G * S * 4 + 12) + 15) / 16 * 16;    // bytes per bulk copy (covers any 4-byte start shift)
    static constexpr int RAW_PITCH_W = ROW_COPY / 4;
    static constexpr int RAW_STAGE_W = L * RAW_PITCH_W;
    static constexpr int NSTAGE = NSTAGE_;
    // tile[col][d]: pitch P = 2 mod 4 complex elements keeps the 128-bit stores of phase A (lanes along col) conflict free:
    // a quarter warp's 8 rows start 2P mod 32 = 4*odd words apart
    static constexpr int PMIN = DF * HP;
    static constexpr int P = PMIN % 4 == 2 ? PMIN : PMIN + 2;
    static constexpr int TILE = COLS_A * P;                  // complex elements
    static_assert(L % 2 == 0, "phase A pairs neighbouring Doppler bins into 128-bit stores");
    static_assert(DF == 1 || (DF == 2 && LH % 2 == 1), "the split needs coprime factors 2 and L/2");
    static constexpr size_t smem_bytes() {
        return (size_t)NTILE * TILE * sizeof(cf) + (size_t)NSTAGE * RAW_STAGE_W * 4 + 16 * sizeof(unsigned long long);
    }
};

template <class C>
struct RdRegParams {
    const uint32_t* iq;        // int16 pairs as 32-bit words: [frame][chirp][rx][sample]
    cf* rd_out;                // [frame][V][S][L] or null
    float* mag_out;            // see radar_range_doppler
    unsigned n_items;          // frames * items_per_frame
    int mag_antenna;
    float2 winL[C::L];         // (w, w) pairs, [h][m] = Doppler window of input l = (h*LH + DF*m) mod L: one LDC.64 per packed multiply
    float winS[C::S];
};

// sqrt to 1 ulp in one MUFU (the IEEE sqrtf costs a Newton step and a guarded slow path per cell, 20 KB of code here)
RB_HD float rd_reg_sqrt(float x) {
#ifdef __CUDA_ARCH__
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
#else
    return sqrtf(x);
#endif
}

// ---- phase A: Doppler transform of residue class h of one (antenna, sample) column ----------------------------------------
// raw: stage base (words), row l at raw + l*RAW_PITCH_W; col_shift = word offset of this column inside a row.  Input m of
// class h is row l = (h*LH + DF*m) mod L: DF*m + h*LH while that stays below L, DF*m + h*LH - L after - two base pointers keep
// every row offset a compile-time immediate for both classes (one copy of the code serves both).
template <class C>
RB_HD void rd_reg_doppler_load(const RdRegParams<C>& a, const uint32_t* raw, int col_shift, int h, cf (&v)[C::LH]) {
    const uint32_t* lo = raw + col_shift + h * (C::LH * C::RAW_PITCH_W);
    const uint32_t* hi = lo - C::L * C::RAW_PITCH_W;
    const float2* win = a.winL + h * C::LH;
#pragma unroll
    for (int m = 0; m < C::LH; ++m) {
        // class 1 wraps when DF*m + LH >= L; class 0 never does (h = 0 makes lo and hi + L rows the same address)
        const bool wraps = C::DF * m + C::LH >= C::L;
        const uint32_t w = wraps ? (h ? hi : lo)[C::DF * m * C::RAW_PITCH_W] : lo[C::DF * m * C::RAW_PITCH_W];
        const cf x = iq_word(w);
#ifdef __CUDA_ARCH__
        pk64 r;
        asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(x)), "l"(pk(cmake(win[m].x, win[m].y))));
        v[m] = unpk(r);
#else
        v[m] = cscale(x, win[m].x);
#endif
    }
}
template <class C>
RB_HD void rd_reg_doppler_finish(float ws, cf (&v)[C::LH]) {
    Dft<C::LH>::run(v);
#pragma unroll
    for (int k = 0; k < C::LH; ++k) v[k] = cscale(v[k], ws);
}
// DF == 1: tile row in fftshifted order (frequency bin k lands at d = (k + L/2) mod L); DF == 2: the class's LH partial
// bins in natural order at row + h*HP (the combine and the shift happen in phase B's addressing)
template <class C>
RB_HD void rd_reg_store_row(const cf (&v)[C::LH], cf* row, int h) {
    if (C::DF == 1) {
        float4* dst = reinterpret_cast<float4*>(row);
#pragma unroll
        for (int d = 0; d + 1 < C::L; d += 2) {
            const int k0 = (d + C::L - C::L / 2) % C::L % C::LH, k1 = (d + 1 + C::L - C::L / 2) % C::L % C::LH;
            dst[d / 2] = make_float4(v[k0].x, v[k0].y, v[k1].x, v[k1].y);
        }
    } else {
        cf* half = row + h * C::HP;
        float4* dst = reinterpret_cast<float4*>(half);
#pragma unroll
        for (int k = 0; k + 1 < C::LH; k += 2) dst[k / 2] = make_float4(v[k].x, v[k].y, v[k + 1].x, v[k + 1].y);
        if (C::LH % 2) half[C::LH - 1] = v[C::LH - 1];
    }
}

// ---- phase B: range transform of column (antenna g, Doppler position d) -----------------------------------------------
// DF == 2: position d holds frequency bin k = (d + L - L/2) mod L = Y0[k mod LH] + (-1)^k Y1[k mod LH]
template <class C>
RB_HD void rd_reg_range_load(const cf* tile, int g, int d, cf (&v)[C::S]) {
    if (C::DF == 1) {
        const cf* p = tile + (size_t)g * C::S * C::P + d;
#pragma unroll
        for (int s = 0; s < C::S; ++s) v[s] = p[s * C::P];
    } else {
        const int k = (d + C::L - C::L / 2) % C::L;
        const float sign = (k & 1) ? -1.0f : 1.0f;
        const cf* p = tile + (size_t)g * C::S * C::P + k % C::LH;
#pragma unroll
        for (int s = 0; s < C::S; ++s) v[s] = cfma(sign, p[s * C::P + C::HP], p[s * C::P]);
    }
}
// out: &rd_out[frame][v0 + g][0][d] (or null), mag: &mag_out[..][0][d] (or null)
template <class C>
RB_HD void rd_reg_range_store(cf (&v)[C::S], cf* out, float* mag) {
    Dft<C::S>::run(v);
    if (out) {
#pragma unroll
        for (int k = 0; k < C::S; ++k) out[k * C::L] = v[k];
    }
    if (mag) {
#pragma unroll
        for (int k = 0; k < C::S; ++k) mag[k * C::L] = rd_reg_sqrt(v[k].x * v[k].x + v[k].y * v[k].y);
    }
}

}  // namespace rb
