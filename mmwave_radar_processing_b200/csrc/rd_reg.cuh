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
// (the prime-factor lengths 35, 40, 45, 63, 70 live in fft_radix.cuh)
// 25 = 5 x 5 by Cooley-Tukey in registers: n = 5*n1 + n2, k = k1 + 5*k2, twiddles W25^(n2*k1) as literals
template <> struct Dft<25> {
    RB_HD static void run(cf* v) {
        cf a[5][5];
#pragma unroll
        for (int n2 = 0; n2 < 5; ++n2) {
#pragma unroll
            for (int n1 = 0; n1 < 5; ++n1) a[n2][n1] = v[5 * n1 + n2];
            Dft<5>::run(a[n2]);
        }
        a[1][1] = cmul(a[1][1], cmake(0.96858316112863107605f, -0.24868988716485479484f));   // W25^1
        a[1][2] = cmul(a[1][2], cmake(0.87630668004386358394f, -0.48175367410171532345f));   // W25^2
        a[1][3] = cmul(a[1][3], cmake(0.72896862742141155245f, -0.68454710592868861507f));   // W25^3
        a[1][4] = cmul(a[1][4], cmake(0.53582679497899654564f, -0.84432792550201507531f));   // W25^4
        a[2][1] = cmul(a[2][1], cmake(0.87630668004386358394f, -0.48175367410171532345f));   // W25^2
        a[2][2] = cmul(a[2][2], cmake(0.53582679497899654564f, -0.84432792550201507531f));   // W25^4
        a[2][3] = cmul(a[2][3], cmake(0.06279051952931352654f, -0.99802672842827155897f));   // W25^6
        a[2][4] = cmul(a[2][4], cmake(-0.42577929156507271502f, -0.90482705246601946580f));   // W25^8
        a[3][1] = cmul(a[3][1], cmake(0.72896862742141155245f, -0.68454710592868861507f));   // W25^3
        a[3][2] = cmul(a[3][2], cmake(0.06279051952931352654f, -0.99802672842827155897f));   // W25^6
        a[3][3] = cmul(a[3][3], cmake(-0.63742398974868974548f, -0.77051324277578925326f));   // W25^9
        a[3][4] = cmul(a[3][4], cmake(-0.99211470131447776488f, -0.12533323356430453588f));   // W25^12
        a[4][1] = cmul(a[4][1], cmake(0.53582679497899654564f, -0.84432792550201507531f));   // W25^4
        a[4][2] = cmul(a[4][2], cmake(-0.42577929156507271502f, -0.90482705246601946580f));   // W25^8
        a[4][3] = cmul(a[4][3], cmake(-0.99211470131447776488f, -0.12533323356430453588f));   // W25^12
        a[4][4] = cmul(a[4][4], cmake(-0.63742398974868952344f, 0.77051324277578936428f));   // W25^16
#pragma unroll
        for (int k1 = 0; k1 < 5; ++k1) {
            cf y[5] = {a[0][k1], a[1][k1], a[2][k1], a[3][k1], a[4][k1]};
            Dft<5>::run(y);
#pragma unroll
            for (int k2 = 0; k2 < 5; ++k2) v[k1 + 5 * k2] = y[k2];
        }
    }
};
template <> struct Dft<50> : DftPfa<2, 25> {};

// DF = 2: the Doppler column is split over two threads by the prime-factor map L = 2 * (L/2), L/2 odd: thread h transforms
// the L/2 inputs l = (h*L/2 + 2m) mod L and X[k] = Y0[k mod L/2] + (-1)^k Y1[k mod L/2] is formed by phase B as it loads
// (twice the A warps with half the registers and half the straight-line code each; the code of all roles then fits the
// 32 KB instruction cache level, which the undivided version overflowed: 44 % of its phase-A cycles were no-instruction stalls).
// BF = 2 or 4: the range column is split over BF threads by the prime-factor map S = BF * (S/BF), S/BF odd: thread class k_BF
// reads all S tile values, folds them with the trivial radix-BF butterfly of ITS output class into S/BF values, transforms
// those and stores the bins k = k_BF (mod BF) - no twiddles, BF times the B warps with 1/BF of the registers (S = 100 does
// not fit one thread's registers at all).  The class is a = k_BF >> 1 (run time: a sign and a base pointer) and, for BF = 4,
// ODD = k_BF & 1 (two copies of the code).
template <int S_, int L_, int CR_, int G_, int DF_ = 1, int NSTAGE_ = 3, int NTILE_ = 1, int BF_ = 1>
struct RdRegShape {
    static constexpr int S = S_, L = L_, CR = CR_, G = G_, DF = DF_, NTILE = NTILE_, BF = BF_;
    static constexpr int SH = S / BF;                        // points per phase-B thread
    static constexpr int E_SH = BF * pfa_inv(BF % (SH > 1 ? SH : 2), SH > 1 ? SH : 2);     // 1 (mod SH), 0 (mod BF)
    static constexpr int E_BF = BF == 1 ? 0 : SH * pfa_inv(SH % (BF > 1 ? BF : 2), BF > 1 ? BF : 2);   // 1 (mod BF), 0 (mod SH)
    static_assert(BF == 1 || ((BF == 2 || BF == 4) && SH % 2 == 1 && SH * BF == S), "the range split needs coprime factors");
    static constexpr int LH = L / DF;                        // points per phase-A thread
    // DF = 2 with an even half (L = 100): no coprime split exists, so the halves are the even / odd samples (plain decimation
    // in time) and phase B folds them with the column's own twiddle: X[k] = E[k mod LH] + W_L^k O[k mod LH] (a per-thread
    // constant: four FMAs per value instead of one packed FMA)
    static constexpr bool TW = DF == 2 && LH % 2 == 0;
    static constexpr int HP = (LH + 1) / 2 * 2;              // tile offset between the halves (16-byte aligned)
    static constexpr int ROW = CR * S;                       // words between consecutive loops of one TX slot
    static constexpr int COLS_A = G * S, COLS_B = G * L;
    static constexpr int NA1 = (COLS_A + 31) / 32;           // phase-A warps per residue class
    static constexpr int NB1 = (COLS_B + 31) / 32;           // phase-B warps per output class
    static constexpr int NA = NA1 * DF, NB = NB1 * BF;       // warps per role
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
    static_assert(DF == 1 || DF == 2, "Doppler column: whole or in two halves");
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
    float2 winL[C::L];         // (w, w) pairs, [h][m] = Doppler window of input m of class h (see rd_reg_doppler_input): one LDC.64 per packed multiply
    float winS[C::S];
    cf twL[C::TW ? C::L : 1];  // W_L^k (twiddle mode of the Doppler split only)
};
// input index of sample m of residue class h
template <class C>
RB_HD constexpr int rd_reg_doppler_input(int h, int m) { return C::TW ? C::DF * m + h : (h * C::LH + C::DF * m) % C::L; }

// Doppler beam sharpening on top of the same phases (kernels_rd_reg.cu: rd_reg_dbs_kernel): phase B beamforms the Doppler
// transformed samples of the frame's antennas per output angle in shared memory and runs one range transform per angle.
struct RdRegDbs {
    const int32_t* vel_bin;     // [n_out] Doppler column (shifted axis) of output angle i
    const int32_t* ang_bin;     // [n_out] coarse angle bin of output angle i
    const int32_t* ant_list;    // [N] virtual antennas of the subset, ascending, or null (all CR antennas)
    const float* ant_weight;    // [N] or null
    const cf* twA;              // [A] W_A^j
    float* out;                 // [frames][S][n_out]
    int n_out, N, A;
    unsigned n_frames;
};

// int16 pair -> complex float on the conversion unit (two I2F.S16, the second with the .H1 selector): half the issue slots
// of the LOP3/PRMT/LOP3/FADD2 form of iq_word and nothing on the FMA / ALU pipes, which phase A keeps busy; the conversion
// unit is otherwise idle in this kernel (70 words per item and thread = 140 quarter-rate instructions per A warp and item)
RB_HD cf iq_word_cvt(uint32_t w) {
#ifdef __CUDA_ARCH__
    short lo, hi;
    asm("mov.b32 {%0,%1}, %2;" : "=h"(lo), "=h"(hi) : "r"(w));
    float a, b;
    asm("cvt.rn.f32.s16 %0, %1;" : "=f"(a) : "h"(lo));
    asm("cvt.rn.f32.s16 %0, %1;" : "=f"(b) : "h"(hi));
    return cmake(a, b);
#else
    return cmake((float)(short)(w & 0xffffu), (float)(short)(w >> 16));
#endif
}

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
    // twiddle mode: class h starts at row h and never wraps
    const uint32_t* lo = raw + col_shift + h * ((C::TW ? 1 : C::LH) * C::RAW_PITCH_W);
    const uint32_t* hi = lo - C::L * C::RAW_PITCH_W;
    const float2* win = a.winL + h * C::LH;
#pragma unroll
    for (int m = 0; m < C::LH; ++m) {
        // class 1 wraps when DF*m + LH >= L; class 0 never does (h = 0 makes lo and hi + L rows the same address)
        const bool wraps = !C::TW && C::DF * m + C::LH >= C::L;
        const uint32_t w = wraps ? (h ? hi : lo)[C::DF * m * C::RAW_PITCH_W] : lo[C::DF * m * C::RAW_PITCH_W];
        // the conversion unit serves one A warp per scheduler well (6843: 0.547 -> 0.519 ms); with two they queue (1843: 0.257 -> 0.262)
        const cf x = C::NA <= 4 ? iq_word_cvt(w) : iq_word(w);
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
// Twiddle mode of the Doppler split (L = 100: even / odd samples): X[k] = E[k mod LH] + W_L^k O[k mod LH] and
// W_L^(m + LH) = -W_L^m, so the odd class multiplies its LH partial bins by W_L^m HERE, once, and phase B only adds or
// subtracts (bins k < LH / k >= LH).  Folding in phase B instead cost four FMAs per value, for every range sample of every
// column and again in every output class of the range split: four times the multiplies.
template <class C>
RB_HD void rd_reg_doppler_twiddle(const RdRegParams<C>& a, int h, cf (&v)[C::LH]) {
    if (C::TW && h) {
#pragma unroll
        for (int m = 0; m < C::LH; ++m) v[m] = cmul(v[m], a.twL[m]);
    }
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

// ---- phase B: range transform of column (antenna g, Doppler position d), output class (a, ODD) ------------------------------
// DF == 2: position d holds frequency bin k = (d + L - L/2) mod L = Y0[k mod LH] + sign Y1[k mod LH]; coprime halves:
// sign = (-1)^k; twiddle mode (Y1 already carries W_L^(k mod LH), rd_reg_doppler_twiddle): sign = -1 for k >= LH
template <class C>
struct RdRegColumn {                      // one tile column as phase B sees it: value(s) applies the Doppler combine
    const cf* p;
    float sign;
    RB_HD RdRegColumn(const cf* tile, int g, int d, const cf*) {
        sign = 1.0f;
        if (C::DF == 1) p = tile + (size_t)g * C::S * C::P + d;
        else {
            const int k = (d + C::L - C::L / 2) % C::L;
            sign = (C::TW ? k >= C::LH : (k & 1)) ? -1.0f : 1.0f;
            p = tile + (size_t)g * C::S * C::P + k % C::LH;
        }
    }
    RB_HD cf value(int s) const {
        if (C::DF == 1) return p[s * C::P];
        return cfma(sign, p[s * C::P + C::HP], p[s * C::P]);
    }
};
// sigma = +1 for a = 0, -1 for a = 1.  Input m of class k_BF: sum_q x[(SH*q + BF*m) mod S] * W_BF^(q*k_BF)
template <class C, int ODD>
RB_HD void rd_reg_range_load(const RdRegParams<C>& a, const cf* tile, int g, int d, float sigma, cf (&v)[C::SH]) {
    const RdRegColumn<C> col(tile, g, d, a.twL);
#pragma unroll
    for (int m = 0; m < C::SH; ++m) {
        if (C::BF == 1) v[m] = col.value(m);
        else if (C::BF == 2) v[m] = cfma(sigma, col.value((C::SH + 2 * m) % C::S), col.value((2 * m) % C::S));
        else {
            const cf x0 = col.value((4 * m) % C::S), x1 = col.value((C::SH + 4 * m) % C::S);
            const cf x2 = col.value((2 * C::SH + 4 * m) % C::S), x3 = col.value((3 * C::SH + 4 * m) % C::S);
            // k_4 = 0, 2: (x0 + x2) +- (x1 + x3);  k_4 = 1, 3: (x0 - x2) -+ i (x1 - x3)
            if (ODD == 0) v[m] = cfma(sigma, cadd(x1, x3), cadd(x0, x2));
            else v[m] = cfma(sigma, csub_mi(x1, x3), csub(x0, x2));
        }
    }
}
// the same input combination for a plain column p[s * pitch] (rd_reg_dbs_kernel: the beamformed samples of one output angle)
template <class C, int ODD>
RB_HD void rd_reg_range_load_plain(const cf* p, int pitch, float sigma, cf (&v)[C::SH]) {
#pragma unroll
    for (int m = 0; m < C::SH; ++m) {
        if (C::BF == 1) v[m] = p[m * pitch];
        else if (C::BF == 2) v[m] = cfma(sigma, p[((C::SH + 2 * m) % C::S) * pitch], p[((2 * m) % C::S) * pitch]);
        else {
            const cf x0 = p[((4 * m) % C::S) * pitch], x1 = p[((C::SH + 4 * m) % C::S) * pitch];
            const cf x2 = p[((2 * C::SH + 4 * m) % C::S) * pitch], x3 = p[((3 * C::SH + 4 * m) % C::S) * pitch];
            if (ODD == 0) v[m] = cfma(sigma, cadd(x1, x3), cadd(x0, x2));
            else v[m] = cfma(sigma, csub_mi(x1, x3), csub(x0, x2));
        }
    }
}
// Output kh of the class is bin k = (a*S/2 + ODD*E_BF + E_SH*kh) mod S: r = (ODD*E_BF + E_SH*kh) mod S is a compile-time
// constant and the a*S/2 term with its wrap is folded into two base pointers (lo = base + a*S/2 rows for r < S/2, hi = base
// - a*S/2 rows for r >= S/2; a = 0: both are the base).  out_*: &rd_out[frame][ant][0][d] +- a*(S/2)*L (or null); mag likewise.
template <class C, int ODD>
RB_HD void rd_reg_range_store(cf (&v)[C::SH], cf* out_lo, cf* out_hi, float* mag_lo, float* mag_hi) {
    Dft<C::SH>::run(v);
    if (out_lo) {
#pragma unroll
        for (int kh = 0; kh < C::SH; ++kh) {
            const int r = (ODD * C::E_BF + C::E_SH * kh) % C::S;
            (C::BF > 1 && r >= C::S / 2 ? out_hi : out_lo)[r * C::L] = v[kh];
        }
    }
    if (mag_lo) {
#pragma unroll
        for (int kh = 0; kh < C::SH; ++kh) {
            const int r = (ODD * C::E_BF + C::E_SH * kh) % C::S;
            (C::BF > 1 && r >= C::S / 2 ? mag_hi : mag_lo)[r * C::L] = rd_reg_sqrt(v[kh].x * v[kh].x + v[kh].y * v[kh].y);
        }
    }
}

}  // namespace rb
