// Shape-specialised fused range-Doppler kernel (K1+K2) for the int16 ingest layout.
//
// Same arithmetic as rd_fused_kernel in kernels_rd.cu (reference: processors/virtual_array_reformater.py:44-65,
// processors/range_doppler_resp.py:49-110, range_doppler_detection/range_doppler_detector.py:62-80) but with the
// transform lengths, their radix splits and the CTA size as template parameters, so every index computation folds
// into immediates, every loop unrolls and the butterflies run on packed FFMA2/FADD2 (fft_radix.cuh).
//
// One CTA per (frame, virtual antenna).  Shared memory holds the raw int16 words of the antenna (raw[l][s], 4 B per
// sample) and the complex tile A[s][l] (odd row pitch):
//   load : cp.async of the antenna's L rows (4*S bytes each, 16/8/4-byte pieces by alignment) - no register round trip
//   R0   : radix RS1 over s straight from the raw words: exact int->float, Hann(S)*Hann(L), butterfly, twiddles
//          W_S^(n2*k1); lanes along n2 (raw reads contiguous, tile writes at the odd pitch: conflict free)
//   R1   : radix RS2 over s, in place, lanes along l; range bin k = k1 + RS1*k2 ends in row RS2*k1 + k2
//   D0   : radix RL1 over l, in place, twiddles W_L^(n2*k1); lanes along the rows (odd pitch: conflict free)
//   D1   : radix RL2 over l; lanes along the Doppler bins so the results leave the registers straight to global
//          memory as runs of RL1 consecutive bins at [k][fftshift(d)] (+ |.| for the CFAR map of one antenna)
// Shared-memory traffic is 56 B per sample (the data path of the SM, not the FMA pipe, is what this kernel fills next
// to HBM), so the layout work above is about keeping every one of those accesses conflict free.
// Two-factor splits cover every length <= 256 whose factors have a hand-written butterfly; other shapes use the
// generic kernel.
#include "radar_internal.h"
#include "rd_core.cuh"

namespace rb {

// pitch (in 32-bit words) of a raw row: multiple of the copy vector, and congruent to about RS2 modulo the 32 banks so
// the RS2 consecutive words that the lanes of neighbouring rows read land in different banks
constexpr int raw_pitch(int S, int RS2, int VEC) {
    for (int sp = S;; ++sp)
        if (sp % VEC == 0 && (sp % 32) >= (RS2 % 32) && (sp % 32) <= (RS2 % 32) + 3) return sp;
}

template <int S_, int L_, int RS1_, int RS2_, int RL1_, int RL2_, int NT_>
struct RdShape {
    static constexpr int S = S_, L = L_, RS1 = RS1_, RS2 = RS2_, RL1 = RL1_, RL2 = RL2_, NT = NT_;
    static constexpr int Lp = L | 1;
    static constexpr int VEC = S % 4 == 0 ? 4 : (S % 2 == 0 ? 2 : 1);     // words per cp.async
    static constexpr int SP = raw_pitch(S, RS2, VEC);
    static_assert(RS1 * RS2 == S && RL1 * RL2 == L, "radix split must multiply to the length");
    static constexpr size_t smem_bytes() {
        return arena_bytes((size_t)S * Lp, sizeof(cf)) + arena_bytes((size_t)L * SP, 4) + arena_bytes(S, sizeof(cf)) +
               arena_bytes(L, sizeof(cf)) + arena_bytes(S, 4) + arena_bytes(L, 4);
    }
};

struct RdSpecArgs {
    const uint32_t* iq;          // int16 pairs as words: [frame][chirp][rx][sample]
    int R, cpl;
    int n_out_ant;
    const int32_t* ant_list;
    const float* winS; const float* winL;
    const cf* twS; const cf* twL;
    cf* rd_out; float* mag_out; int mag_antenna;
};

__device__ __forceinline__ cf iq_word_to_cf(uint32_t w) {
    return cmake((float)(short)(w & 0xffffu), (float)(short)(w >> 16));
}

// One in-place radix-R decimation-in-frequency task on elements base + stride*M*i, twiddles tw[step*k].
template <int R, bool TWIDDLE>
__device__ __forceinline__ void dif_task(cf* p, int stride_m, const cf* tw, int step) {
    cf v[R];
#pragma unroll
    for (int i = 0; i < R; ++i) v[i] = p[i * stride_m];
    Dft<R>::run(v);
    if (TWIDDLE) {
#pragma unroll
        for (int k = 1; k < R; ++k) v[k] = cmul(v[k], tw[step * k]);
    }
#pragma unroll
    for (int k = 0; k < R; ++k) p[k * stride_m] = v[k];
}

template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gmem_src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    if constexpr (BYTES == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(d), "l"(gmem_src), "n"(BYTES) : "memory");
}

template <class C>
__global__ void __launch_bounds__(C::NT) rd_spec_kernel(const __grid_constant__ RdSpecArgs a) {
    constexpr int S = C::S, L = C::L, Lp = C::Lp, NT = C::NT, SP = C::SP, VEC = C::VEC;
    constexpr int RS1 = C::RS1, RS2 = C::RS2, RL1 = C::RL1, RL2 = C::RL2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SmemArena ar(smem_raw);
    cf* tile = ar.take<cf>(S * Lp);
    uint32_t* raw = ar.take<uint32_t>(L * SP);
    cf* twS = ar.take<cf>(S);
    cf* twL = ar.take<cf>(L);
    float* winS = ar.take<float>(S);
    float* winL = ar.take<float>(L);
    const int tid = threadIdx.x;

    const int frame = blockIdx.x / a.n_out_ant;
    const int vo = blockIdx.x - frame * a.n_out_ant;
    const int v = a.ant_list ? a.ant_list[vo] : vo;
    const int tx = v / a.R, rx = v - tx * a.R;
    // word index of (loop l, sample s): ((frame*C + l*cpl + tx)*R + rx)*S + s
    const long long row0 = ((long long)frame * (a.cpl * L) + tx) * a.R + rx;
    const long long row_stride_words = (long long)a.cpl * a.R * S;      // between consecutive loops
    const uint32_t* src = a.iq + row0 * S;

    // ---- raw rows -> shared memory (asynchronous), tables meanwhile ----------------------------------------------
    {
        constexpr int VPR = S / VEC;                    // copies per row
        constexpr int NV = VPR * L;
        for (int i = tid; i < NV; i += NT) {
            const int l = i / VPR, q = i - l * VPR;
            cp_async<4 * VEC>(raw + l * SP + VEC * q, src + l * row_stride_words + VEC * q);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int i = tid; i < S; i += NT) { twS[i] = a.twS[i]; winS[i] = a.winS[i]; }
    for (int i = tid; i < L; i += NT) { twL[i] = a.twL[i]; winL[i] = a.winL[i]; }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- range FFT stage 0: raw words -> window -> radix RS1 -> tile; lanes along n2 -------------------------------
    {
        constexpr int T0 = RS2 * L;
        for (int t = tid; t < T0; t += NT) {
            const int l = t / RS2, n2 = t - l * RS2;
            const uint32_t* rsrc = raw + l * SP + n2;
            const float wl = winL[l];
            cf vv[RS1];
#pragma unroll
            for (int i = 0; i < RS1; ++i) vv[i] = cscale(iq_word_to_cf(rsrc[RS2 * i]), winS[n2 + RS2 * i] * wl);
            Dft<RS1>::run(vv);
            if constexpr (RS2 > 1) {
#pragma unroll
                for (int k = 1; k < RS1; ++k) vv[k] = cmul(vv[k], twS[n2 * k]);
            }
            cf* dst = tile + n2 * Lp + l;
#pragma unroll
            for (int k = 0; k < RS1; ++k) dst[k * RS2 * Lp] = vv[k];
        }
        __syncthreads();
        if constexpr (RS2 > 1) {
            constexpr int T1 = RS1 * L;
            for (int t = tid; t < T1; t += NT) {
                const int k1 = t / L, l = t - k1 * L;
                dif_task<RS2, false>(tile + (RS2 * k1) * Lp + l, Lp, twS, 0);
            }
            __syncthreads();
        }
    }

    // ---- Doppler FFT (axis l): stage 0 lanes along the rows, last stage lanes along the output bins ------------------------------------------------------------
    if constexpr (RL2 > 1) {
        constexpr int T0 = S * RL2;
#pragma unroll
        for (int t = tid; t < T0; t += NT) {
            const int n2 = t / S, p = t - n2 * S;
            dif_task<RL1, true>(tile + p * Lp + n2, RL2, twL, n2);
        }
        __syncthreads();
    }
    {
        // last stage: radix RLAST on elements base + i; output bin d = k1 + RL1*k2 (k1 = 0 when single stage)
        constexpr int RLAST = RL2 > 1 ? RL2 : RL1;
        constexpr int NK1 = RL2 > 1 ? RL1 : 1;
        constexpr int T1 = S * NK1;
        constexpr int H = L / 2;
        const long long tile_base = ((long long)frame * a.n_out_ant + vo) * (S * L);
        cf* rd_dst = a.rd_out ? a.rd_out + tile_base : nullptr;
        float* mag_dst = nullptr;
        if (a.mag_out) {
            if (a.mag_antenna < 0) mag_dst = a.mag_out + tile_base;
            else if (a.mag_antenna == vo) mag_dst = a.mag_out + (long long)frame * (S * L);
        }
#pragma unroll
        for (int t = tid; t < T1; t += NT) {
            const int p = t / NK1, k1 = t - p * NK1;
            cf vv[RLAST];
            const cf* src_row = tile + p * Lp + RLAST * k1;
#pragma unroll
            for (int i = 0; i < RLAST; ++i) vv[i] = src_row[i];
            Dft<RLAST>::run(vv);
            // row p holds range bin k = k1r + RS1*k2r with p = RS2*k1r + k2r
            const int k1r = p / RS2, k2r = p - k1r * RS2;
            const int k = k1r + RS1 * k2r;
            const int out_row = k * L;
#pragma unroll
            for (int k2 = 0; k2 < RLAST; ++k2) {
                int pos;
                if constexpr (H % NK1 == 0) {
                    pos = k1 + NK1 * ((k2 + H / NK1) % RLAST);          // fftshift folded at compile time
                } else {
                    pos = k1 + NK1 * k2 + H;
                    if (pos >= L) pos -= L;
                }
                if (rd_dst) rd_dst[out_row + pos] = vv[k2];
                if (mag_dst) mag_dst[out_row + pos] = sqrtf(vv[k2].x * vv[k2].x + vv[k2].y * vv[k2].y);
            }
        }
    }
}

template <class C>
static int launch_spec(radar_ctx* ctx, const RdSpecArgs& a, int n_frames, cudaStream_t st) {
    constexpr size_t smem = C::smem_bytes();
    static_assert(smem <= 227 * 1024, "tile exceeds shared memory");
    if (smem > 48 * 1024)
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(rd_spec_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long grid = (long long)n_frames * a.n_out_ant;
    if (grid > 0x7fffffffLL) return set_error(ctx, RADAR_E_INVALID, "too many frames in one call");
    rd_spec_kernel<C><<<(unsigned)grid, C::NT, smem, st>>>(a);
    ctx->last_rd_kernel = "rd_spec_kernel";
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// Returns RADAR_OK when a specialised kernel was launched, 1 when the shape has no specialisation (caller falls back).
int launch_range_doppler_spec(radar_ctx* ctx, const void* adc, int n_frames, const int32_t* ant_list, int n_ant,
                              int window_flags, cf* rd_out, float* mag_out, int mag_antenna, cudaStream_t st) {
    if (reinterpret_cast<uintptr_t>(adc) % 16 != 0) return 1;      // cp.async pieces assume a 16-byte aligned batch
    RdSpecArgs a;
    a.iq = reinterpret_cast<const uint32_t*>(adc);
    a.R = ctx->shape.num_rx; a.cpl = ctx->shape.chirps_per_loop;
    a.n_out_ant = ant_list ? n_ant : ctx->V;
    a.ant_list = ant_list;
    a.winS = (window_flags & RADAR_WIN_RANGE) ? ctx->fftS.hann : ctx->fftS.ones;
    a.winL = (window_flags & RADAR_WIN_DOPPLER) ? ctx->fftL.hann : ctx->fftL.ones;
    a.twS = ctx->fftS.tw; a.twL = ctx->fftL.tw;
    a.rd_out = rd_out; a.mag_out = mag_out; a.mag_antenna = mag_antenna;
    const int S = ctx->shape.num_samples, L = ctx->shape.num_loops;
#define RB_SPEC(s, l, rs1, rs2, rl1, rl2, nt) \
    if (S == s && L == l) return launch_spec<RdShape<s, l, rs1, rs2, rl1, rl2, nt>>(ctx, a, n_frames, st);
    RB_SPEC(100, 30, 10, 10, 6, 5, 160)      // 1843_RadVel
    RB_SPEC(63, 70, 9, 7, 10, 7, 128)        // 6843_RadVel_ods_20Hz
    RB_SPEC(90, 100, 10, 9, 10, 10, 320)     // RadSAR_1843_official, 1843_RadVel_10Hz_lr
    RB_SPEC(63, 100, 9, 7, 10, 10, 224)      // 6843_RadVel_ods_10Hz / 5Hz, 1843_RadVel_20Hz, RadSAR_6843
    RB_SPEC(90, 80, 10, 9, 16, 5, 320)       // 1843_RadVel_10Hz_lr_SAR
    RB_SPEC(70, 40, 10, 7, 8, 5, 160)        // 6843_hand_ods_30Hz
    RB_SPEC(64, 40, 8, 8, 8, 5, 160)         // RadCloud
    RB_SPEC(130, 50, 13, 10, 10, 5, 256)     // 1843_RaGNNarok_UAV_5m
    RB_SPEC(100, 100, 10, 10, 10, 10, 320)   // RadSAR
#undef RB_SPEC
    return 1;
}

}  // namespace rb
