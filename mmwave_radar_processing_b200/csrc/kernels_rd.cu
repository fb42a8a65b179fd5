// Range-Doppler family: fused int16-deinterleave + TDM virtual-array gather + Hann windows + range FFT +
// Doppler FFT (+fftshift, +magnitude) with the whole (S x L) antenna tile resident in shared memory;
// the coarse range profile; the range-angle heat map; Doppler beam sharpening.
//
// Reference arithmetic reproduced (paths relative to mmwave_radar_processing/ in the reference):
//   processors/virtual_array_reformater.py:44-65, processors/range_doppler_resp.py:49-110,
//   processors/range_resp.py:32-57, processors/range_angle_resp.py:55-122,
//   processors/range_angle_resp_dbs_enhanced.py:137-304.
#include "radar_internal.h"
#include "rd_core.cuh"

namespace rb {

constexpr int kRdThreads = 256;

// ------------------------------------------------------------------------------------------------
// ADC element fetch for the two ingest layouts (see include/radar_b200.h)
// ------------------------------------------------------------------------------------------------
struct AdcView {
    const void* base;
    int in_kind;
    int R, S, cpl, L, V, C;
};

// x[v][s][l] of frame f as complex float (exact for int16).
__device__ __forceinline__ cf adc_fetch(const AdcView& a, long long frame, int v, int s, int l) {
    if (a.in_kind == RADAR_IN_IQ16) {
        const int tx = v / a.R, rx = v - tx * a.R;
        const long long row = ((frame * a.C + (long long)l * a.cpl + tx) * a.R + rx) * a.S + s;
        const unsigned w = __ldg(reinterpret_cast<const unsigned*>(a.base) + row);
        return cmake((float)(short)(w & 0xffffu), (float)(short)(w >> 16));
    } else {
        const long long idx = ((frame * a.V + v) * a.S + s) * (long long)a.L + l;
        return __ldg(reinterpret_cast<const cf*>(a.base) + idx);
    }
}

// ------------------------------------------------------------------------------------------------
// K1+K2 fused range-Doppler: one CTA per (frame, output antenna)
// ------------------------------------------------------------------------------------------------
struct RdArgs {
    AdcView adc;
    int n_out_ant;
    const int32_t* ant_list;
    const float* ant_weight;
    const float* winS;
    const float* winL;
    const cf* twS;
    const cf* twL;
    const uint16_t* posS;
    const uint16_t* posL;
    cf* rd_out;
    float* mag_out;
    int mag_antenna;
    int range_only;            // RADAR_RD_RANGE_ONLY: stop after the range transform, chirp axis left in loop order
    FftPlan planS, planL;
    FastDiv divS, divL;
};

struct RdSmem {
    cf* tile; cf* twS; cf* twL; float* winS; float* winL; uint16_t* posS; uint16_t* srcL;
};

__device__ __forceinline__ RdSmem rd_carve(unsigned char* raw, int S, int L) {
    SmemArena ar(raw);
    RdSmem sm;
    sm.tile = ar.take<cf>(S * odd_pitch(L));
    sm.twS = ar.take<cf>(S);
    sm.twL = ar.take<cf>(L);
    sm.winS = ar.take<float>(S);
    sm.winL = ar.take<float>(L);
    sm.posS = ar.take<uint16_t>(S);
    sm.srcL = ar.take<uint16_t>(L);
    return sm;
}
__host__ __device__ static inline size_t rd_smem_bytes(int S, int L) {
    return arena_bytes((size_t)S * (L | 1), sizeof(cf)) + arena_bytes(S, sizeof(cf)) + arena_bytes(L, sizeof(cf)) +
           arena_bytes(S, 4) + arena_bytes(L, 4) + arena_bytes(S, 2) + arena_bytes(L, 2);
}

__device__ __forceinline__ void rd_load_tables(const RdSmem& sm, const RdArgs& a) {
    const int S = a.adc.S, L = a.adc.L;
    load_table(sm.twS, a.twS, S);
    load_table(sm.twL, a.twL, L);
    load_table(sm.winS, a.winS, S);
    load_table(sm.winL, a.winL, L);
    load_table(sm.posS, a.posS, S);
    // fftshift folded into the gather: output column d holds frequency bin (d - L/2) mod L
    for (int d = threadIdx.x; d < L; d += blockDim.x) {
        int f = d - L / 2;
        if (f < 0) f += L;
        sm.srcL[d] = a.range_only ? (uint16_t)d : a.posL[f];
    }
}

// Load one antenna's (S x L) samples with the separable window, run both FFTs in place.
// On return tile[posS[k]*Lp + srcL[d]] = RD[k][d].
__device__ __forceinline__ void rd_compute_tile(const RdSmem& sm, const RdArgs& a, long long frame, int v, float weight) {
    const int S = a.adc.S, L = a.adc.L, Lp = odd_pitch(L);
    const int total = S * L;
    if (a.adc.in_kind == RADAR_IN_IQ16) {
        for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
            const int l = (int)a.divS.div((uint32_t)idx);
            const int s = idx - l * S;
            const cf x = adc_fetch(a.adc, frame, v, s, l);
            const float w = sm.winS[s] * sm.winL[l] * weight;
            sm.tile[s * Lp + l] = cmake(x.x * w, x.y * w);
        }
    } else {
        for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
            const int s = (int)a.divL.div((uint32_t)idx);
            const int l = idx - s * L;
            const cf x = adc_fetch(a.adc, frame, v, s, l);
            const float w = sm.winS[s] * sm.winL[l] * weight;
            sm.tile[s * Lp + l] = cmake(x.x * w, x.y * w);
        }
    }
    __syncthreads();
    fft_axis(sm.tile, /*strideF=*/Lp, /*strideO=*/1, /*O=*/L, a.divL, a.planS, sm.twS);   // range
    if (!a.range_only) fft_axis(sm.tile, /*strideF=*/1, /*strideO=*/Lp, /*O=*/S, a.divS, a.planL, sm.twL);   // Doppler
}

__global__ void __launch_bounds__(kRdThreads) rd_fused_kernel(const __grid_constant__ RdArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = a.adc.S, L = a.adc.L, Lp = odd_pitch(L);
    RdSmem sm = rd_carve(smem_raw, S, L);
    rd_load_tables(sm, a);
    const long long frame = blockIdx.x / a.n_out_ant;
    const int vo = blockIdx.x - (int)frame * a.n_out_ant;
    const int v = a.ant_list ? a.ant_list[vo] : vo;
    const float weight = a.ant_weight ? a.ant_weight[vo] : 1.0f;
    __syncthreads();
    rd_compute_tile(sm, a, frame, v, weight);

    const int total = S * L;
    const long long tile_base = ((long long)frame * a.n_out_ant + vo) * total;
    const bool want_mag_all = a.mag_out && a.mag_antenna < 0;
    const bool want_mag_one = a.mag_out && a.mag_antenna == vo;
    float* mag_dst = want_mag_all ? a.mag_out + tile_base : (want_mag_one ? a.mag_out + frame * total : nullptr);
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
        const int k = (int)a.divL.div((uint32_t)idx);
        const int d = idx - k * L;
        const cf val = sm.tile[sm.posS[k] * Lp + sm.srcL[d]];
        if (a.rd_out) a.rd_out[tile_base + idx] = val;
        if (mag_dst) mag_dst[idx] = sqrtf(val.x * val.x + val.y * val.y);
    }
}

static int check_in_kind(radar_ctx* ctx, int in_kind) {
    if (in_kind != RADAR_IN_IQ16 && in_kind != RADAR_IN_CUBE64)
        return set_error(ctx, RADAR_E_INVALID, "unknown in_kind %d", in_kind);
    return RADAR_OK;
}

static AdcView make_view(const radar_ctx* ctx, const void* adc, int in_kind) {
    AdcView v;
    v.base = adc; v.in_kind = in_kind;
    v.R = ctx->shape.num_rx; v.S = ctx->shape.num_samples; v.cpl = ctx->shape.chirps_per_loop;
    v.L = ctx->shape.num_loops; v.V = ctx->V; v.C = ctx->C;
    return v;
}

template <typename K>
static int opt_in_smem(radar_ctx* ctx, K kernel, size_t bytes, const char* what) {
    if (bytes > (size_t)ctx->max_smem_optin)
        return set_error(ctx, RADAR_E_UNSUPPORTED, "%s needs %zu bytes of shared memory per CTA (device limit %d)",
                         what, bytes, ctx->max_smem_optin);
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return RADAR_OK;
}

int launch_range_doppler(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, const int32_t* ant_list,
                         int n_ant, const float* ant_weight, int window_flags, cf* rd_out, float* mag_out,
                         int mag_antenna, cudaStream_t st) {
    if (int rc = check_in_kind(ctx, in_kind)) return rc;
    if (n_frames <= 0) return RADAR_OK;
    RdArgs a;
    a.adc = make_view(ctx, adc, in_kind);
    a.n_out_ant = ant_list ? n_ant : ctx->V;
    if (a.n_out_ant <= 0) return set_error(ctx, RADAR_E_INVALID, "empty antenna list");
    if (mag_out && mag_antenna >= a.n_out_ant) return set_error(ctx, RADAR_E_INVALID, "mag_antenna %d out of range", mag_antenna);
    const bool range_only = (window_flags & RADAR_RD_RANGE_ONLY) != 0;
    a.range_only = range_only ? 1 : 0;
    if (!range_only && in_kind == RADAR_IN_IQ16 && !ant_weight && !ant_list && ctx->force_generic == 0 && ctx->rd_reg) {
        const int rc = launch_range_doppler_reg(ctx, adc, n_frames, window_flags, rd_out, mag_out, mag_antenna, st);
        if (rc <= 0) return rc;            // launched (0) or failed (<0); 1 = no register-resident instantiation for this shape
    }
    if (!range_only && in_kind == RADAR_IN_IQ16 && !ant_weight && !ant_list && ctx->force_generic == 0) {
        const int rc = launch_range_doppler_stream(ctx, adc, n_frames, window_flags, rd_out, mag_out, mag_antenna, st);
        if (rc <= 0) return rc;            // launched (0) or failed (<0); 1 = no streaming instantiation for this shape
    }
    if (!range_only && in_kind == RADAR_IN_IQ16 && !ant_weight && ctx->force_generic != 1) {
        const int rc = launch_range_doppler_spec(ctx, adc, n_frames, ant_list, n_ant, window_flags, rd_out, mag_out, mag_antenna, st);
        if (rc <= 0) return rc;            // launched (0) or failed (<0); 1 = no specialisation for this shape
    }
    a.ant_list = ant_list;
    a.ant_weight = ant_weight;
    a.winS = (window_flags & RADAR_WIN_RANGE) ? ctx->fftS.hann : ctx->fftS.ones;
    a.winL = (window_flags & RADAR_WIN_DOPPLER) ? ctx->fftL.hann : ctx->fftL.ones;
    a.twS = ctx->fftS.tw; a.twL = ctx->fftL.tw;
    a.posS = ctx->fftS.pos; a.posL = ctx->fftL.pos;
    a.rd_out = rd_out; a.mag_out = mag_out; a.mag_antenna = mag_antenna;
    a.planS = ctx->fftS.plan; a.planL = ctx->fftL.plan;
    a.divS = make_fastdiv((uint32_t)a.adc.S); a.divL = make_fastdiv((uint32_t)a.adc.L);
    const size_t smem = rd_smem_bytes(a.adc.S, a.adc.L);
    if (int rc = opt_in_smem(ctx, rd_fused_kernel, smem, "range-Doppler tile")) return rc;
    const long long grid = (long long)n_frames * a.n_out_ant;
    if (grid > 0x7fffffffLL) return set_error(ctx, RADAR_E_INVALID, "too many frames in one call");
    rd_fused_kernel<<<(unsigned)grid, kRdThreads, smem, st>>>(a);
    ctx->last_rd_kernel = "rd_fused_kernel";
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ------------------------------------------------------------------------------------------------
// a10: coarse range profile - one CTA per frame, tile [s][v]
// ------------------------------------------------------------------------------------------------
struct ProfArgs {
    AdcView adc;
    int loop_idx;
    const float* winS; const cf* twS; const uint16_t* posS;
    float* out;
    FftPlan planS; FastDiv divS, divV;
};

__global__ void __launch_bounds__(kRdThreads) range_profile_kernel(const __grid_constant__ ProfArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = a.adc.S, V = a.adc.V, Vp = odd_pitch(V);
    SmemArena ar(smem_raw);
    cf* tile = ar.take<cf>(S * Vp);
    cf* tw = ar.take<cf>(S);
    float* win = ar.take<float>(S);
    uint16_t* pos = ar.take<uint16_t>(S);
    load_table(tw, a.twS, S);
    load_table(win, a.winS, S);
    load_table(pos, a.posS, S);
    __syncthreads();
    const long long frame = blockIdx.x;
    for (int idx = threadIdx.x; idx < S * V; idx += blockDim.x) {
        const int v = (int)a.divS.div((uint32_t)idx);
        const int s = idx - v * S;
        const cf x = adc_fetch(a.adc, frame, v, s, a.loop_idx);
        tile[s * Vp + v] = cscale(x, win[s]);
    }
    __syncthreads();
    fft_axis(tile, Vp, 1, V, a.divV, a.planS, tw);
    for (int k = threadIdx.x; k < S; k += blockDim.x) {
        const cf* row = tile + pos[k] * Vp;
        double acc = 0.0;
        for (int v = 0; v < V; ++v) acc += (double)sqrtf(row[v].x * row[v].x + row[v].y * row[v].y);
        a.out[frame * S + k] = (float)(acc / (double)V);
    }
}

int launch_range_profile(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx, float* out,
                         cudaStream_t st) {
    if (int rc = check_in_kind(ctx, in_kind)) return rc;
    if (loop_idx < 0 || loop_idx >= ctx->shape.num_loops) return set_error(ctx, RADAR_E_INVALID, "loop_idx %d out of range", loop_idx);
    if (n_frames <= 0) return RADAR_OK;
    ProfArgs a;
    a.adc = make_view(ctx, adc, in_kind);
    a.loop_idx = loop_idx;
    a.winS = ctx->fftS.hann; a.twS = ctx->fftS.tw; a.posS = ctx->fftS.pos;
    a.out = out;
    a.planS = ctx->fftS.plan;
    a.divS = make_fastdiv((uint32_t)a.adc.S); a.divV = make_fastdiv((uint32_t)a.adc.V);
    const int S = a.adc.S, V = a.adc.V;
    const size_t smem = arena_bytes((size_t)S * (V | 1), sizeof(cf)) + arena_bytes(S, sizeof(cf)) + arena_bytes(S, 4) + arena_bytes(S, 2);
    if (int rc = opt_in_smem(ctx, range_profile_kernel, smem, "range profile tile")) return rc;
    range_profile_kernel<<<(unsigned)n_frames, kRdThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ------------------------------------------------------------------------------------------------
// K3a: range-angle heat map - one CTA per frame; range FFT of one chirp per antenna, then a
// zero-padded angle DFT evaluated directly (K = n_ant <= V terms per output cell)
// ------------------------------------------------------------------------------------------------
struct RaArgs {
    AdcView adc;
    int loop_idx;
    const int32_t* ant_list; int n_ant;
    int A;
    const float* winS; const float* winV;     // winV indexed by ORIGINAL antenna (may be all-ones)
    const cf* twS; const uint16_t* posS; const cf* twA;
    float* out;
    long long n_frames; int frames_per_pass;
    int fft4;             // ra_angle_stage_fft4 (2/4/8 antennas, power-of-two bins, A/N >= 4, 16-byte aligned output)
    FftPlan planS; FastDiv divS, divN, divW;
};

// sqrt to 1 ulp in one MUFU (the IEEE sqrtf costs a Newton step and a guarded slow path per output cell; the heat map is
// compared at 1e-4 relative L2)
__device__ __forceinline__ float sqrt_fast(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Zero-padded angle DFT of one range row + fftshift + magnitude: out[k][ao] = | sum_j row[j] w^j |, w = W_A^((ao - A/2) mod A),
// evaluated by Horner's rule.  One warp per range bin: the row is read once (broadcast) and kept in registers, lanes run
// along the angle bin so the stores are coalesced.  NC > 0: antenna count compiled in.  NB > 0: A == 32*NB, every lane keeps
// the NB twiddles of its angle bins in registers for the whole kernel.
template <int NC, int NB>
__device__ __forceinline__ void ra_angle_stage(const cf* tile, int Np, const uint16_t* pos, const cf* twA, int S, int A,
                                               float* dst, int n_rt = 0) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int N = NC > 0 ? NC : n_rt;
    cf wreg[NB > 0 ? NB : 1];
    if (NB > 0) {
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            int ap = lane + 32 * b - A / 2;          // fftshift: column ao holds bin (ao - A/2) mod A
            if (ap < 0) ap += A;
            wreg[b] = twA[ap];
        }
    }
    for (int k = warp; k < S; k += nw) {
        const cf* row = tile + pos[k] * Np;
        cf r[NC > 0 ? NC : 1];
        if (NC > 0) {
#pragma unroll
            for (int j = 0; j < NC; ++j) r[j] = row[j];
        }
        float* out = dst + k * A + lane;
        auto one = [&](cf w) {
            cf acc;
            if (NC > 0) {
                acc = r[NC - 1];
#pragma unroll
                for (int j = NC - 2; j >= 0; --j)
                    acc = cmake(fmaf(acc.x, w.x, fmaf(-acc.y, w.y, r[j].x)), fmaf(acc.x, w.y, fmaf(acc.y, w.x, r[j].y)));
            } else {
                acc = row[N - 1];
                for (int j = N - 2; j >= 0; --j) {
                    const cf x = row[j];
                    acc = cmake(fmaf(acc.x, w.x, fmaf(-acc.y, w.y, x.x)), fmaf(acc.x, w.y, fmaf(acc.y, w.x, x.y)));
                }
            }
            return sqrt_fast(acc.x * acc.x + acc.y * acc.y);
        };
        if (NB > 0) {
            float m[NB > 0 ? NB : 1];
#pragma unroll
            for (int b = 0; b < NB; ++b) m[b] = one(wreg[b]);
#pragma unroll
            for (int b = 0; b < NB; ++b) out[32 * b] = m[b];
        } else {
            for (int ao = lane; ao < A; ao += 32) {
                int ap = ao - A / 2;
                if (ap < 0) ap += A;
                out[ao - lane] = one(twA[ap]);
            }
        }
    }
}
// Same output as ra_angle_stage for NC in {2, 4, 8} antennas and A a power of two: the zero-padded A-point transform of NC
// inputs splits as bin a' = r + (A/NC) q:  X[a'] = sum_j (x_j W_A^(j r)) W_NC^(j q), i.e. NC-1 twiddles and ONE NC-point
// register butterfly give the NC bins of a residue r - about 3x fewer operations than NC-term Horner sums per bin.
// One lane per (range bin, residue); a warp covers 32 consecutive residues (whole rows when A/NC <= 32), and for a fixed q
// neighbouring lanes store neighbouring bins.
template <int NC>
__device__ __forceinline__ void ra_angle_stage_fft(const cf* tile, int Np, const uint16_t* pos, const cf* twA, int S, int A,
                                                   float* dst) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int M = A / NC, total = S * M;
    const int m_shift = 31 - __clz(M);
    const bool fixed_r = M <= 32;                       // then r = lane mod M for every task of this lane
    cf wj[NC];
#pragma unroll
    for (int j = 1; j < NC; ++j) wj[j] = twA[(j * (lane & (M - 1))) & (A - 1)];
    for (int t0 = warp * 32; t0 < total; t0 += nw * 32) {
        const int t = t0 + lane;
        if (t >= total) break;
        const int k = t >> m_shift, r = t & (M - 1);
        const cf* row = tile + pos[k] * Np;
        cf v[NC];
        v[0] = row[0];
#pragma unroll
        for (int j = 1; j < NC; ++j) v[j] = cmul(row[j], fixed_r ? wj[j] : twA[(j * r) & (A - 1)]);
        Dft<NC>::run(v);
        float* out = dst + k * A;
#pragma unroll
        for (int q = 0; q < NC; ++q) out[(r + M * q + A / 2) & (A - 1)] = sqrt_fast(v[q].x * v[q].x + v[q].y * v[q].y);
    }
}

// Four neighbouring residues per lane (A/NC >= 4): the NC - 1 twiddles of a residue come from a transposed table
// twT[j-1][r] = W_A^(j r) (neighbouring lanes read neighbouring entries: no bank conflicts, 128-bit loads), the four
// butterflies share the row loads and the index arithmetic, and the four bins r .. r+3 of every q leave as ONE 128-bit store.
// ncu on the one-residue form at A = 1024 (profiles/r3e): issue-bound, 26 thread-instructions per output cell, 26 % of them
// address arithmetic, a third of the shared-memory wavefronts bank conflicts of the stride-j twiddle reads.
template <int NC>
__device__ __forceinline__ void ra_angle_stage_fft4(const cf* tile, int Np, const uint16_t* pos, const cf* twT, int S, int A,
                                                    float* dst) {
    const int M = A / NC, M4 = M >> 2, total = S * M4;
    const int m_shift = 31 - __clz(M4);
    for (int t = threadIdx.x; t < total; t += blockDim.x) {
        const int k = t >> m_shift, r = (t & (M4 - 1)) << 2;
        const cf* row = tile + pos[k] * Np;
        cf v[4][NC];
        const cf x0 = row[0];
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i][0] = x0;
#pragma unroll
        for (int j = 1; j < NC; ++j) {
            const cf x = row[j];
            const float4 wa = *reinterpret_cast<const float4*>(twT + (j - 1) * M + r);
            const float4 wb = *reinterpret_cast<const float4*>(twT + (j - 1) * M + r + 2);
            v[0][j] = cmul(x, cmake(wa.x, wa.y));
            v[1][j] = cmul(x, cmake(wa.z, wa.w));
            v[2][j] = cmul(x, cmake(wb.x, wb.y));
            v[3][j] = cmul(x, cmake(wb.z, wb.w));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) Dft<NC>::run(v[i]);
        float* out = dst + k * A;
#pragma unroll
        for (int q = 0; q < NC; ++q) {
            float4 m;
            m.x = sqrt_fast(v[0][q].x * v[0][q].x + v[0][q].y * v[0][q].y);
            m.y = sqrt_fast(v[1][q].x * v[1][q].x + v[1][q].y * v[1][q].y);
            m.z = sqrt_fast(v[2][q].x * v[2][q].x + v[2][q].y * v[2][q].y);
            m.w = sqrt_fast(v[3][q].x * v[3][q].x + v[3][q].y * v[3][q].y);
            *reinterpret_cast<float4*>(out + ((r + M * q + A / 2) & (A - 1))) = m;
        }
    }
}

template <int NC>
__device__ __forceinline__ void ra_angle_stage_any(const cf* tile, int Np, const uint16_t* pos, const cf* twA, int S, int A,
                                                   float* dst, int n_rt = 0, const cf* twT = nullptr) {
    if constexpr (NC == 2 || NC == 4 || NC == 8) {
        if (twT) { ra_angle_stage_fft4<NC>(tile, Np, pos, twT, S, A, dst); return; }
        if (A >= 32 && (A & (A - 1)) == 0) { ra_angle_stage_fft<NC>(tile, Np, pos, twA, S, A, dst); return; }
    }
    if (A == 64) ra_angle_stage<NC, 2>(tile, Np, pos, twA, S, A, dst, n_rt);
    else if (A == 128) ra_angle_stage<NC, 4>(tile, Np, pos, twA, S, A, dst, n_rt);
    else if (A == 32) ra_angle_stage<NC, 1>(tile, Np, pos, twA, S, A, dst, n_rt);
    else ra_angle_stage<NC, 0>(tile, Np, pos, twA, S, A, dst, n_rt);
}

// FC frames per CTA pass: the range FFT runs over FC*N columns at once (the generic stage executor puts lanes along the
// columns, so a single frame's 4 antennas would leave most of the CTA idle) and the tables are staged once per CTA.
__global__ void __launch_bounds__(kRdThreads) range_angle_kernel(const __grid_constant__ RaArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = a.adc.S, N = a.n_ant, FC = a.frames_per_pass, W = FC * N, P = odd_pitch(W), A = a.A;
    SmemArena ar(smem_raw);
    cf* tile = ar.take<cf>(S * P);
    cf* tw = ar.take<cf>(S);
    cf* twA = ar.take<cf>(A);
    float* win = ar.take<float>(S);
    uint16_t* pos = ar.take<uint16_t>(S);
    cf* twT = a.fft4 ? ar.take<cf>((N - 1) * (A / N)) : nullptr;         // transposed twiddles of ra_angle_stage_fft4
    load_table(tw, a.twS, S);
    load_table(twA, a.twA, A);
    if (twT) {
        const int M = A / N;
        for (int i = threadIdx.x; i < (N - 1) * M; i += blockDim.x) {
            const int j = i / M + 1, r = i - (j - 1) * M;
            twT[i] = a.twA[(j * r) & (A - 1)];
        }
    }
    load_table(win, a.winS, S);
    load_table(pos, a.posS, S);
    for (int i = threadIdx.x; i < S * P; i += blockDim.x) tile[i] = cmake(0.f, 0.f);
    __syncthreads();
    for (long long f0 = (long long)blockIdx.x * FC; f0 < a.n_frames; f0 += (long long)gridDim.x * FC) {
        const int nf = (int)min((long long)FC, a.n_frames - f0);
        for (int idx = threadIdx.x; idx < S * N * nf; idx += blockDim.x) {
            const int c = (int)a.divS.div((uint32_t)idx);          // column = f*N + j
            const int s = idx - c * S;
            const int f = (int)a.divN.div((uint32_t)c), j = c - f * N;
            const int v = a.ant_list ? a.ant_list[j] : j;
            const cf x = adc_fetch(a.adc, f0 + f, v, s, a.loop_idx);
            tile[s * P + c] = cscale(x, win[s] * a.winV[v]);
        }
        __syncthreads();
        fft_axis(tile, P, 1, W, a.divW, a.planS, tw);
        for (int f = 0; f < nf; ++f) {
            float* dst = a.out + (f0 + f) * (long long)S * A;
            const cf* t = tile + f * N;
            switch (N) {
                case 1: ra_angle_stage_any<1>(t, P, pos, twA, S, A, dst); break;
                case 2: ra_angle_stage_any<2>(t, P, pos, twA, S, A, dst, 0, twT); break;
                case 3: ra_angle_stage_any<3>(t, P, pos, twA, S, A, dst); break;
                case 4: ra_angle_stage_any<4>(t, P, pos, twA, S, A, dst, 0, twT); break;
                case 8: ra_angle_stage_any<8>(t, P, pos, twA, S, A, dst, 0, twT); break;
                case 12: ra_angle_stage_any<12>(t, P, pos, twA, S, A, dst); break;
                default: ra_angle_stage_any<0>(t, P, pos, twA, S, A, dst, N); break;
            }
        }
        __syncthreads();
    }
}

int launch_range_angle(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx,
                       const int32_t* ant_list, int n_ant, int num_angle_bins, int window_flags, float* out,
                       cudaStream_t st) {
    if (int rc = check_in_kind(ctx, in_kind)) return rc;
    if (loop_idx < 0 || loop_idx >= ctx->shape.num_loops) return set_error(ctx, RADAR_E_INVALID, "loop_idx %d out of range", loop_idx);
    const int N = ant_list ? n_ant : ctx->V;
    if (N <= 0) return set_error(ctx, RADAR_E_INVALID, "empty antenna list");
    if (num_angle_bins < N || num_angle_bins > 8192)
        return set_error(ctx, RADAR_E_INVALID, "num_angle_bins %d must be in [n_ant=%d, 8192]", num_angle_bins, N);
    if (n_frames <= 0) return RADAR_OK;
    AngleTables at;
    if (int rc = get_angle_tables(ctx, num_angle_bins, &at)) return rc;
    float* onesV = nullptr;
    if (int rc = get_hann(ctx, -ctx->V, &onesV)) return rc;     // negative length = all-ones table
    RaArgs a;
    a.adc = make_view(ctx, adc, in_kind);
    a.loop_idx = loop_idx;
    a.ant_list = ant_list; a.n_ant = N; a.A = num_angle_bins;
    a.winS = (window_flags & RADAR_WIN_RANGE) ? ctx->fftS.hann : ctx->fftS.ones;
    a.winV = (window_flags & RADAR_WIN_ANTENNA) ? ctx->hannV : onesV;
    a.twS = ctx->fftS.tw; a.posS = ctx->fftS.pos; a.twA = at.tw;
    a.out = out;
    a.planS = ctx->fftS.plan;
    const int S = a.adc.S;
    // frames per pass: up to a warp's worth of FFT columns and a tile under 64 KB; among the halvings of that, the one whose
    // pass count quantises best over the resident CTA slots (a pass of fc frames costs about fc + 1/2 frame times: the angle
    // stage is per frame, the range FFT loses lane occupancy as the columns shrink).  The grid is then balanced: every CTA
    // gets the same number of passes, to within one.
    int fc_max = 32 / N; if (fc_max < 1) fc_max = 1; if (fc_max > 8) fc_max = 8;
    while (fc_max > 1 && (size_t)S * ((fc_max * N) | 1) * sizeof(cf) > 64 * 1024) --fc_max;
    if (fc_max > n_frames) fc_max = n_frames;
    const bool pow2 = (num_angle_bins & (num_angle_bins - 1)) == 0;
    a.fft4 = ((N == 2 || N == 4 || N == 8) && pow2 && num_angle_bins / N >= 4 && num_angle_bins <= 2048 && ctx->ra_fft4 &&
              reinterpret_cast<uintptr_t>(out) % 16 == 0) ? 1 : 0;
    auto smem_for = [&](int fc) {
        return arena_bytes((size_t)S * ((fc * N) | 1), sizeof(cf)) + arena_bytes(S, sizeof(cf)) +
               arena_bytes(num_angle_bins, sizeof(cf)) + arena_bytes(S, 4) + arena_bytes(S, 2) +
               (a.fft4 ? arena_bytes((size_t)(N - 1) * (num_angle_bins / N), sizeof(cf)) : 0);
    };
    if (int rc = opt_in_smem(ctx, range_angle_kernel, smem_for(fc_max), "range-angle tile")) return rc;
    int fc = fc_max;
    long long slots = ctx->sm_count;
    double best_cost = 0.0;
    for (int cand = fc_max; cand >= 1; cand = cand > 1 ? (cand + 1) / 2 : 0) {
        if (ctx->ra_fc > 0 && cand != ctx->ra_fc && !(cand == fc_max && ctx->ra_fc > fc_max)) continue;
        int resident = 0;
        RB_CUDA_OK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, range_angle_kernel, kRdThreads, smem_for(cand)));
        if (resident < 1) resident = 1;
        const long long sl = (long long)resident * ctx->sm_count;
        const long long passes = ((long long)n_frames + cand - 1) / cand;
        const double cost = (double)((passes + sl - 1) / sl) * ((double)cand + 0.5);
        if (best_cost == 0.0 || cost < best_cost) { best_cost = cost; fc = cand; slots = sl; }
    }
    a.n_frames = n_frames; a.frames_per_pass = fc;
    a.divS = make_fastdiv((uint32_t)S); a.divN = make_fastdiv((uint32_t)N); a.divW = make_fastdiv((uint32_t)(fc * N));
    const size_t smem = smem_for(fc);
    const long long passes = ((long long)n_frames + fc - 1) / fc;
    const long long per_cta = (passes + slots - 1) / slots;
    const long long grid = (passes + per_cta - 1) / per_cta;
    range_angle_kernel<<<(unsigned)grid, kRdThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// ------------------------------------------------------------------------------------------------
// K5b: Doppler beam sharpening - one CTA per frame.  For each antenna of the subset the full RD tile is
// formed in shared memory (same code as K1+K2) and only the selected (range, Doppler-bin) cells are
// combined with the angle steering coefficient of each output angle.
// ------------------------------------------------------------------------------------------------
struct DbsArgs {
    RdArgs rd;
    int A, n_out;
    const int32_t* vel_bin; const int32_t* ang_bin;
    const cf* twA;
    cf* acc_global;      // used when the accumulator does not fit in shared memory
    int acc_in_smem;
    float* out;
    FastDiv divOut;
};

// 1024 threads: one CTA per frame holds a whole antenna tile (149 KB for 90 x 100) so only one CTA fits an SM; ncu on the
// 256-thread launch showed it latency-bound at 8 warps per SM (issue 23 %, long scoreboard 3.9 of 8 stall cycles).
// 256 / 512 / 768 / 1024 threads: 1.92 / 1.18 / 0.98 / 0.89 ms per 1024 frames of RadSAR_1843_official.
constexpr int kDbsThreads = 1024;
__global__ void __launch_bounds__(kDbsThreads) dbs_kernel(const __grid_constant__ DbsArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = a.rd.adc.S, L = a.rd.adc.L, Lp = odd_pitch(L);
    RdSmem sm = rd_carve(smem_raw, S, L);
    unsigned char* after = smem_raw + rd_smem_bytes(S, L);
    SmemArena ar(after);
    cf* twA = ar.take<cf>(a.A);
    int* selL = ar.take<int>(a.n_out);       // tile column holding Doppler bin vel_bin[i]
    int* selA = ar.take<int>(a.n_out);       // shifted angle frequency of ang_bin[i]
    const long long frame = blockIdx.x;
    const int total = S * a.n_out;
    cf* acc = a.acc_in_smem ? ar.take<cf>(total) : a.acc_global + frame * (long long)total;
    rd_load_tables(sm, a.rd);
    load_table(twA, a.twA, a.A);
    __syncthreads();
    for (int i = threadIdx.x; i < a.n_out; i += blockDim.x) {
        selL[i] = sm.srcL[a.vel_bin[i]];
        int f = a.ang_bin[i] - a.A / 2;
        if (f < 0) f += a.A;
        selA[i] = f;
    }
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) acc[idx] = cmake(0.f, 0.f);
    __syncthreads();
    for (int j = 0; j < a.rd.n_out_ant; ++j) {
        const int v = a.rd.ant_list ? a.rd.ant_list[j] : j;
        const float weight = a.rd.ant_weight ? a.rd.ant_weight[j] : 1.0f;
        rd_compute_tile(sm, a.rd, frame, v, weight);
        for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
            const int k = (int)a.divOut.div((uint32_t)idx);
            const int i = idx - k * a.n_out;
            const cf x = sm.tile[sm.posS[k] * Lp + selL[i]];
            const cf w = twA[(int)(((long long)j * selA[i]) % a.A)];
            cf c = acc[idx];
            c.x = fmaf(x.x, w.x, fmaf(-x.y, w.y, c.x));
            c.y = fmaf(x.x, w.y, fmaf(x.y, w.x, c.y));
            acc[idx] = c;
        }
        __syncthreads();
    }
    float* dst = a.out + frame * (long long)total;
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
        const cf c = acc[idx];
        dst[idx] = sqrtf(c.x * c.x + c.y * c.y);
    }
}

// K5b on top of the register-resident range-Doppler kernel: for the raw int16 layout and the shapes rd_reg_kernel serves,
// the windowed range-Doppler cube of a block of frames is written to scratch by that kernel (0.6-0.8 of the HBM roofline
// instead of the shared-memory FFT of dbs_kernel, one CTA per SM) and dbs_combine_kernel forms
//   out[f][k][i] = | sum_j hann_N[j] RD[f][v_j][k][vel_bin[i]] W_A^(j * (ang_bin[i] - A/2)) |
// from it, block of frames by block of frames.
struct DbsCombineArgs {
    const cf* cube;      // [frames][V][S][L]
    int V, S, L, N, A, n_out;
    const int32_t* ant_list; const float* ant_weight;
    const int32_t* vel_bin; const int32_t* ang_bin;
    const cf* twA;
    float* out;          // [frames][S][n_out]
    FastDiv divOut;
};

constexpr int kDbsCombineThreads = 256;
constexpr int kDbsCombineMaxAnt = 16;
__global__ void __launch_bounds__(kDbsCombineThreads) dbs_combine_kernel(const __grid_constant__ DbsCombineArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cf* wtab = reinterpret_cast<cf*>(smem_raw);                        // [N][n_out]: antenna weight x steering coefficient
    int* col = reinterpret_cast<int*>(wtab + (size_t)a.N * a.n_out);   // [n_out]: Doppler column of output angle i
    __shared__ long long ant_off[kDbsCombineMaxAnt];
    for (int i = threadIdx.x; i < a.n_out; i += blockDim.x) col[i] = a.vel_bin[i];
    for (int j = threadIdx.x; j < a.N; j += blockDim.x)
        ant_off[j] = (long long)(a.ant_list ? a.ant_list[j] : j) * a.S * a.L;
    for (int e = threadIdx.x; e < a.N * a.n_out; e += blockDim.x) {
        const int j = e / a.n_out, i = e - j * a.n_out;
        int f = a.ang_bin[i] - a.A / 2;
        if (f < 0) f += a.A;
        const cf w = a.twA[(int)(((long long)j * f) % a.A)];
        const float wt = a.ant_weight ? a.ant_weight[j] : 1.0f;
        wtab[e] = cmake(w.x * wt, w.y * wt);
    }
    __syncthreads();
    const long long frame = blockIdx.x;
    const cf* src = a.cube + frame * (long long)a.V * a.S * a.L;
    float* dst = a.out + frame * (long long)a.S * a.n_out;
    const int total = a.S * a.n_out;
    for (int idx = threadIdx.x + blockIdx.y * blockDim.x; idx < total; idx += blockDim.x * gridDim.y) {
        const int k = (int)a.divOut.div((uint32_t)idx);
        const int i = idx - k * a.n_out;
        const cf* p = src + (size_t)k * a.L + col[i];
        cf c = cmake(0.f, 0.f);
        for (int j = 0; j < a.N; ++j) {
            const cf x = __ldcg(p + ant_off[j]);
            const cf w = wtab[j * a.n_out + i];
            c.x = fmaf(x.x, w.x, fmaf(-x.y, w.y, c.x));
            c.y = fmaf(x.x, w.y, fmaf(x.y, w.x, c.y));
        }
        dst[idx] = sqrtf(c.x * c.x + c.y * c.y);
    }
}

// 0 = done, 1 = not applicable (caller runs dbs_kernel), < 0 = error
static int launch_dbs_two_pass(radar_ctx* ctx, const void* adc, int n_frames, const int32_t* ant_list, int N, int num_angle_bins,
                               const int32_t* vel_bin, const int32_t* ang_bin, int n_out, const cf* twA, const float* hannN, float* out,
                               cudaStream_t st) {
    if (!ctx->rd_reg || ctx->dbs_mode == 1 || N > kDbsCombineMaxAnt) return 1;
    if (reinterpret_cast<uintptr_t>(adc) & 15u) return 1;
    if (ctx->dbs_mode != 2) {               // RADAR_B200_DBS=2: keep the cube round trip (comparison)
        const int rc = launch_dbs_reg(ctx, adc, n_frames, ant_list, N, num_angle_bins, vel_bin, ang_bin, n_out, twA, hannN, out, st);
        if (rc <= 0) return rc;
    }
    const int S = ctx->shape.num_samples, L = ctx->shape.num_loops, V = ctx->V;
    const size_t frame_cells = (size_t)V * S * L;
    const size_t frame_in_bytes = (size_t)ctx->C * ctx->shape.num_rx * S * 4;
    if (frame_in_bytes & 15u) return 1;
    // frames per block: as many as 1 GiB of scratch holds.  Measured on RadSAR_1843_official, 1024 frames: blocks of 64 / 113
    // (L2-sized) / 256 / 1024 frames take 0.591 / 0.474 / 0.362 / 0.305 ms - the launches and the tail of the persistent
    // range-Doppler kernel cost more than the cube's trip through HBM.  RADAR_B200_DBS_CHUNK overrides.
    long long chunk = ctx->dbs_chunk > 0 ? ctx->dbs_chunk : (long long)((1ull << 30) / (frame_cells * sizeof(cf)));
    if (chunk < 16) chunk = 16;
    if (chunk > n_frames) chunk = n_frames;
    const size_t smem = ((size_t)N * n_out * sizeof(cf) + (size_t)n_out * sizeof(int) + 15) & ~(size_t)15;
    if (smem > 96 * 1024) return 1;
    const char* keep = ctx->last_rd_kernel;
    for (long long f0 = 0; f0 < n_frames; f0 += chunk) {
        const int nf = (int)((n_frames - f0 < chunk) ? n_frames - f0 : chunk);
        if (f0 == 0) {
            if (int rc = ensure_scratch(ctx, (size_t)chunk * frame_cells * sizeof(cf))) return rc;
        }
        cf* cube = reinterpret_cast<cf*>(ctx->scratch);
        const int rc = launch_range_doppler_reg(ctx, reinterpret_cast<const unsigned char*>(adc) + (size_t)f0 * frame_in_bytes, nf,
                                                RADAR_WIN_RANGE | RADAR_WIN_DOPPLER, cube, nullptr, 0, st);
        ctx->last_rd_kernel = keep;                         // radar_last_rd_kernel reports radar_range_doppler calls only
        if (rc != 0) return f0 == 0 ? rc : set_error(ctx, RADAR_E_CUDA, "range-Doppler kernel refused a later block of frames");
        DbsCombineArgs c{cube, V, S, L, N, num_angle_bins, n_out, ant_list, hannN, vel_bin, ang_bin, twA,
                         out + (size_t)f0 * S * n_out, make_fastdiv((uint32_t)n_out)};
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(dbs_combine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const int ysplit = (S * n_out + 8 * kDbsCombineThreads - 1) / (8 * kDbsCombineThreads);
        dbs_combine_kernel<<<dim3((unsigned)nf, (unsigned)(ysplit < 1 ? 1 : ysplit)), kDbsCombineThreads, smem, st>>>(c);
        RB_CUDA_OK(ctx, cudaGetLastError());
    }
    return 0;
}

int launch_dbs(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, const int32_t* ant_list, int n_ant,
               int num_angle_bins, const int32_t* vel_bin, const int32_t* ang_bin, int n_out, float* out,
               cudaStream_t st) {
    if (int rc = check_in_kind(ctx, in_kind)) return rc;
    const int N = ant_list ? n_ant : ctx->V;
    if (N <= 0 || n_out <= 0) return set_error(ctx, RADAR_E_INVALID, "empty antenna list or output grid");
    if (num_angle_bins < N || num_angle_bins > 8192)
        return set_error(ctx, RADAR_E_INVALID, "num_angle_bins %d must be in [n_ant=%d, 8192]", num_angle_bins, N);
    if (n_frames <= 0) return RADAR_OK;
    AngleTables at;
    if (int rc = get_angle_tables(ctx, num_angle_bins, &at)) return rc;
    float* hannN = nullptr;
    if (int rc = get_hann(ctx, N, &hannN)) return rc;           // Hann over the SUBSET's length (ref :176-177)
    if (in_kind == RADAR_IN_IQ16) {
        const int rc = launch_dbs_two_pass(ctx, adc, n_frames, ant_list, N, num_angle_bins, vel_bin, ang_bin, n_out, at.tw, hannN, out, st);
        if (rc <= 0) return rc;
    }
    DbsArgs a;
    a.rd.adc = make_view(ctx, adc, in_kind);
    a.rd.n_out_ant = N; a.rd.ant_list = ant_list; a.rd.ant_weight = hannN;
    a.rd.winS = ctx->fftS.hann; a.rd.winL = ctx->fftL.hann;
    a.rd.twS = ctx->fftS.tw; a.rd.twL = ctx->fftL.tw; a.rd.posS = ctx->fftS.pos; a.rd.posL = ctx->fftL.pos;
    a.rd.rd_out = nullptr; a.rd.mag_out = nullptr; a.rd.mag_antenna = 0; a.rd.range_only = 0;
    a.rd.planS = ctx->fftS.plan; a.rd.planL = ctx->fftL.plan;
    a.rd.divS = make_fastdiv((uint32_t)a.rd.adc.S); a.rd.divL = make_fastdiv((uint32_t)a.rd.adc.L);
    a.A = num_angle_bins; a.n_out = n_out;
    a.vel_bin = vel_bin; a.ang_bin = ang_bin; a.twA = at.tw; a.out = out;
    a.divOut = make_fastdiv((uint32_t)n_out);
    const int S = a.rd.adc.S, L = a.rd.adc.L;
    const size_t base = rd_smem_bytes(S, L) + arena_bytes(num_angle_bins, sizeof(cf)) + 2 * arena_bytes(n_out, 4);
    const size_t acc_bytes = arena_bytes((size_t)S * n_out, sizeof(cf));
    size_t smem = base + acc_bytes;
    a.acc_in_smem = 1; a.acc_global = nullptr;
    if (smem > (size_t)ctx->max_smem_optin) {
        smem = base; a.acc_in_smem = 0;
        if (int rc = ensure_scratch(ctx, (size_t)n_frames * S * n_out * sizeof(cf))) return rc;
        a.acc_global = reinterpret_cast<cf*>(ctx->scratch);
    }
    if (int rc = opt_in_smem(ctx, dbs_kernel, smem, "DBS tile")) return rc;
    dbs_kernel<<<(unsigned)n_frames, kDbsThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

}  // namespace rb
