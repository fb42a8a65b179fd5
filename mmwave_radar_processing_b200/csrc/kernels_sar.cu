// K5a: multi-frame synthetic-array (SAR) beamformer.
//
// Reference arithmetic reproduced (processors/simple_synthetic_array_beamformer_processor_multiFrame.py:499-585):
//   out[:, g] = FFT_S( hanning(S) * sum_n hamming(N)[n] * x[s, n] * exp(+2 pi i (d_g . p_n) / lambda) )
// i.e. a complex GEMM  X[S x N] * W[N x G]  whose right operand is generated on the fly from the element
// positions and the pointing vectors, followed by a windowed range FFT per direction.
//
// One CTA per slab of kSarDirs directions.  The steering phase (d.p)/lambda reaches 1e2..1e4 turns, so it is
// reduced to its fractional part in fp64 before the fp32 sincospi.  The aperture is
// streamed through shared memory in chunks of kSarChunk elements; partial beams accumulate in registers per
// chunk and in the shared-memory tile across chunks; the tile then goes through the same shared-memory
// mixed-radix FFT as the range-Doppler kernel.  (FP32 FMA pipe; the tcgen05 split-precision variant is the
// planned follow-up - see DESIGN.md.)
#include "radar_internal.h"
#include "rd_core.cuh"

namespace rb {

constexpr int kSarThreads = 256;
constexpr int kSarDirs = 16;     // directions per CTA
constexpr int kSarChunk = 32;    // aperture elements per shared-memory chunk

struct SarArgs {
    const cf* aperture;          // [S][N]
    int S, N;
    const double* pos;           // [3][N]
    const double* dirs;          // [3][G]
    int G;
    double inv_lambda;
    int g_begin, g_end;
    const float* winS; const cf* twS; const uint16_t* posS;
    cf* out;                     // [S][g_end-g_begin]
    FftPlan planS;
    FastDiv divDirs;
};

__global__ void __launch_bounds__(kSarThreads) sar_kernel(const __grid_constant__ SarArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = a.S, Gp = kSarDirs | 1;
    SmemArena ar(smem_raw);
    cf* tile = ar.take<cf>(S * Gp);                  // beams [s][g_local]
    cf* xs = ar.take<cf>(S * kSarChunk);             // aperture chunk [s][n_local]
    cf* ws = ar.take<cf>(kSarChunk * kSarDirs);      // steering chunk [n_local][g_local]
    cf* tw = ar.take<cf>(S);
    float* win = ar.take<float>(S);
    uint16_t* pos = ar.take<uint16_t>(S);
    load_table(tw, a.twS, S);
    load_table(win, a.winS, S);
    load_table(pos, a.posS, S);
    const int g0 = a.g_begin + blockIdx.x * kSarDirs;
    const int n_dirs = min(kSarDirs, a.g_end - g0);
    const int cells = S * kSarDirs;
    for (int i = threadIdx.x; i < S * Gp; i += blockDim.x) tile[i] = cmake(0.f, 0.f);
    __syncthreads();

    for (int n0 = 0; n0 < a.N; n0 += kSarChunk) {
        const int nc = min(kSarChunk, a.N - n0);
        // aperture chunk (coalesced along n)
        for (int i = threadIdx.x; i < S * kSarChunk; i += blockDim.x) {
            const int s = i / kSarChunk, nl = i - s * kSarChunk;
            xs[i] = nl < nc ? __ldg(a.aperture + (long long)s * a.N + n0 + nl) : cmake(0.f, 0.f);
        }
        // steering chunk: hamming(N)[n] * exp(+2 pi i frac(d.p/lambda))
        for (int i = threadIdx.x; i < kSarChunk * kSarDirs; i += blockDim.x) {
            const int nl = i / kSarDirs, gl = i - nl * kSarDirs;
            cf w = cmake(0.f, 0.f);
            if (nl < nc && gl < n_dirs) {
                const int n = n0 + nl, g = g0 + gl;
                const double turns = (a.dirs[g] * a.pos[n] + a.dirs[a.G + g] * a.pos[a.N + n] +
                                      a.dirs[2 * a.G + g] * a.pos[2 * a.N + n]) * a.inv_lambda;
                const float frac = (float)(turns - rint(turns));
                float sn, cs;
                sincospif(2.0f * frac, &sn, &cs);
                const double ham = a.N > 1 ? 0.54 - 0.46 * cospi(2.0 * (double)n / (double)(a.N - 1)) : 1.0;
                w = cmake(cs * (float)ham, sn * (float)ham);
            }
            ws[i] = w;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < cells; i += blockDim.x) {
            const int s = (int)a.divDirs.div((uint32_t)i);
            const int gl = i - s * kSarDirs;
            cf acc = tile[s * Gp + gl];
            const cf* xr = xs + s * kSarChunk;
#pragma unroll 8
            for (int nl = 0; nl < kSarChunk; ++nl) {
                const cf x = xr[nl], w = ws[nl * kSarDirs + gl];
                acc.x = fmaf(x.x, w.x, fmaf(-x.y, w.y, acc.x));
                acc.y = fmaf(x.x, w.y, fmaf(x.y, w.x, acc.y));
            }
            tile[s * Gp + gl] = acc;
        }
        __syncthreads();
    }
    // Hann over range samples, then the range FFT of every beam
    for (int i = threadIdx.x; i < cells; i += blockDim.x) {
        const int s = (int)a.divDirs.div((uint32_t)i);
        const int gl = i - s * kSarDirs;
        tile[s * Gp + gl] = cscale(tile[s * Gp + gl], win[s]);
    }
    __syncthreads();
    fft_axis(tile, Gp, 1, kSarDirs, a.divDirs, a.planS, tw);
    const int width = a.g_end - a.g_begin;
    for (int i = threadIdx.x; i < cells; i += blockDim.x) {
        const int k = (int)a.divDirs.div((uint32_t)i);
        const int gl = i - k * kSarDirs;
        if (gl < n_dirs) a.out[(long long)k * width + (g0 - a.g_begin) + gl] = tile[pos[k] * Gp + gl];
    }
}

int launch_sar_image(radar_ctx* ctx, const cf* aperture, int num_samples, int n_elems, const double* positions,
                     const double* dirs, int n_dirs, double lambda_m, int g_begin, int g_end, cf* out, cudaStream_t st) {
    if (num_samples != ctx->shape.num_samples)
        return set_error(ctx, RADAR_E_INVALID, "SAR aperture has %d samples, ctx was built for %d", num_samples, ctx->shape.num_samples);
    if (n_elems <= 0 || n_dirs <= 0 || !(lambda_m > 0)) return set_error(ctx, RADAR_E_INVALID, "empty aperture / grid or bad wavelength");
    if (g_begin < 0 || g_end > n_dirs || g_begin > g_end) return set_error(ctx, RADAR_E_INVALID, "voxel slice [%d,%d) outside [0,%d)", g_begin, g_end, n_dirs);
    if (g_begin == g_end) return RADAR_OK;
    if (ctx->sar_mode != 1) {              // tensor-core path (kernels_sar_tc.cu); 1 = shape not handled there
        const int rc = launch_sar_image_tc(ctx, aperture, num_samples, n_elems, positions, dirs, n_dirs, lambda_m, g_begin, g_end, out, st);
        if (rc <= 0) return rc;
    }
    SarArgs a;
    a.aperture = aperture; a.S = num_samples; a.N = n_elems; a.pos = positions; a.dirs = dirs; a.G = n_dirs;
    a.inv_lambda = 1.0 / lambda_m;
    a.g_begin = g_begin; a.g_end = g_end;
    a.winS = ctx->fftS.hann; a.twS = ctx->fftS.tw; a.posS = ctx->fftS.pos;
    a.out = out;
    a.planS = ctx->fftS.plan;
    a.divDirs = make_fastdiv(kSarDirs);
    const int S = num_samples;
    const size_t smem = arena_bytes((size_t)S * (kSarDirs | 1), sizeof(cf)) + arena_bytes((size_t)S * kSarChunk, sizeof(cf)) +
                        arena_bytes(kSarChunk * kSarDirs, sizeof(cf)) + arena_bytes(S, sizeof(cf)) + arena_bytes(S, 4) + arena_bytes(S, 2);
    if (smem > (size_t)ctx->max_smem_optin) return set_error(ctx, RADAR_E_UNSUPPORTED, "SAR tile exceeds shared memory");
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(sar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (g_end - g_begin + kSarDirs - 1) / kSarDirs;
    sar_kernel<<<grid, kSarThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    // radar_ctx_set_image_mirrors on this (comparison / fallback) path: strided copies of the finished tile
    const size_t row_bytes = (size_t)(g_end - g_begin) * sizeof(cf);
    for (int m = 0; m < ctx->n_image_mirrors; ++m)
        RB_CUDA_OK(ctx, cudaMemcpy2DAsync(ctx->image_mirrors[m] + ctx->image_col0, (size_t)ctx->image_row_stride * sizeof(cf), out, row_bytes,
                                          row_bytes, (size_t)S, cudaMemcpyDeviceToDevice, st));
    return RADAR_OK;
}

}  // namespace rb
