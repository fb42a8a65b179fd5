// Persistent streaming range-Doppler kernel: device wrapper and launcher for the phases in rd_stream.cuh.
#include <cmath>
#include "radar_internal.h"
#include "rd_stream.cuh"

namespace rb {

// Item i of the batch = antenna group (i mod IPF) of frame (i div IPF), IPF = V/G groups per frame (V = CR).  The group's
// first virtual antenna v0 = tx*R + rx0 starts at word v0*S of the frame's first chirp block; its output tiles are
// consecutive in rd_out, so the output offset is linear in i.
template <class C>
__device__ __forceinline__ const uint32_t* rd_stream_item_src(const RdStreamParams<C>& a, unsigned item) {
    constexpr unsigned IPF = C::CR / C::G;
    const unsigned frame = item / IPF, grp = item - frame * IPF;
    return a.iq + (size_t)frame * ((size_t)C::ROW * C::L) + grp * (C::G * C::S);
}

template <class C>
__global__ void __launch_bounds__(C::NT, C::MINB) rd_stream_kernel(const __grid_constant__ RdStreamParams<C> a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cf* tiles = reinterpret_cast<cf*>(smem_raw);
    float* winS_s = reinterpret_cast<float*>(smem_raw + (size_t)C::NBUF * C::G * C::TILE * sizeof(cf));
    const int tid = threadIdx.x;
    for (int i = tid; i < C::S; i += C::NT) winS_s[i] = a.winS[i];

    auto finish_item = [&](unsigned item, cf* tile) {          // range stage 1 + stores of one item
        constexpr unsigned IPF = C::CR / C::G;
        const unsigned frame = item / IPF, grp = item - frame * IPF;
        const long long out_base = (long long)item * (C::G * C::S * C::L);
        float* mag = a.mag_out;
        long long mag_base = out_base;
        int mag_g = -1;
        if (mag && a.mag_antenna >= 0) {
            mag_g = a.mag_antenna - (int)grp * C::G;
            if (mag_g < 0 || mag_g >= C::G) mag = nullptr;
            mag_base = (long long)frame * (C::S * C::L);
        }
        rd_stream_range_s1<C>(tid, tile, a.rd_out, out_base, mag, mag_base, mag_g);
    };
    const unsigned n_items = (unsigned)a.n_items;
    unsigned item = blockIdx.x;
    int buf = 0;
    if constexpr (C::DF > 1) {
        // Class 0 is loaded at the top of the item, class 1 as soon as class 0 is converted (it lands while class 0 is
        // transformed); what travels one item ahead is an L2 prefetch, which costs no registers - holding next-item words
        // across phase B made ptxas spill them right after their load, i.e. wait for them (long-scoreboard stall on an STL).
        if (item < n_items) {
            rd_stream_prefetch_half<C, 0>(tid, rd_stream_item_src<C>(a, item));
            rd_stream_prefetch_half<C, 1>(tid, rd_stream_item_src<C>(a, item));
        }
        __syncthreads();
        for (; item < n_items; item += gridDim.x) {
            cf* tile = tiles + buf * (C::G * C::TILE);
            uint32_t w[C::DL1];
            cf v[C::DL1];
            const uint32_t* src = rd_stream_item_src<C>(a, item);
            rd_stream_load_half<C, 0>(tid, src, w);
            const unsigned next = item + gridDim.x;
            if (next < n_items) {
                const uint32_t* nsrc = rd_stream_item_src<C>(a, next);
                rd_stream_prefetch_half<C, 0>(tid, nsrc);
                rd_stream_prefetch_half<C, 1>(tid, nsrc);
            }
            rd_stream_convert_half<C, 0>(a, tid, w, v);
            asm volatile("" ::: "memory");                  // class 1 is requested only now, into the registers class 0 just freed
            rd_stream_load_half<C, 1>(tid, src, w);
            rd_stream_transform_half<C, 0>(tid, v, tile, winS_s);
            rd_stream_convert_half<C, 1>(a, tid, w, v);
            rd_stream_transform_half<C, 1>(tid, v, tile, winS_s);
            __syncthreads();
            rd_stream_range_s0_fused<C>(a, tid, tile);
            __syncthreads();
            finish_item(item, tile);
            if constexpr (C::NBUF == 2) buf ^= 1; else __syncthreads();
        }
    } else {
        uint32_t w[C::RAW_WORDS];
        if (item < n_items) rd_stream_load_raw<C>(tid, rd_stream_item_src<C>(a, item), w);
        __syncthreads();
        for (; item < n_items; item += gridDim.x) {
            cf* tile = tiles + buf * (C::G * C::TILE);
            rd_stream_doppler_s0<C>(a, tid, w, tile, winS_s);
            // the raw words of the NEXT item travel while this one goes through the remaining stages
            const unsigned next = item + gridDim.x;
            if (next < n_items) rd_stream_load_raw<C>(tid, rd_stream_item_src<C>(a, next), w);
            __syncthreads();
            if constexpr (C::DL2 > 1) {
                rd_stream_doppler_s1<C>(tid, tile, winS_s);
                __syncthreads();
            }
            rd_stream_range_s0<C>(a, tid, tile);
            __syncthreads();
            finish_item(item, tile);
            // NBUF == 2: no barrier here - the next item is built in the other tile buffer, and the barriers of that item
            // order its successor's writes after every read of this one
            if constexpr (C::NBUF == 2) buf ^= 1; else __syncthreads();
        }
    }
}

static std::vector<float> host_hann(int n, bool on) {
    std::vector<float> w((size_t)n, 1.0f);
    if (on && n > 1)
        for (int k = 0; k < n; ++k) w[(size_t)k] = (float)(0.5 - 0.5 * std::cos(2.0 * M_PI * (double)k / (double)(n - 1)));
    return w;
}

template <class C>
static int launch_stream(radar_ctx* ctx, const void* adc, int n_frames, int window_flags, cf* rd_out, float* mag_out,
                         int mag_antenna, cudaStream_t st) {
    constexpr size_t smem = C::smem_bytes();
    static_assert(smem <= 227 * 1024, "tiles exceed shared memory");
    static_assert(sizeof(RdStreamParams<C>) <= 30000, "parameter block too large (CUDA 12.1+: 32 764 bytes)");
    RdStreamParams<C> a;
    a.iq = reinterpret_cast<const uint32_t*>(adc);
    a.rd_out = rd_out; a.mag_out = mag_out; a.mag_antenna = mag_antenna;
    a.R = ctx->shape.num_rx; a.cpl = ctx->shape.chirps_per_loop; a.V = ctx->V;
    a.items_per_frame = ctx->V / C::G;
    if ((long long)n_frames * a.items_per_frame > 0x7fffffffLL) return set_error(ctx, RADAR_E_INVALID, "too many frames in one call");
    a.n_items = (long long)n_frames * a.items_per_frame;
    const std::vector<float> wl = host_hann(C::L, window_flags & RADAR_WIN_DOPPLER);
    const std::vector<float> ws = host_hann(C::S, window_flags & RADAR_WIN_RANGE);
    for (int i = 0; i < C::L; ++i) a.winL[i] = wl[(size_t)i];
    for (int i = 0; i < C::S; ++i) a.winS[i] = ws[(size_t)i];
    for (int n2 = 0; n2 < C::RS2; ++n2)
        for (int k = 0; k < C::RS1; ++k) {
            const double ang = -2.0 * M_PI * (double)((n2 * k) % C::S) / (double)C::S;
            a.twS[n2 * C::RS1 + k] = cmake((float)std::cos(ang), (float)std::sin(ang));
        }
    for (int n2 = 0; n2 < C::DL2; ++n2)
        for (int k = 0; k < C::DL1; ++k) {
            const double ang = -2.0 * M_PI * (double)((n2 * k) % C::L) / (double)C::L;
            a.twD[n2 * C::DL1 + k] = cmake((float)std::cos(ang), (float)std::sin(ang));
        }
    static int resident_by_dev[64] = {0};         // CTAs per SM for this instantiation (queried once per device)
    int& resident = resident_by_dev[ctx->device & 63];
    if (!resident) {
        if (smem > 48 * 1024)
            RB_CUDA_OK(ctx, cudaFuncSetAttribute(rd_stream_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(rd_stream_kernel<C>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        int occ = 0;
        RB_CUDA_OK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rd_stream_kernel<C>, C::NT, smem));
        if (occ < 1) return set_error(ctx, RADAR_E_UNSUPPORTED, "rd_stream_kernel does not fit on this device");
        resident = occ;
    }
    long long grid = (long long)ctx->sm_count * resident;
    if (grid > a.n_items) grid = a.n_items;
    rd_stream_kernel<C><<<(unsigned)grid, C::NT, smem, st>>>(a);
    ctx->last_rd_kernel = "rd_stream_kernel";
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// Returns RADAR_OK when a streaming kernel was launched, 1 when the shape has no instantiation (caller falls back).
int launch_range_doppler_stream(radar_ctx* ctx, const void* adc, int n_frames, int window_flags, cf* rd_out,
                                float* mag_out, int mag_antenna, cudaStream_t st) {
    const int S = ctx->shape.num_samples, L = ctx->shape.num_loops, R = ctx->shape.num_rx;
    if (reinterpret_cast<uintptr_t>(adc) % 4 != 0) return 1;
    const int CR = ctx->shape.chirps_per_loop * R;
    const int variant = ctx->stream_variant;
#define RB_STREAM(var, s, l, cr, rs1, rs2, dl1, dl2, g, nt, minb, df)                                       \
    if (variant == var && S == s && L == l && CR == cr && R % g == 0)                                       \
        return launch_stream<RdStreamShape<s, l, cr, rs1, rs2, dl1, dl2, g, nt, minb, df>>(ctx, adc, n_frames, window_flags, \
                                                                                          rd_out, mag_out, mag_antenna, st);
#include "rd_stream_shapes.inc"
#undef RB_STREAM
    return 1;
}

}  // namespace rb
