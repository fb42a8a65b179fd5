// Register-resident range-Doppler kernel: warp-specialised device wrapper and launcher for the phases in rd_reg.cuh.
//
// One persistent CTA per SM.  Item i = antenna group (i mod IPF) of frame (i div IPF); a group is G consecutive RX
// antennas of one TX slot, whose raw words form one contiguous run of G*S words in every chirp row.
//   TMA warp : per item L bulk copies (one per loop, 16-byte aligned window around the run) into stage j mod NSTAGE of the
//              raw ring; waits on raw_empty, completes on raw_full (expect_tx).
//   A warps  : one thread per (antenna, sample) column: ring -> registers (stage released as soon as the words are read),
//              L-point Doppler transform, row of the tile; waits on tile_empty, arrives on tile_full.
//   B warps  : one thread per (antenna, Doppler bin) column: tile -> registers (tile released at once), S-point range
//              transform, stores to rd_out / the magnitude map.
#include <cmath>
#include "radar_internal.h"
#include "rd_reg.cuh"
#include "mbar.cuh"

namespace rb {

template <class C>
__global__ void __launch_bounds__(C::NT, 1) rd_reg_kernel(const __grid_constant__ RdRegParams<C> a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cf* tiles = reinterpret_cast<cf*>(smem_raw);
    uint32_t* ring = reinterpret_cast<uint32_t*>(smem_raw + (size_t)C::NTILE * C::TILE * sizeof(cf));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)C::NTILE * C::TILE * sizeof(cf) + (size_t)C::NSTAGE * C::RAW_STAGE_W * 4);
    uint64_t* raw_full = bars;
    uint64_t* raw_empty = bars + C::NSTAGE;
    uint64_t* tile_full = bars + 2 * C::NSTAGE;
    uint64_t* tile_empty = tile_full + C::NTILE;
    static_assert(2 * C::NSTAGE + 2 * C::NTILE <= 16, "barrier block");
    constexpr unsigned IPF = C::CR / C::G;
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // tells the compiler the role branches are warp uniform
    if (tid == 0) {
        for (int s = 0; s < C::NSTAGE; ++s) {
            mb::init(raw_full + s, 1);
            mb::init(raw_empty + s, C::NA * 32);
        }
        for (int t = 0; t < C::NTILE; ++t) {
            mb::init(tile_full + t, C::NA * 32);
            mb::init(tile_empty + t, C::NB * 32);
        }
        mb::fence_init();
    }
    __syncthreads();
    const unsigned n_items = a.n_items;

    if (warp == C::NA + C::NB) {
        // ===== TMA warp: one lane issues the L row copies of an item =====
        if (lane == 0) {
            unsigned j = 0;
            for (unsigned item = blockIdx.x; item < n_items; item += gridDim.x, ++j) {
                const unsigned st = j % C::NSTAGE, round = j / C::NSTAGE;
                mb::wait(raw_empty + st, (round & 1) ^ 1);
                const unsigned frame = item / IPF, grp = item - frame * IPF;
                const unsigned off = grp * (C::G * C::S * 4);                      // byte offset of the run inside a chirp row
                const unsigned a0 = off & ~15u;
                const unsigned bytes = ((off - a0) + C::G * C::S * 4 + 15u) & ~15u;
                mb::arrive_expect_tx(raw_full + st, bytes * C::L);
                const unsigned char* src = reinterpret_cast<const unsigned char*>(a.iq) + (size_t)frame * ((size_t)C::ROW * C::L * 4) + a0;
                uint32_t* dst = ring + st * C::RAW_STAGE_W;
#pragma unroll 5
                for (int l = 0; l < C::L; ++l)
                    mb::bulk_g2s(dst + l * C::RAW_PITCH_W, src + (size_t)l * (C::ROW * 4), bytes, raw_full + st);
            }
        }
    } else if (warp < C::NA) {
        // ===== phase A: Doppler =====
        const int h = C::DF == 1 ? 0 : warp / C::NA1;
        const int col = (warp - h * C::NA1) * 32 + lane;
        const bool active = col < C::COLS_A;
        const int colc = active ? col : C::COLS_A - 1;
        const float ws = a.winS[colc % C::S];
        unsigned j = 0;
        for (unsigned item = blockIdx.x; item < n_items; item += gridDim.x, ++j) {
            const unsigned st = j % C::NSTAGE, round = j / C::NSTAGE;
            const unsigned tb = j % C::NTILE, tround = j / C::NTILE;
            const unsigned grp = item % IPF;
            const int shift = (int)((grp * (C::G * C::S * 4)) & 15u) >> 2;
            cf v[C::LH];
            mb::wait(raw_full + st, round & 1);
            rd_reg_doppler_load<C>(a, ring + st * C::RAW_STAGE_W, shift + colc, h, v);
            mb::arrive(raw_empty + st);
            rd_reg_doppler_finish<C>(ws, v);
            rd_reg_doppler_twiddle<C>(a, h, v);
            mb::wait(tile_empty + tb, (tround & 1) ^ 1);
            if (active) rd_reg_store_row<C>(v, tiles + (size_t)tb * C::TILE + (size_t)col * C::P, h);
            mb::arrive(tile_full + tb);
        }
    } else {
        // ===== phase B: range =====
        const int wb = warp - C::NA;
        const int cls = wb / C::NB1;                                  // output class k_BF of this warp (warp uniform)
        const int ca = C::BF == 4 ? cls >> 1 : (C::BF == 2 ? cls : 0), odd = C::BF == 4 ? cls & 1 : 0;
        const float sigma = ca ? -1.0f : 1.0f;
        const int col = (wb - cls * C::NB1) * 32 + lane;
        const bool active = col < C::COLS_B;
        const int colc = active ? col : C::COLS_B - 1;
        const int g = colc / C::L, d = colc - g * C::L;
        const long long half = (long long)ca * (C::S / 2) * C::L;    // a*S/2 rows of the output
        unsigned j = 0;
        for (unsigned item = blockIdx.x; item < n_items; item += gridDim.x, ++j) {
            const unsigned tb = j % C::NTILE, tround = j / C::NTILE;
            const unsigned frame = item / IPF, grp = item - frame * IPF;
            const int ant = (int)grp * C::G + g;
            cf v[C::SH];
            mb::wait(tile_full + tb, tround & 1);
            if (odd) rd_reg_range_load<C, 1>(a, tiles + (size_t)tb * C::TILE, g, d, sigma, v);
            else rd_reg_range_load<C, 0>(a, tiles + (size_t)tb * C::TILE, g, d, sigma, v);
            mb::arrive(tile_empty + tb);
            const size_t o = ((size_t)frame * C::CR + ant) * (C::S * C::L) + d;
            cf* out = (a.rd_out && active) ? a.rd_out + o : nullptr;
            float* mag = nullptr;
            if (a.mag_out && active) {
                if (a.mag_antenna < 0) mag = a.mag_out + o;
                else if (a.mag_antenna == ant) mag = a.mag_out + (size_t)frame * (C::S * C::L) + d;
            }
            if (odd) rd_reg_range_store<C, 1>(v, out ? out + half : nullptr, out ? out - half : nullptr, mag ? mag + half : nullptr, mag ? mag - half : nullptr);
            else rd_reg_range_store<C, 0>(v, out ? out + half : nullptr, out ? out - half : nullptr, mag ? mag + half : nullptr, mag ? mag - half : nullptr);
        }
    }
}

// ---- Doppler beam sharpening without the range-Doppler cube ---------------------------------------------------------
//   out[f][k][i] = | sum_j w_j RD[f][v_j][k][vel_bin[i]] W_A^(j (ang_bin[i] - A/2)) |     (processors/range_angle_resp_dbs_enhanced.py:137-304)
// The range transform is linear, so the antennas are combined BEFORE it: Z[s][i] = sum_j w_ji Y_j[s][vel_bin[i]] with Y_j the
// windowed, Doppler-transformed samples phase A leaves in the tile, and one range transform per output angle (n_out of
// them) replaces one per antenna and Doppler column (CR * L).  The same three roles as rd_reg_kernel; a CTA owns whole frames
// (frame = blockIdx.x, += gridDim.x) and walks their antenna groups in order, so Z (S x n_out complex values) stays in its
// shared memory.  Phase B, per item: every thread adds its cells of Z (lanes along the output angles: neighbouring angles read
// the same or neighbouring tile columns, Z is contiguous; the first antenna of the subset stores, so nothing is zeroed); after the last antenna of the
// frame, between two barriers of the phase-B warps, one thread per (output angle, output class) transforms its column of Z in
// registers and writes the magnitudes.
template <class C, int ODD>
__device__ __forceinline__ void rd_reg_dbs_range(const cf* zcol, int n_out, int ca, float* out_col) {
    cf v[C::SH];
    rd_reg_range_load_plain<C, ODD>(zcol, n_out, ca ? -1.0f : 1.0f, v);
    Dft<C::SH>::run(v);
    const int half_rows = ca * (C::S / 2);
#pragma unroll
    for (int kh = 0; kh < C::SH; ++kh) {
        const int r = (ODD * C::E_BF + C::E_SH * kh) % C::S;
        const int k = (C::BF > 1 && r >= C::S / 2) ? r - half_rows : r + half_rows;
        out_col[(size_t)k * n_out] = sqrtf(v[kh].x * v[kh].x + v[kh].y * v[kh].y);
    }
}

constexpr int kDbsSlots = 4;             // output angles per lane: n_out <= 128
template <class C>
__global__ void __launch_bounds__(C::NT, 1) rd_reg_dbs_kernel(const __grid_constant__ RdRegParams<C> a, const __grid_constant__ RdRegDbs b) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cf* tiles = reinterpret_cast<cf*>(smem_raw);
    uint32_t* ring = reinterpret_cast<uint32_t*>(smem_raw + (size_t)C::NTILE * C::TILE * sizeof(cf));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)C::NTILE * C::TILE * sizeof(cf) + (size_t)C::NSTAGE * C::RAW_STAGE_W * 4);
    uint64_t* raw_full = bars;
    uint64_t* raw_empty = bars + C::NSTAGE;
    uint64_t* tile_full = bars + 2 * C::NSTAGE;
    uint64_t* tile_empty = tile_full + C::NTILE;
    cf* Z = reinterpret_cast<cf*>(smem_raw + C::smem_bytes());                    // [S][n_out] beamformed samples of the frame
    cf* wtab = Z + (size_t)C::S * b.n_out;                                          // [CR][n_out]: antenna weight x steering coefficient of a virtual antenna
    int* colL = reinterpret_cast<int*>(wtab + (size_t)C::CR * b.n_out);            // [n_out] Doppler column of output angle i
    int* jsub = colL + b.n_out;                                                     // [CR] antenna is in the subset; then first / last antenna
    constexpr unsigned IPF = C::CR / C::G;
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    if (tid == 0) {
        for (int s = 0; s < C::NSTAGE; ++s) {
            mb::init(raw_full + s, 1);
            mb::init(raw_empty + s, C::NA * 32);
        }
        for (int t = 0; t < C::NTILE; ++t) {
            mb::init(tile_full + t, C::NA * 32);
            mb::init(tile_empty + t, C::NB * 32);
        }
        mb::fence_init();
    }
    // tables of the call
    for (int v = tid; v < C::CR; v += C::NT) jsub[v] = -1;
    for (int i = tid; i < b.n_out; i += C::NT) {
        const int d = b.vel_bin[i];
        colL[i] = d < 0 ? 0 : (d >= C::L ? C::L - 1 : d);
    }
    // the coefficient of virtual antenna v for output angle i: position j of v in the subset sets the steering phase; a
    // list that names an antenna twice (or out of order) gets the sum of its positions' coefficients - the same image
    for (int e = tid; e < C::CR * b.n_out; e += C::NT) {
        const int v = e / b.n_out, i = e - v * b.n_out;
        int f = b.ang_bin[i] - b.A / 2;
        if (f < 0) f += b.A;
        cf acc = cmake(0.f, 0.f);
        for (int j = 0; j < b.N; ++j) {
            if ((b.ant_list ? b.ant_list[j] : j) != v) continue;
            const cf w = b.twA[(int)(((long long)j * f) % b.A)];
            const float wt = b.ant_weight ? b.ant_weight[j] : 1.0f;
            acc.x += w.x * wt; acc.y += w.y * wt;
        }
        wtab[e] = acc;
    }
    __syncthreads();
    if (tid == 0) {
        int lo = C::CR, hi = -1;
        for (int j = 0; j < b.N; ++j) {
            const int v = b.ant_list ? b.ant_list[j] : j;
            if (v < 0 || v >= C::CR) continue;
            jsub[v] = 1;
            lo = v < lo ? v : lo;
            hi = v > hi ? v : hi;
        }
        jsub[C::CR] = lo; jsub[C::CR + 1] = hi;
    }
    __syncthreads();
    const unsigned n_frames = b.n_frames;

    if (warp == C::NA + C::NB) {
        // ===== TMA warp =====
        if (lane == 0) {
            unsigned j = 0;
            for (unsigned frame = blockIdx.x; frame < n_frames; frame += gridDim.x)
                for (unsigned grp = 0; grp < IPF; ++grp, ++j) {
                    const unsigned st = j % C::NSTAGE, round = j / C::NSTAGE;
                    mb::wait(raw_empty + st, (round & 1) ^ 1);
                    const unsigned off = grp * (C::G * C::S * 4);
                    const unsigned a0 = off & ~15u;
                    const unsigned bytes = ((off - a0) + C::G * C::S * 4 + 15u) & ~15u;
                    mb::arrive_expect_tx(raw_full + st, bytes * C::L);
                    const unsigned char* src = reinterpret_cast<const unsigned char*>(a.iq) + (size_t)frame * ((size_t)C::ROW * C::L * 4) + a0;
                    uint32_t* dst = ring + st * C::RAW_STAGE_W;
#pragma unroll 5
                    for (int l = 0; l < C::L; ++l)
                        mb::bulk_g2s(dst + l * C::RAW_PITCH_W, src + (size_t)l * (C::ROW * 4), bytes, raw_full + st);
                }
        }
    } else if (warp < C::NA) {
        // ===== phase A: Doppler (as in rd_reg_kernel) =====
        const int h = C::DF == 1 ? 0 : warp / C::NA1;
        const int col = (warp - h * C::NA1) * 32 + lane;
        const bool active = col < C::COLS_A;
        const int colc = active ? col : C::COLS_A - 1;
        const float ws = a.winS[colc % C::S];
        unsigned j = 0;
        for (unsigned frame = blockIdx.x; frame < n_frames; frame += gridDim.x)
            for (unsigned grp = 0; grp < IPF; ++grp, ++j) {
                const unsigned st = j % C::NSTAGE, round = j / C::NSTAGE;
                const unsigned tb = j % C::NTILE, tround = j / C::NTILE;
                const int shift = (int)((grp * (C::G * C::S * 4)) & 15u) >> 2;
                cf v[C::LH];
                mb::wait(raw_full + st, round & 1);
                rd_reg_doppler_load<C>(a, ring + st * C::RAW_STAGE_W, shift + colc, h, v);
                mb::arrive(raw_empty + st);
                rd_reg_doppler_finish<C>(ws, v);
                rd_reg_doppler_twiddle<C>(a, h, v);
                mb::wait(tile_empty + tb, (tround & 1) ^ 1);
                if (active) rd_reg_store_row<C>(v, tiles + (size_t)tb * C::TILE + (size_t)col * C::P, h);
                mb::arrive(tile_full + tb);
            }
    } else {
        // ===== phase B: beamform per output angle, range transform after the frame's last antenna =====
        // warp w takes the samples s = w, w + NB, ...; its lanes take the output angles i = lane + 32 q: per angle the tile
        // position, the Doppler combine (see RdRegColumn) and the steering coefficient sit in registers, the kDbsSlots cells
        // of a sample are independent (instruction-level parallelism for the shared-memory read-modify-writes)
        constexpr int nB = C::NB * 32;
        const int tB = tid - C::NA * 32, wB = tB >> 5;
        const int n_out = b.n_out;
        const int ant_first = jsub[C::CR], ant_last = jsub[C::CR + 1];
        int pos[kDbsSlots];
        cf cw[kDbsSlots];
#pragma unroll
        for (int q = 0; q < kDbsSlots; ++q) {
            const int i = lane + 32 * q;
            const int d = i < n_out ? colL[i] : 0;
            pos[q] = d; cw[q] = cmake(1.0f, 0.0f);
            if (C::DF == 2) {
                const int k = (d + C::L - C::L / 2) % C::L;
                pos[q] = k % C::LH;
                cw[q] = cmake((C::TW ? k >= C::LH : (k & 1)) ? -1.0f : 1.0f, 0.0f);
            }
        }
        unsigned j = 0;
        for (unsigned frame = blockIdx.x; frame < n_frames; frame += gridDim.x)
            for (unsigned grp = 0; grp < IPF; ++grp, ++j) {
                const unsigned tb = j % C::NTILE, tround = j / C::NTILE;
                const cf* tile = tiles + (size_t)tb * C::TILE;
                mb::wait(tile_full + tb, tround & 1);
#pragma unroll
                for (int g = 0; g < C::G; ++g) {
                    const int ant = (int)grp * C::G + g;
                    const int js = jsub[ant];
                    if (js < 0) continue;                                   // uniform over the CTA
                    const bool first = ant == ant_first;
                    cf w[kDbsSlots];
#pragma unroll
                    for (int q = 0; q < kDbsSlots; ++q) w[q] = lane + 32 * q < n_out ? wtab[(size_t)ant * n_out + lane + 32 * q] : cmake(0.f, 0.f);
                    for (int s = wB; s < C::S; s += C::NB) {
                        const cf* row = tile + ((size_t)g * C::S + s) * C::P;
                        cf* zrow = Z + (size_t)s * n_out + lane;
#pragma unroll
                        for (int q = 0; q < kDbsSlots; ++q) {
                            if (lane + 32 * q < n_out) {
                                cf x = row[pos[q]];
                                if (C::DF == 2) {                           // the Doppler combine of RdRegColumn::value
                                    const cf o = row[pos[q] + C::HP];
                                    x = cfma(cw[q].x, o, x);
                                }
                                cf c = first ? cmake(0.f, 0.f) : zrow[32 * q];
                                c.x = fmaf(x.x, w[q].x, fmaf(-x.y, w[q].y, c.x));
                                c.y = fmaf(x.x, w[q].y, fmaf(x.y, w[q].x, c.y));
                                zrow[32 * q] = c;
                            }
                        }
                    }
                }
                mb::arrive(tile_empty + tb);
                if ((int)grp == ant_last / C::G) {
                    asm volatile("bar.sync 1, %0;" ::"n"(nB) : "memory");          // every cell of the frame is complete
                    float* out_frame = b.out + (size_t)frame * C::S * n_out;
                    for (int t = tB; t < n_out * C::BF; t += nB) {
                        const int cls = t / n_out, i = t - cls * n_out;
                        const int ca = C::BF == 4 ? cls >> 1 : (C::BF == 2 ? cls : 0), odd = C::BF == 4 ? cls & 1 : 0;
                        if (odd) rd_reg_dbs_range<C, 1>(Z + i, n_out, ca, out_frame + i);
                        else rd_reg_dbs_range<C, 0>(Z + i, n_out, ca, out_frame + i);
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(nB) : "memory");          // Z may be overwritten by the next frame
                }
            }
    }
}

static void host_hann_f(int n, bool on, float* w) {
    for (int k = 0; k < n; ++k)
        w[k] = (on && n > 1) ? (float)(0.5 - 0.5 * std::cos(2.0 * M_PI * (double)k / (double)(n - 1))) : 1.0f;
}

template <class C>
static int launch_reg(radar_ctx* ctx, const void* adc, int n_frames, int window_flags, cf* rd_out, float* mag_out,
                      int mag_antenna, cudaStream_t st) {
    constexpr size_t smem = C::smem_bytes();
    static_assert(smem <= 227 * 1024, "tile + ring exceed shared memory");
    static_assert(sizeof(RdRegParams<C>) <= 4000, "parameter block too large");
    static_assert((C::ROW * 4) % 16 == 0, "chirp rows must keep the 16-byte alignment of the bulk copies");
    static_assert(C::CR % C::G == 0, "antenna groups must tile the virtual array");
    RdRegParams<C> a;
    a.iq = reinterpret_cast<const uint32_t*>(adc);
    a.rd_out = rd_out; a.mag_out = mag_out; a.mag_antenna = mag_antenna;
    constexpr int IPF = C::CR / C::G;
    if ((long long)n_frames * IPF > 0x7fffffffLL) return set_error(ctx, RADAR_E_INVALID, "too many frames in one call");
    a.n_items = (unsigned)((long long)n_frames * IPF);
    float wl[C::L];
    host_hann_f(C::L, window_flags & RADAR_WIN_DOPPLER, wl);
    for (int h = 0; h < C::DF; ++h)
        for (int m = 0; m < C::LH; ++m) {
            const float w = wl[rd_reg_doppler_input<C>(h, m)];
            a.winL[h * C::LH + m] = make_float2(w, w);
        }
    if (C::TW)
        for (int k = 0; k < C::L; ++k) {
            const double ang = -2.0 * M_PI * (double)k / (double)C::L;
            a.twL[k] = cmake((float)std::cos(ang), (float)std::sin(ang));
        }
    host_hann_f(C::S, window_flags & RADAR_WIN_RANGE, a.winS);
    static bool configured[64] = {false};
    if (!configured[ctx->device & 63]) {
        RB_CUDA_OK(ctx, cudaFuncSetAttribute(rd_reg_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[ctx->device & 63] = true;
    }
    unsigned grid = (unsigned)ctx->sm_count;
    if (grid > a.n_items) grid = a.n_items;
    rd_reg_kernel<C><<<grid, C::NT, smem, st>>>(a);
    ctx->last_rd_kernel = "rd_reg_kernel";
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

template <class C>
static int launch_reg_dbs(radar_ctx* ctx, const void* adc, int n_frames, const RdRegDbs& b0, cudaStream_t st) {
    static_assert(sizeof(RdRegParams<C>) + sizeof(RdRegDbs) <= 4000, "parameter block too large");
    static_assert((C::ROW * 4) % 16 == 0, "chirp rows must keep the 16-byte alignment of the bulk copies");
    static_assert(C::CR % C::G == 0, "antenna groups must tile the virtual array");
    const size_t extra = ((size_t)C::S * b0.n_out + (size_t)C::CR * b0.n_out) * sizeof(cf) + ((size_t)b0.n_out + C::CR + 2) * sizeof(int);
    const size_t smem = C::smem_bytes() + ((extra + 15) & ~(size_t)15);
    if (smem > (size_t)ctx->max_smem_optin || b0.n_out > 32 * kDbsSlots) return 1;
    RdRegParams<C> a;
    a.iq = reinterpret_cast<const uint32_t*>(adc);
    a.rd_out = nullptr; a.mag_out = nullptr; a.mag_antenna = 0; a.n_items = 0;
    float wl[C::L];
    host_hann_f(C::L, true, wl);
    for (int h = 0; h < C::DF; ++h)
        for (int m = 0; m < C::LH; ++m) {
            const float w = wl[rd_reg_doppler_input<C>(h, m)];
            a.winL[h * C::LH + m] = make_float2(w, w);
        }
    if (C::TW)
        for (int k = 0; k < C::L; ++k) {
            const double ang = -2.0 * M_PI * (double)k / (double)C::L;
            a.twL[k] = cmake((float)std::cos(ang), (float)std::sin(ang));
        }
    host_hann_f(C::S, true, a.winS);
    RdRegDbs b = b0;
    b.n_frames = (unsigned)n_frames;
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(rd_reg_dbs_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    unsigned grid = (unsigned)ctx->sm_count;
    if (grid > (unsigned)n_frames) grid = (unsigned)n_frames;
    rd_reg_dbs_kernel<C><<<grid, C::NT, smem, st>>>(a, b);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

// Doppler beam sharpening fused into the register-resident kernel: RADAR_OK when launched, 1 when the shape has no
// instantiation or the accumulator does not fit beside the tile and the ring (caller falls back).
int launch_dbs_reg(radar_ctx* ctx, const void* adc, int n_frames, const int32_t* ant_list, int N, int num_angle_bins,
                   const int32_t* vel_bin, const int32_t* ang_bin, int n_out, const cf* twA, const float* ant_weight, float* out,
                   cudaStream_t st) {
    const int S = ctx->shape.num_samples, L = ctx->shape.num_loops, R = ctx->shape.num_rx;
    if (reinterpret_cast<uintptr_t>(adc) % 16 != 0) return 1;
    const int CR = ctx->shape.chirps_per_loop * R;
    RdRegDbs b{vel_bin, ang_bin, ant_list, ant_weight, twA, out, n_out, N, num_angle_bins, 0u};
#define RB_REG_DBS(s, l, cr, g, df, nstage, ntile, bf)                                                        \
    if (S == s && L == l && CR == cr && R % g == 0)                                                           \
        return launch_reg_dbs<RdRegShape<s, l, cr, g, df, nstage, ntile, bf>>(ctx, adc, n_frames, b, st);
    // two ring stages: the accumulator needs the room (ring depth 2 / 3 / 4 measured equal on the range-Doppler kernel)
    RB_REG_DBS(90, 100, 8, 1, 2, 2, 1, 2)
    RB_REG_DBS(63, 70, 12, 2, 1, 2, 1, 1)
    RB_REG_DBS(100, 30, 8, 2, 1, 2, 1, 4)
#undef RB_REG_DBS
    return 1;
}

// Returns RADAR_OK when the register-resident kernel was launched, 1 when the shape has no instantiation (caller falls back).
int launch_range_doppler_reg(radar_ctx* ctx, const void* adc, int n_frames, int window_flags, cf* rd_out,
                             float* mag_out, int mag_antenna, cudaStream_t st) {
    const int S = ctx->shape.num_samples, L = ctx->shape.num_loops, R = ctx->shape.num_rx;
    if (reinterpret_cast<uintptr_t>(adc) % 16 != 0) return 1;
    const int CR = ctx->shape.chirps_per_loop * R;
    const int variant = ctx->rd_reg;
#define RB_REG(var, s, l, cr, g, df, nstage, ntile, bf)                                                       \
    if (variant == var && S == s && L == l && CR == cr && R % g == 0)                                         \
        return launch_reg<RdRegShape<s, l, cr, g, df, nstage, ntile, bf>>(ctx, adc, n_frames, window_flags, rd_out, mag_out, mag_antenna, st);
    // RB_REG(variant (RADAR_B200_RD_REG; 1 is the default), S, L, CR, G, DF, ring stages, tile buffers, BF)
    RB_REG(1, 63, 70, 12, 2, 1, 3, 1, 1)      // 6843 ODS: 0.547 ms per 4096 frames (0.75 of measured HBM), 166 registers, 10 warps
    RB_REG(4, 63, 70, 12, 2, 1, 2, 2, 1)      //   two tile buffers, two ring stages: 0.530
    RB_REG(5, 63, 70, 12, 2, 1, 4, 1, 1)      //   four ring stages
    RB_REG(2, 63, 70, 12, 2, 2, 3, 1, 1)      //   split column: 14 warps are allocated as 16 -> 128 registers, phase B spills: 0.886
    RB_REG(3, 63, 70, 12, 2, 2, 2, 2, 1)      //   the same with two tile buffers: 0.976
    RB_REG(1, 100, 30, 8, 2, 1, 3, 1, 4)     // 1843_RadVel: 30-point Doppler per thread, 100-point range split 4 x 25
    RB_REG(2, 100, 30, 8, 2, 1, 2, 2, 4)
    RB_REG(1, 64, 40, 4, 2, 1, 3, 1, 1)      // RadCloud: 40-point Doppler, 64-point range (8 x 8) per thread
    RB_REG(1, 90, 100, 8, 1, 2, 3, 1, 2)     // RadSAR_1843_official: Doppler in even / odd halves of 50 folded with the column's twiddle, range 2 x 45
#undef RB_REG
    return 1;
}

}  // namespace rb
