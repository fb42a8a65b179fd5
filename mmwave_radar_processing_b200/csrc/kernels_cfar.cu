// CFAR family + ordered detection compaction: one CTA per map, the map resident in shared memory.
//
// Reference arithmetic reproduced (paths relative to mmwave_radar_processing/ in the reference):
//   detectors/base.py:38-65, :208-230 (strict x > threshold, np.where order),
//   detectors/ca_cfar.py:11-155, detectors/os_cfar.py:11-195, detectors/go_so_cfar.py:11-123.
//
// * CA / GO / SO: one thread per cell under test, training sums accumulated in fp64.
// * OS, mask only (the streaming path): one thread per cell under test counts the training cells below a per-cell bound
//   and stops as soon as the decision is fixed (see os_count_cells).
// * OS with noise estimates requested: one warp per cell under test; every lane keeps its share of the training window in registers
//   and the exact k-th order statistic is built bit by bit (MSB->LSB) from warp-wide counts
//   (__reduce_add_sync), which is an exact selection on the fp32 bit patterns (non-negative floats order
//   like their unsigned encodings).
// * decision in fp64: (double)x > alpha*(double)noise - the same expression the host layer uses to publish
//   `thresholds`.
// * compaction: warp ballots + popcounts, an exclusive scan over the warps' totals, then an ordered
//   scatter - row-major (range, Doppler) like np.where, one global counter write per map.
#include "radar_internal.h"
#include "rd_core.cuh"

namespace rb {

constexpr int kCfarThreads = 256;
constexpr int kCfarWarps = kCfarThreads / 32;

struct CfarArgs {
    const float* maps;
    int rows, cols;
    int kind, tr, td, gr, gd;
    int n_train;          // cells entering the estimate (CA/OS: all; GO/SO: one side)
    int k_rank;           // OS: 1-based rank
    double alpha;
    const uint8_t* row_select;
    double* noise_out;
    uint8_t* mask_out;
    int32_t* dets_out;
    int32_t* count_out;
    int max_dets;
    int pitch;            // row pitch of the map in shared memory: == nvd (mod 32) so consecutive cells under test, which run
                          // across valid rows, read consecutive banks (>= cols)
    FastDiv div_cols, div_vd, div_wd;
    int ph1_extra;        // OS counting path: full rows counted before the first decision, beyond the minimum that can reject
};

template <int MAXV>
__device__ __forceinline__ void os_cells(const CfarArgs& a, const float* map, uint8_t* mask, double* noise_dst) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cr = a.tr + a.gr, cd = a.td + a.gd;
    const int wr = 2 * cr + 1, wd = 2 * cd + 1;
    const int nvr = a.rows - 2 * cr, nvd = a.cols - 2 * cd;
    // per-lane window offsets, identical for every cell under test
    int off[MAXV];
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int w = lane + 32 * i;
        off[i] = -1;
        if (w < wr * wd) {
            const int y = (int)a.div_wd.div((uint32_t)w);
            const int x = w - y * wd;
            const bool guard = (y >= a.tr && y <= a.tr + 2 * a.gr) && (x >= a.td && x <= a.td + 2 * a.gd);
            if (!guard) off[i] = y * a.pitch + x;
        }
    }
    const int n_cut = nvr * nvd;
    for (int c = warp; c < n_cut; c += kCfarWarps) {
        const int vr = (int)a.div_vd.div((uint32_t)c);
        const int vd = c - vr * nvd;
        const int r = vr + cr, d = vd + cd;
        if (a.row_select && !a.row_select[r]) continue;
        const int base = vr * a.pitch + vd;     // top-left corner of the window
        unsigned vals[MAXV];
#pragma unroll
        for (int i = 0; i < MAXV; ++i) vals[i] = off[i] >= 0 ? __float_as_uint(map[base + off[i]]) : 0xffffffffu;
        unsigned res = 0;
#pragma unroll 1
        for (int bit = 30; bit >= 0; --bit) {
            const unsigned cand = res | (1u << bit);
            int cnt = 0;
#pragma unroll
            for (int i = 0; i < MAXV; ++i) cnt += (vals[i] < cand) ? 1 : 0;
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (cnt < a.k_rank) res = cand;     // fewer than k cells below cand: the k-th smallest is >= cand
        }
        if (lane == 0) {
            const int cell = r * a.cols + d;
            const double noise = (double)__uint_as_float(res);
            mask[cell] = ((double)map[r * a.pitch + d] > a.alpha * noise) ? 1 : 0;
            if (noise_dst) noise_dst[cell] = noise;
        }
    }
}


// ---- OS-CFAR decision without the order statistic ---------------------------------------------------------------
// The detector only needs  x > alpha * v_k  (v_k = k-th smallest training cell).  Rounded multiplication by a positive
// constant is monotone, so that holds iff at least k training cells c satisfy  alpha*(double)c < (double)x.  For
// non-negative floats that predicate is a plain threshold  c < T  with T the smallest float for which it fails, found
// from x/alpha by stepping at most a couple of ulps.  The count needs no selection at all and can stop as soon as the
// outcome is fixed (count >= k: detection; count + cells still unseen < k: none) - for noise cells that is after about a
// third of the window.  The mask is identical to the one the exact order statistic produces (tests compare both paths).
__device__ __forceinline__ float os_compare_bound(float x, double alpha) {
    const double xd = (double)x;
    float q = (float)(xd / alpha);
    if (!(q >= 0.0f)) q = 0.0f;                                   // negative or NaN cell under test: nothing is below it
    while (q < __int_as_float(0x7f800000) && alpha * (double)q < xd) q = __uint_as_float(__float_as_uint(q) + 1u);
    while (q > 0.0f) {
        const float p = __uint_as_float(__float_as_uint(q) - 1u);
        if (alpha * (double)p < xd) break;
        q = p;
    }
    return q;
}

__device__ __forceinline__ int count_below(const float* __restrict__ row, int begin, int end, float t) {
    int n = 0;
#pragma unroll 4
    for (int i = begin; i < end; ++i) n += (row[i] < t) ? 1 : 0;
    return n;
}
// compile-time span: every load is base + immediate and the whole row is in flight at once
template <int BEGIN, int END>
__device__ __forceinline__ int count_below_fixed(const float* __restrict__ row, float t) {
    float v[END - BEGIN > 0 ? END - BEGIN : 1];
#pragma unroll
    for (int i = BEGIN; i < END; ++i) v[i - BEGIN] = row[i];
    int n = 0;
#pragma unroll
    for (int i = BEGIN; i < END; ++i)       // one compare + one predicated add per cell (the compiler's select form takes three)
        asm("{.reg .pred p; setp.lt.f32 p, %1, %2; @p add.s32 %0, %0, 1;}" : "+r"(n) : "f"(v[i - BEGIN]), "f"(t));
    return n;
}

// TD, GD > 0: Doppler-axis window compiled in (the common detector settings); TD == 0: run-time widths.
//
// Two phases against warp divergence (ncu: a warp ran ~19 window rows because one hard cell keeps 31 decided lanes
// waiting, while the average cell is decided after ~6): phase 1 counts the first ph1 full rows for EVERY cell with no
// exit test in between (uniform, all loads independent) and decides the easy cells (noise: too few cells below the bound
// are still possible); the undecided ones are appended to a shared-memory work list (warp-aggregated) with their count
// and bound, and phase 2 walks that list with dense warps, row by row with the early exit.
template <int TD, int GD>
__device__ __forceinline__ void os_count_cells(const CfarArgs& a, const float* map, uint8_t* mask, int* work_cell,
                                               float* work_t, int* work_n) {
    const int td = TD > 0 ? TD : a.td, gd = TD > 0 ? GD : a.gd;
    const int cr = a.tr + a.gr, cd = td + gd;
    const int wr = 2 * cr + 1, wd = 2 * cd + 1;
    const int nvr = a.rows - 2 * cr, nvd = a.cols - 2 * cd;
    const int n_cut = nvr * nvd;
    const int g_lo = a.tr, g_hi = a.tr + 2 * a.gr;                 // window rows crossing the guard band
    const int seg2 = td + 2 * gd + 1;
    // rows of phase 1: enough that "all of them above the bound" already rules a detection out, plus two, all full rows
    int ph1 = (a.n_train - a.k_rank + wd) / wd + a.ph1_extra;
    ph1 = ph1 < g_lo ? ph1 : g_lo;
    const int lane = threadIdx.x & 31;
    for (int c0 = 0; c0 < n_cut; c0 += blockDim.x) {
        const int c = c0 + threadIdx.x;
        bool undecided = false;
        int cnt = 0, cell = 0;
        float t = 0.0f;
        if (c < n_cut) {
            const int vr = (int)a.div_vd.div((uint32_t)c);
            const int vd = c - vr * nvd;
            const int r = vr + cr;
            if (!(a.row_select && !a.row_select[r])) {
                cell = r * a.cols + vd + cd;
                const float* row = map + vr * a.pitch + vd;           // top-left corner of the window
                t = os_compare_bound(row[cr * a.pitch + cd], a.alpha);
                for (int y = 0; y < ph1; ++y, row += a.pitch) {
                    if (TD > 0) cnt += count_below_fixed<0, 2 * (TD + GD) + 1>(row, t);
                    else cnt += count_below(row, 0, wd, t);
                }
                const int unseen = a.n_train - ph1 * wd;
                if (cnt >= a.k_rank) mask[cell] = 1;
                else if (cnt + unseen >= a.k_rank) undecided = true;  // else: no detection, mask stays 0
            }
        }
        const unsigned vote = __ballot_sync(0xffffffffu, undecided);
        if (vote) {
            int base = 0;
            if (lane == 0) base = atomicAdd(work_n, __popc(vote));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (undecided) {
                const int slot = base + __popc(vote & ((1u << lane) - 1u));
                work_cell[slot] = c | (cnt << 16);
                work_t[slot] = t;
            }
        }
    }
    __syncthreads();
    const int n_work = *work_n;
    for (int i = threadIdx.x; i < n_work; i += blockDim.x) {
        const int packed = work_cell[i];
        const int c = packed & 0xffff;
        int cnt = packed >> 16;
        const float t = work_t[i];
        const int vr = (int)a.div_vd.div((uint32_t)c);
        const int vd = c - vr * nvd;
        const int cell = (vr + cr) * a.cols + vd + cd;
        const float* row = map + (vr + ph1) * a.pitch + vd;
        int unseen = a.n_train - ph1 * wd;
        for (int y = ph1; y < wr; ++y, row += a.pitch) {
            if (y < g_lo || y > g_hi) {
                if (TD > 0) cnt += count_below_fixed<0, 2 * (TD + GD) + 1>(row, t);
                else cnt += count_below(row, 0, wd, t);
                unseen -= wd;
            } else {
                if (TD > 0) cnt += count_below_fixed<0, TD>(row, t) + count_below_fixed<TD + 2 * GD + 1, 2 * (TD + GD) + 1>(row, t);
                else cnt += count_below(row, 0, td, t) + count_below(row, seg2, wd, t);
                unseen -= 2 * td;
            }
            if (cnt >= a.k_rank || cnt + unseen < a.k_rank) break;
        }
        mask[cell] = cnt >= a.k_rank ? 1 : 0;
    }
}

__device__ __forceinline__ void os_count_dispatch(const CfarArgs& a, const float* map, uint8_t* mask, int* wc, float* wt, int* wn) {
    // Doppler-axis (train, guard) of the reference's detector settings (tests/verify_processors.py:121-143,
    // gui_configs/processor_params.yaml:41-88)
    if (a.td == 4 && a.gd == 1) os_count_cells<4, 1>(a, map, mask, wc, wt, wn);
    else if (a.td == 5 && a.gd == 2) os_count_cells<5, 2>(a, map, mask, wc, wt, wn);
    else if (a.td == 5 && a.gd == 3) os_count_cells<5, 3>(a, map, mask, wc, wt, wn);
    else if (a.td == 12 && a.gd == 4) os_count_cells<12, 4>(a, map, mask, wc, wt, wn);
    else os_count_cells<0, 0>(a, map, mask, wc, wt, wn);
}

__global__ void __launch_bounds__(kCfarThreads, 4) cfar_kernel(const __grid_constant__ CfarArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int warp_counts[kCfarWarps];
    const int n_cells = a.rows * a.cols;
    SmemArena ar(smem_raw);
    float* map = ar.take<float>(a.rows * a.pitch);
    uint8_t* mask = ar.take<uint8_t>(n_cells);
    const int n_work_max = max(a.rows - 2 * (a.tr + a.gr), 0) * max(a.cols - 2 * (a.td + a.gd), 0);   // cells under test
    int* work_cell = ar.take<int>(n_work_max);           // OS counting path: undecided cells after phase 1
    float* work_t = ar.take<float>(n_work_max);
    __shared__ int work_n;
    if (threadIdx.x == 0) work_n = 0;
    const long long m = blockIdx.x;
    const float* src = a.maps + m * n_cells;
    double* noise_dst = a.noise_out ? a.noise_out + m * n_cells : nullptr;
    const uint8_t* rsel = a.row_select ? a.row_select + m * a.rows : nullptr;
    CfarArgs la = a;
    la.row_select = rsel;
    for (int i = threadIdx.x; i < n_cells; i += blockDim.x) {
        { const int y = (int)a.div_cols.div((uint32_t)i); map[y * a.pitch + (i - y * a.cols)] = src[i]; }
        mask[i] = 0;
        if (noise_dst) noise_dst[i] = 0.0;
    }
    __syncthreads();

    const int cr = a.tr + a.gr, cd = a.td + a.gd;
    const int nvr = a.rows - 2 * cr, nvd = a.cols - 2 * cd;
    if (nvr > 0 && nvd > 0) {      // window larger than the map: no valid cell (reference swallows the ValueError)
        if (a.kind == RADAR_CFAR_OS && !noise_dst && a.alpha > 0.0) {
            os_count_dispatch(la, map, mask, work_cell, work_t, &work_n);
        } else if (a.kind == RADAR_CFAR_OS) {
            const int wcells = (2 * cr + 1) * (2 * cd + 1);
            if (wcells <= 32 * 8) os_cells<8>(la, map, mask, noise_dst);
            else if (wcells <= 32 * 16) os_cells<16>(la, map, mask, noise_dst);
            else os_cells<32>(la, map, mask, noise_dst);
        } else {
            const int n_cut = nvr * nvd;
            for (int c = threadIdx.x; c < n_cut; c += blockDim.x) {
                const int vr = (int)a.div_vd.div((uint32_t)c);
                const int vd = c - vr * nvd;
                const int r = vr + cr, d = vd + cd;
                if (rsel && !rsel[r]) continue;
                double noise;
                if (a.kind == RADAR_CFAR_CA) {
                    double sum = 0.0;
                    for (int y = r - cr; y <= r + cr; ++y) {
                        const float* row = map + y * a.pitch;
                        const bool inner = (y >= r - a.gr && y <= r + a.gr);
                        if (inner) {
                            for (int x = d - cd; x < d - a.gd; ++x) sum += (double)row[x];
                            for (int x = d + a.gd + 1; x <= d + cd; ++x) sum += (double)row[x];
                        } else {
                            for (int x = d - cd; x <= d + cd; ++x) sum += (double)row[x];
                        }
                    }
                    noise = sum / (double)a.n_train;
                } else {                            // GO / SO: 1-D only (tr == gr == 0)
                    const float* row = map + r * a.pitch;
                    double left = 0.0, right = 0.0;
                    for (int x = d - cd; x < d - a.gd; ++x) left += (double)row[x];
                    for (int x = d + a.gd + 1; x <= d + cd; ++x) right += (double)row[x];
                    left /= (double)a.td;
                    right /= (double)a.td;
                    noise = (a.kind == RADAR_CFAR_GO) ? fmax(left, right) : fmin(left, right);
                }
                const int cell = r * a.cols + d;
                mask[cell] = ((double)map[r * a.pitch + d] > a.alpha * noise) ? 1 : 0;
                if (noise_dst) noise_dst[cell] = noise;
            }
        }
    }
    __syncthreads();

    if (a.mask_out) {
        uint8_t* dst = a.mask_out + m * n_cells;
        for (int i = threadIdx.x; i < n_cells; i += blockDim.x) dst[i] = mask[i];
    }
    if (!a.dets_out && !a.count_out) return;

    // ordered compaction: each warp owns a contiguous, 32-aligned span of cells
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int span = (((n_cells + kCfarWarps - 1) / kCfarWarps) + 31) & ~31;
    const int begin = warp * span;
    const int end = min(begin + span, n_cells);
    int cnt = 0;
    for (int i = begin + lane; i < begin + span; i += 32) {
        const bool hit = (i < end) && mask[i];
        cnt += __popc(__ballot_sync(0xffffffffu, hit));
    }
    if (lane == 0) warp_counts[warp] = cnt;
    __syncthreads();
    int offset = 0, total = 0;
    for (int w = 0; w < kCfarWarps; ++w) {
        if (w < warp) offset += warp_counts[w];
        total += warp_counts[w];
    }
    if (threadIdx.x == 0 && a.count_out) a.count_out[m] = total;
    if (!a.dets_out) return;
    int32_t* dets = a.dets_out + m * (long long)a.max_dets * 2;
    for (int i = begin + lane; i < begin + span; i += 32) {
        const bool hit = (i < end) && mask[i];
        const unsigned ballot = __ballot_sync(0xffffffffu, hit);
        if (hit) {
            const int slot = offset + __popc(ballot & ((1u << lane) - 1u));
            if (slot < a.max_dets) {
                const int r = (int)a.div_cols.div((uint32_t)i);
                dets[2 * slot] = r;
                dets[2 * slot + 1] = i - r * a.cols;
            }
        }
        offset += __popc(ballot);
    }
}

static int resolve_params(radar_ctx* ctx, const radar_cfar_params& p, bool one_d, CfarArgs* a) {
    a->kind = p.kind;
    a->tr = one_d ? 0 : p.train_r;  a->gr = one_d ? 0 : p.guard_r;
    a->td = p.train_d;              a->gd = p.guard_d;
    if (a->tr < 0 || a->td < 0 || a->gr < 0 || a->gd < 0) return set_error(ctx, RADAR_E_INVALID, "negative CFAR window");
    const int wr = 2 * (a->tr + a->gr) + 1, wd = 2 * (a->td + a->gd) + 1;
    const int n_all = wr * wd - (2 * a->gr + 1) * (2 * a->gd + 1);
    if (n_all <= 0) return set_error(ctx, RADAR_E_INVALID, "CFAR window has no training cells");
    switch (p.kind) {
        case RADAR_CFAR_CA:
            a->n_train = n_all;
            a->alpha = p.alpha > 0 ? p.alpha : radar_cfar_ca_alpha(n_all, p.pfa);
            break;
        case RADAR_CFAR_GO: case RADAR_CFAR_SO:
            if (!one_d) return set_error(ctx, RADAR_E_INVALID, "GO/SO-CFAR exist only in 1-D (reference go_so_cfar.py)");
            a->n_train = a->td;
            a->alpha = p.alpha > 0 ? p.alpha : radar_cfar_ca_alpha(a->td, p.pfa);
            break;
        case RADAR_CFAR_OS: {
            a->n_train = n_all;
            int k = p.os_rank > 0 ? p.os_rank : (int)(p.rho * (double)n_all);
            if (p.os_rank <= 0) k = k < 1 ? 1 : (k > n_all ? n_all : k);
            if (k < 1 || k > n_all) return set_error(ctx, RADAR_E_INVALID, "k_rank must be between 1 and %d, got %d", n_all, k);
            if (wr * wd > 32 * 32) return set_error(ctx, RADAR_E_UNSUPPORTED, "OS-CFAR window of %d cells exceeds 1024", wr * wd);
            a->k_rank = k;
            a->alpha = p.alpha;
            break;
        }
        default:
            return set_error(ctx, RADAR_E_INVALID, "unknown CFAR kind %d", p.kind);
    }
    a->div_wd = make_fastdiv((uint32_t)wd);
    a->ph1_extra = ctx->cfar_ph1 >= 0 ? ctx->cfar_ph1 : 2;
    return RADAR_OK;
}

static int launch_cfar(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const radar_cfar_params& p,
                       bool one_d, const uint8_t* row_select, double* noise_out, uint8_t* mask_out,
                       int32_t* dets_out, int32_t* count_out, int max_dets, cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return set_error(ctx, RADAR_E_INVALID, "empty map");
    if (dets_out && max_dets <= 0) return set_error(ctx, RADAR_E_INVALID, "max_dets must be positive");
    if (n_maps <= 0) return RADAR_OK;
    CfarArgs a{};
    if (int rc = resolve_params(ctx, p, one_d, &a)) return rc;
    a.maps = maps; a.rows = rows; a.cols = cols;
    a.row_select = row_select;
    a.noise_out = noise_out; a.mask_out = mask_out; a.dets_out = dets_out; a.count_out = count_out; a.max_dets = max_dets;
    a.div_cols = make_fastdiv((uint32_t)cols);
    const int nvd = cols - 2 * (a.td + a.gd);
    a.div_vd = make_fastdiv((uint32_t)(nvd > 0 ? nvd : 1));
    const size_t n_cells = (size_t)rows * cols;
    a.pitch = cols;
    if (nvd > 0) while ((a.pitch - nvd) % 32 != 0) ++a.pitch;
    const int nvr = rows - 2 * (a.tr + a.gr);
    const size_t n_work_max = (size_t)(nvr > 0 ? nvr : 0) * (size_t)(nvd > 0 ? nvd : 0);
    const size_t smem = arena_bytes((size_t)rows * a.pitch, 4) + arena_bytes(n_cells, 1) + 2 * arena_bytes(n_work_max, 4);
    if (smem > (size_t)ctx->max_smem_optin)
        return set_error(ctx, RADAR_E_UNSUPPORTED, "CFAR map of %d x %d cells exceeds shared memory", rows, cols);
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(cfar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cfar_kernel<<<(unsigned)n_maps, kCfarThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    return RADAR_OK;
}

int launch_cfar_2d(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const radar_cfar_params& p,
                   double* noise_out, uint8_t* mask_out, int32_t* dets_out, int32_t* count_out, int max_dets,
                   cudaStream_t st) {
    return launch_cfar(ctx, maps, n_maps, rows, cols, p, false, nullptr, noise_out, mask_out, dets_out, count_out, max_dets, st);
}

int launch_cfar_1d_rows(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const radar_cfar_params& p,
                        const uint8_t* row_select, double* noise_out, uint8_t* mask_out, int32_t* dets_out,
                        int32_t* count_out, int max_dets, cudaStream_t st) {
    return launch_cfar(ctx, maps, n_maps, rows, cols, p, true, row_select, noise_out, mask_out, dets_out, count_out, max_dets, st);
}

}  // namespace rb
