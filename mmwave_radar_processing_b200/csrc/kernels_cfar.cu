// CFAR family + ordered detection compaction: one CTA per map (or per tile of a map that exceeds one CTA's shared memory),
// the map resident in shared memory (one bulk copy of the TMA engine when the layout is unpadded).
//
// Reference arithmetic reproduced (paths relative to mmwave_radar_processing/ in the reference):
//   detectors/base.py:38-65, :208-230 (strict x > threshold, np.where order),
//   detectors/ca_cfar.py:11-155, detectors/os_cfar.py:11-195, detectors/go_so_cfar.py:11-123.
//
// * CA: separable sliding box sums in fp64 with error-free accumulation (ca_box_cells); GO / SO (1-D) and CA with
//   RADAR_B200_CFAR_BOX=0: one thread per cell under test, training sums accumulated in fp64.
// * OS, mask only (the streaming path): a ladder of four levels around the expected order statistic - the window counts at
//   the levels are carried as packed bytes by sliding sums and decide all but a fraction of a percent of the cells; the rest
//   get an exact count (os_ladder_cells).  Detector settings the packed fields cannot hold, row_select, or
//   RADAR_B200_CFAR_LADDER=0: one thread per cell under test counts the training cells below a per-cell bound and stops as
//   soon as the decision is fixed (os_count_cells).
// * OS with noise estimates requested: one warp per cell under test; every lane keeps its share of the training window in registers
//   and the exact k-th order statistic is built bit by bit (MSB->LSB) from warp-wide counts
//   (__reduce_add_sync), which is an exact selection on order-preserving integer keys of the fp32 values.
//   Maps with a negative cell (detect() accepts signed data; the counting paths assume magnitudes) always take this path.
// * decision in fp64: (double)x > alpha*(double)noise - the same expression the host layer uses to publish
//   `thresholds`.
// * compaction: every thread owns a contiguous run of cells (popcounts of whole mask words, shuffle scan, warp totals through
//   shared memory), then an ordered scatter - row-major (range, Doppler) like np.where, one global counter write per map.
//   Tiled maps: the tiles write a full-size mask and cfar_compact_kernel does the same per map.
#include <cmath>
#include "radar_internal.h"
#include "rd_core.cuh"
#include "mbar.cuh"

namespace rb {

constexpr int kCfarThreads = 256;
constexpr int kCfarWarps = kCfarThreads / 32;
constexpr int kMapSpare = 48;         // floats behind the map in shared memory

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
    int ca_box;           // CA: separable sliding box sums (ca_box_cells) instead of one window walk per cell
    int ladder;           // OS mask-only path: 1 = level-ladder screen (os_ladder_cells) before the exact count
    int seg_d, seg_r;     // ladder: cells per sliding segment along Doppler / range
    int nseg_d, nseg_r;   // ladder: segments per row / per column
    FastDiv div_nseg_d;
    float lvl_mul[4];     // ladder: the four levels as multiples of the sampled map median
    // maps that do not fit one CTA's shared memory: rows x cols above is one TILE (valid cells + halo), the map is full_rows x
    // full_cols, tile (i, j) starts at min(i*tile_r, full_rows - rows), min(j*tile_c, full_cols - cols) - edge tiles move
    // inward, so every tile has the same shape and overlapping cells are computed twice with identical results
    int tiled, full_rows, full_cols, tile_r, tile_c, n_tr, n_tc;
    uint8_t* mask_glob;   // tiled: full-size mask the tiles write (mask_out or scratch); cfar_compact_kernel reads it
};

// order-preserving map of all floats (negative ones included: detect() accepts signed data) onto unsigned keys
__device__ __forceinline__ unsigned float_key(float v) {
    const unsigned b = __float_as_uint(v);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float key_float(unsigned k) {
    return __uint_as_float(k ^ ((k >> 31) ? 0x80000000u : 0xffffffffu));
}

template <int MAXV>
__device__ __forceinline__ void os_cells(const CfarArgs& a, const float* map, uint8_t* mask, double* noise_dst, int npitch) {
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
        for (int i = 0; i < MAXV; ++i) vals[i] = off[i] >= 0 ? float_key(map[base + off[i]]) : 0xffffffffu;
        unsigned res = 0;
#pragma unroll 1
        for (int bit = 31; bit >= 0; --bit) {
            const unsigned cand = res | (1u << bit);
            int cnt = 0;
#pragma unroll
            for (int i = 0; i < MAXV; ++i) cnt += (vals[i] < cand) ? 1 : 0;
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (cnt < a.k_rank) res = cand;     // fewer than k cells below cand: the k-th smallest is >= cand
        }
        if (lane == 0) {
            const int cell = r * a.cols + d;
            const double noise = (double)key_float(res);
            mask[cell] = ((double)map[r * a.pitch + d] > a.alpha * noise) ? 1 : 0;
            if (noise_dst) noise_dst[r * npitch + d] = noise;
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


// ---- OS-CFAR decision by a ladder of four levels -------------------------------------------------------------------
// (reference: detectors/os_cfar.py:94-195; the decision x > alpha * v_k is the one os_count_cells makes.)
// For a level l (a float) the number of training cells below l, n(l), brackets the number below the per-cell bound T:
//   alpha*(double)l <  (double)x  =>  l <  T  =>  n(l) <= n(T)      (a "lower" level: n(l) >= k proves a detection)
//   alpha*(double)l >= (double)x  =>  l >= T  =>  n(T) <= n(l)      (an "upper" level: n(l) < k rules one out)
// because T is the smallest float whose rounded product with alpha reaches x and rounded multiplication is monotone.  The
// window counts at FOUR levels are carried as the four bytes of one word: an indicator byte per level and map cell, sliding
// sums along Doppler (full window width in the low 5 bits of each byte, guard width in the high 3) and then along range
// (at most 255 cells per window), a handful of instructions per cell instead of ~100 compares.  The levels sit around the
// expected order statistic of a noise window (multiples of a 64-sample median of the map), so a noise cell is ruled out by
// the level between its T and its window's v_k and a strong cell is confirmed by the level between v_k and T; only cells
// whose T and v_k share a bracket (0.5 % of the 6843 / 1843 bench maps, a few % with alpha = 2) are left to the exact
// count, done by one warp per cell.  The mask is identical to the exact paths' (tests compare all three).
constexpr int kLadderSeg = 16;        // outputs per thread of the compile-time Doppler pass
__device__ __forceinline__ unsigned expand4(unsigned q) { return (q * 0x00204081u) & 0x01010101u; }

// TD > 0: the detector's (train, guard) pairs compiled in (the reference's settings); TD == 0: run-time widths.
template <int TR, int GR, int TD, int GD>
__device__ __forceinline__ void os_ladder_cells(const CfarArgs& a, const float* map, uint8_t* mask, unsigned* lad,
                                                int* work_cell, int* work_n, float* lvl /* [9]: levels, compare bounds, median */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tr = TD > 0 ? TR : a.tr, gr = TD > 0 ? GR : a.gr, td = TD > 0 ? TD : a.td, gd = TD > 0 ? GD : a.gd;
    const int cr = tr + gr, cd = td + gd;
    const int wr = 2 * cr + 1, wd = 2 * cd + 1;
    const int nvr = a.rows - 2 * cr, nvd = a.cols - 2 * cd;
    const int n_cells = a.rows * a.cols;
    // -- levels: median of 64 samples on a regular grid (rank by counting, ties by index: four threads per sample, 16
    //    comparisons each), times lvl_mul
    {
        float* samp = reinterpret_cast<float*>(lad);        // the sliding-sum words are written later
        if (threadIdx.x < 64) {
            const int stride = n_cells / 64;
            samp[threadIdx.x] = map[threadIdx.x * stride + stride / 2];       // ladder path: the map is unpadded
        }
        if (threadIdx.x == 0) lvl[8] = 0.0f;
        __syncthreads();
        const int i = threadIdx.x >> 2, part = threadIdx.x & 3;
        const float v = samp[i];
        int rank = 0;
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
            const int jx = part * 16 + jj;
            const float sj = samp[jx];
            rank += (sj < v || (sj == v && jx < i)) ? 1 : 0;
        }
        rank += __shfl_xor_sync(0xffffffffu, rank, 1);
        rank += __shfl_xor_sync(0xffffffffu, rank, 2);
        if (part == 0 && rank == 31) lvl[8] = v;
        __syncthreads();
        if (threadIdx.x < 4) {
            float l = lvl[8] * a.lvl_mul[threadIdx.x];
            // a negative or NaN median (maps are magnitudes, but detect() takes anything): the level moves to +inf, where it
            // is never a lower level and, with every finite cell below it, never rules a detection out
            if (!(l >= 0.0f)) l = __int_as_float(0x7f800000);
            const double th = a.alpha * (double)l;           // x > th  <=>  alpha*l < x  (l is a lower level for this cell)
            float f = (float)th;                             // largest float <= th: x > th <=> x > f for float x
            if ((double)f > th) f = __uint_as_float(__float_as_uint(f) - 1u);
            lvl[threadIdx.x] = l;
            lvl[4 + threadIdx.x] = f;
        }
    }
    __syncthreads();
    const float l0 = lvl[0], l1 = lvl[1], l2 = lvl[2], l3 = lvl[3];
    // indicator bytes of one cell: byte j = (m < level j)
    auto indicators = [&](float m) -> unsigned {
        unsigned e = m < l0 ? 1u : 0u;
        if (m < l1) e |= 0x100u;
        if (m < l2) e |= 0x10000u;
        if (m < l3) e |= 0x1000000u;
        return e;
    };
    // -- sliding sums along Doppler: lad[r][d] = (cells below each level in [d-cd, d+cd]) | (in [d-gd, d+gd]) << 5, per byte
    if (TD > 0) {
        // compile-time widths: a thread takes SEG outputs of one row; the indicators of its SEG + 2cd cells stay in registers
        constexpr int SEG = kLadderSeg, CD = TD + GD, NQ = SEG + 2 * CD;
        const int nseg = a.nseg_d;
        for (int t = threadIdx.x; t < a.rows * nseg; t += blockDim.x) {
            const int r = (int)a.div_nseg_d.div((uint32_t)t), sg = t - r * nseg;
            const int d0 = CD + sg * SEG;
            // a short last segment reads on into the next row (the map buffer has kMapSpare floats after its last row);
            // those values only reach outputs that are not stored
            static_assert(NQ <= kMapSpare, "spare floats behind the map");
            const float* mr = map + r * a.cols + d0 - CD;
            unsigned e[NQ];
#pragma unroll
            for (int x = 0; x < NQ; ++x) e[x] = indicators(mr[x]);
            unsigned A = 0, G = 0;
#pragma unroll
            for (int x = 0; x <= 2 * CD; ++x) A += e[x];
#pragma unroll
            for (int x = CD - GD; x <= CD + GD; ++x) G += e[x];
            unsigned* out = lad + r * a.cols + d0;
            const int n_out = min(SEG, a.cols - CD - d0);
#pragma unroll
            for (int o = 0; o < SEG; ++o) {
                if (o > 0) {
                    A += e[o + 2 * CD] - e[o - 1];
                    G += e[o + CD + GD] - e[o + CD - GD - 1];
                }
                if (o < n_out) out[o] = A | (G << 5);
            }
        }
    } else {
        const int nseg = a.nseg_d;
        for (int t = threadIdx.x; t < a.rows * nseg; t += blockDim.x) {
            const int r = (int)a.div_nseg_d.div((uint32_t)t), sg = t - r * nseg;
            const int d0 = cd + sg * a.seg_d, d1 = min(d0 + a.seg_d, a.cols - cd);
            const float* mr = map + r * a.cols;
            unsigned* out = lad + r * a.cols;
            unsigned A = 0, G = 0;
            for (int x = d0 - cd; x <= d0 + cd; ++x) A += indicators(mr[x]);
            for (int x = d0 - gd; x <= d0 + gd; ++x) G += indicators(mr[x]);
            out[d0] = A | (G << 5);
            for (int d = d0 + 1; d < d1; ++d) {
                A += indicators(mr[d + cd]) - indicators(mr[d - cd - 1]);
                G += indicators(mr[d + gd]) - indicators(mr[d - gd - 1]);
                out[d] = A | (G << 5);
            }
        }
    }
    __syncthreads();
    // -- sliding sums along range + decision
    {
        const float f0 = lvl[4], f1 = lvl[5], f2 = lvl[6], f3 = lvl[7];
        const unsigned k4 = (unsigned)a.k_rank * 0x01010101u;
        const int nseg = a.nseg_r;
        for (int t = threadIdx.x; t < nseg * nvd; t += blockDim.x) {
            const int sg = (int)a.div_vd.div((uint32_t)t), vd = t - sg * nvd;
            const int d = vd + cd;
            const int r0 = cr + sg * a.seg_r, r1 = min(r0 + a.seg_r, a.rows - cr);
            const unsigned* col = lad + d;
            unsigned W = 0, Wg = 0;
            if (TD > 0) {
#pragma unroll
                for (int y = -(TR + GR); y <= TR + GR; ++y) W += col[(r0 + y) * a.cols] & 0x1f1f1f1fu;
#pragma unroll
                for (int y = -GR; y <= GR; ++y) Wg += (col[(r0 + y) * a.cols] >> 5) & 0x07070707u;
            } else {
                for (int y = r0 - cr; y <= r0 + cr; ++y) W += col[y * a.cols] & 0x1f1f1f1fu;
                for (int y = r0 - gr; y <= r0 + gr; ++y) Wg += (col[y * a.cols] >> 5) & 0x07070707u;
            }
            const unsigned* cr_in = col + (r0 + cr) * a.cols;      // row entering / leaving the window, entering / leaving the guard band
            const unsigned* cr_out = col + (r0 - cr - 1) * a.cols;
            const unsigned* gr_in = col + (r0 + gr) * a.cols;
            const unsigned* gr_out = col + (r0 - gr - 1) * a.cols;
            const float* xp = map + r0 * a.cols + d;
            uint8_t* mp = mask + r0 * a.cols + d;
            for (int r = r0; r < r1; ++r) {
                if (r > r0) {
                    W += (*cr_in & 0x1f1f1f1fu) - (*cr_out & 0x1f1f1f1fu);
                    Wg += ((*gr_in >> 5) & 0x07070707u) - ((*gr_out >> 5) & 0x07070707u);
                }
                const unsigned Wt = W - Wg;                  // training cells below each level (no borrow: Wg <= W per byte)
                const float x = *xp;
                // bytes of the lower levels (alpha*l < x: the count below l is a lower bound of the count below T)
                const unsigned lower = (x > f0 ? 0xffu : 0u) | (x > f1 ? 0xff00u : 0u) | (x > f2 ? 0xff0000u : 0u) | (x > f3 ? 0xff000000u : 0u);
                const unsigned ge = __vcmpgeu4(Wt, k4);      // 0xff per level whose count reaches k
                if (ge & lower) *mp = 1;                     // a lower level already has k cells below it: detection
                else if ((~ge & ~lower) == 0u)               // no upper level rules it out: undecided
                    work_cell[atomicAdd(work_n, 1)] = (r - cr) * nvd + vd;
                cr_in += a.cols; cr_out += a.cols; gr_in += a.cols; gr_out += a.cols; xp += a.cols; mp += a.cols;
            }
        }
    }
    __syncthreads();
    // -- exact count of the undecided cells: eight lanes per cell, each walking every eighth window cell (the decision's own
    //    predicate alpha*(double)c < (double)x per cell was measured slower than deriving the float bound once)
    const int n_work = *work_n;
    if (n_work == 0) return;
    const int sub = lane >> 3, l8 = lane & 7;
    const int wcells = wr * wd;
    int y0 = 0, x0 = l8;                                     // window cell w = l8 + 8j as (row, column)
    while (x0 >= wd) { x0 -= wd; ++y0; }
    for (int i0 = warp * 4; i0 < n_work; i0 += kCfarWarps * 4) {
        const int i = i0 + sub;
        int cnt = 0, cell = 0;
        if (i < n_work) {
            const int c = work_cell[i];
            const int vr = (int)a.div_vd.div((uint32_t)c);
            const int vd = c - vr * nvd;
            cell = (vr + cr) * a.cols + vd + cd;
            const float tb = os_compare_bound(map[cell], a.alpha);
            const float* win = map + vr * a.cols + vd;
            int y = y0, xx = x0;
            for (int w = l8; w < wcells; w += 8) {
                const bool guard = (y >= tr && y <= tr + 2 * gr) && (xx >= td && xx <= td + 2 * gd);
                cnt += (!guard && win[y * a.cols + xx] < tb) ? 1 : 0;
                xx += 8;
                while (xx >= wd) { xx -= wd; ++y; }
            }
        }
        cnt += __shfl_xor_sync(0xffffffffu, cnt, 1);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, 2);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, 4);
        if (i < n_work && l8 == 0) mask[cell] = cnt >= a.k_rank ? 1 : 0;
    }
}


// ---- CA-CFAR by separable sliding box sums in fp64 ---------------------------------------------------------------------
// (reference: detectors/ca_cfar.py:85-155 - mean of the training cells = (window sum - guard-block sum) / N.)
// Pass 1: H[r][d] = sum of the window's columns d-cd .. d+cd of row r, one thread per (row, Doppler segment), sliding.
// Pass 2: one thread per (Doppler bin, range segment): W = sum of H over the window's rows, G = sum of the guard block
// (its row sums formed on the fly from 2*gd+1 map cells), both sliding.  Every add is an error-free TwoSum with the
// round-off carried along, so a strong cell that leaves the window takes its rounding error with it: the sums are the
// correctly rounded window sums whatever the dynamic range (sums of <= 255 floats are exact in double anyway while the
// window spans < 2^21).  ~45 instructions per cell instead of 3 per window cell (153 cells for the 4,4 / 2,2 detector).
struct DD { double s, c; };
__device__ __forceinline__ void dd_add(DD& a, double b) {
    const double t = a.s + b, bb = t - a.s;
    a.c += (a.s - (t - bb)) + (b - bb);
    a.s = t;
}
__device__ __forceinline__ void ca_box_cells(const CfarArgs& a, const float* map, uint8_t* mask, double* hsum, double* noise_dst,
                                              int npitch, const uint8_t* rsel) {
    const int cr = a.tr + a.gr, cd = a.td + a.gd;
    const int nvr = a.rows - 2 * cr, nvd = a.cols - 2 * cd;
    {
        const int nseg = a.nseg_d;
        for (int t = threadIdx.x; t < a.rows * nseg; t += blockDim.x) {
            const int r = (int)a.div_nseg_d.div((uint32_t)t), sg = t - r * nseg;
            const int d0 = cd + sg * a.seg_d, d1 = min(d0 + a.seg_d, a.cols - cd);
            const float* mr = map + r * a.pitch;
            double* out = hsum + r * a.cols;
            DD h{0.0, 0.0};
            for (int x = d0 - cd; x <= d0 + cd; ++x) dd_add(h, (double)mr[x]);
            out[d0] = h.s + h.c;
            for (int d = d0 + 1; d < d1; ++d) {
                dd_add(h, (double)mr[d + cd]);
                dd_add(h, -(double)mr[d - cd - 1]);
                out[d] = h.s + h.c;
            }
        }
    }
    __syncthreads();
    const int nseg = a.nseg_r;
    for (int t = threadIdx.x; t < nseg * nvd; t += blockDim.x) {
        const int sg = (int)a.div_vd.div((uint32_t)t), vd = t - sg * nvd;
        const int d = vd + cd;
        const int r0 = cr + sg * a.seg_r, r1 = min(r0 + a.seg_r, a.rows - cr);
        auto guard_row = [&](int y) {                    // sum of the guard block's cells of map row y (exact: <= 7 floats)
            const float* p = map + y * a.pitch + d - a.gd;
            double g = 0.0;
            for (int x = 0; x <= 2 * a.gd; ++x) g += (double)p[x];
            return g;
        };
        DD w{0.0, 0.0}, g{0.0, 0.0};
        for (int y = r0 - cr; y <= r0 + cr; ++y) dd_add(w, hsum[y * a.cols + d]);
        for (int y = r0 - a.gr; y <= r0 + a.gr; ++y) dd_add(g, guard_row(y));
        for (int r = r0; r < r1; ++r) {
            if (r > r0) {
                dd_add(w, hsum[(r + cr) * a.cols + d]);
                dd_add(w, -hsum[(r - cr - 1) * a.cols + d]);
                dd_add(g, guard_row(r + a.gr));
                dd_add(g, -guard_row(r - a.gr - 1));
            }
            if (rsel && !rsel[r]) continue;
            // the reference divides the sum by N (ca_cfar.py:139-141: sum(...)/num_train_cells); same operation order here
            const double noise = ((w.s - g.s) + (w.c - g.c)) / (double)a.n_train;
            mask[r * a.cols + d] = ((double)map[r * a.pitch + d] > a.alpha * noise) ? 1 : 0;
            if (noise_dst) noise_dst[r * npitch + d] = noise;
        }
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
    float* map = ar.take<float>(a.rows * a.pitch + kMapSpare);   // spare: alignment shift + round-up of the bulk copy, ladder overreads
    uint8_t* mask = ar.take<uint8_t>(n_cells + 32);
    const int n_work_max = max(a.rows - 2 * (a.tr + a.gr), 0) * max(a.cols - 2 * (a.td + a.gd), 0);   // cells under test
    int* work_cell = ar.take<int>(n_work_max);           // OS counting path: undecided cells after phase 1
    float* work_t = ar.take<float>(a.ladder ? n_cells : (a.ca_box ? 2 * n_cells : n_work_max));   // ladder: packed sliding sums (a word per cell); CA: row sums (a double per cell)
    __shared__ int work_n;
    __shared__ float ladder_lvl[9];
    __shared__ uint64_t load_bar;
    if (threadIdx.x == 0) work_n = 0;
    long long m = blockIdx.x;
    int R0 = 0, C0 = 0, src_pitch = a.cols, npitch = a.cols;
    long long map_cells = n_cells, map_rows = a.rows;
    if (a.tiled) {
        const int tiles = a.n_tr * a.n_tc;
        const int t = (int)(blockIdx.x % (unsigned)tiles);
        m = blockIdx.x / (unsigned)tiles;
        R0 = min((t / a.n_tc) * a.tile_r, a.full_rows - a.rows);
        C0 = min((t % a.n_tc) * a.tile_c, a.full_cols - a.cols);
        src_pitch = npitch = a.full_cols;
        map_cells = (long long)a.full_rows * a.full_cols;
        map_rows = a.full_rows;
    }
    const float* src = a.maps + m * map_cells + (long long)R0 * src_pitch + C0;
    double* noise_dst = a.noise_out ? a.noise_out + m * map_cells + (long long)R0 * npitch + C0 : nullptr;
    const uint8_t* rsel = a.row_select ? a.row_select + m * map_rows + R0 : nullptr;
    // map -> shared memory.  Unpadded layout (ladder path): ONE bulk copy of the TMA engine from the 16-byte aligned address
    // below the map (the copy may run a few bytes into the next map, so the last map of the batch takes the plain loop when
    // its end is not aligned); otherwise batches of independent loads (one load latency per batch, not per cell).
    const uintptr_t src_addr = reinterpret_cast<uintptr_t>(src);
    const unsigned shift_b = (unsigned)(src_addr & 15u);
    const unsigned copy_bytes = (shift_b + (unsigned)n_cells * 4u + 15u) & ~15u;
    const bool want_counting = a.kind == RADAR_CFAR_OS && !noise_dst && a.alpha > 0.0;
    int neg = 0;
    const bool bulk = !a.tiled && a.pitch == a.cols && (m + 1 < (long long)gridDim.x || copy_bytes == shift_b + (unsigned)n_cells * 4u);
    if (bulk) {
        map = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(map) + shift_b);
        if (threadIdx.x == 0) {
            mb::init(&load_bar, 1);
            mb::fence_init();
            mb::arrive_expect_tx(&load_bar, copy_bytes);
            mb::bulk_g2s(reinterpret_cast<unsigned char*>(map) - shift_b, reinterpret_cast<const unsigned char*>(src) - shift_b,
                         copy_bytes, &load_bar);
        }
        for (int i = threadIdx.x; i < (n_cells + 3) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(mask)[i] = 0u;
        if (noise_dst)
            for (int i = threadIdx.x; i < n_cells; i += blockDim.x) noise_dst[i] = 0.0;
        if (threadIdx.x == 0) mb::wait(&load_bar, 0);        // one thread polls; the others sleep at the CTA barrier
        __syncthreads();
        if (want_counting) {                 // sign bits OR-ed together (-0.0 and negative NaNs count as negative: exact path, still right)
            int acc = 0;
#pragma unroll 4
            for (int i = threadIdx.x; i < n_cells; i += blockDim.x) acc |= __float_as_int(map[i]);
            neg = acc < 0 ? 1 : 0;
        }
    } else {
        for (int i0 = threadIdx.x; i0 < n_cells; i0 += 6 * kCfarThreads) {
            float v[6];
#pragma unroll
            for (int j = 0; j < 6; ++j) {
                const int i = i0 + j * kCfarThreads;
                const int y = (int)a.div_cols.div((uint32_t)i);
                v[j] = i < n_cells ? __ldg(src + (long long)y * src_pitch + (i - y * a.cols)) : 0.0f;
            }
#pragma unroll
            for (int j = 0; j < 6; ++j) {
                const int i = i0 + j * kCfarThreads;
                if (i < n_cells) {
                    const int y = (int)a.div_cols.div((uint32_t)i);
                    neg |= v[j] < 0.0f ? 1 : 0;
                    map[y * a.pitch + (i - y * a.cols)] = v[j];
                    mask[i] = 0;
                    if (noise_dst && !a.tiled) noise_dst[i] = 0.0;      // tiled: the launcher cleared the whole array
                }
            }
        }
    }
    if (!bulk && threadIdx.x < 4) mask[n_cells + threadIdx.x] = 0;       // the compaction reads whole words
    // the counting OS decisions rely on non-negative cells (magnitude maps); signed data takes the exact selection
    const bool counting = want_counting && __syncthreads_or(neg) == 0;
    if (!want_counting) __syncthreads();

    const int cr = a.tr + a.gr, cd = a.td + a.gd;
    const int nvr = a.rows - 2 * cr, nvd = a.cols - 2 * cd;
    if (nvr > 0 && nvd > 0) {      // window larger than the map: no valid cell (reference swallows the ValueError)
        if (counting && a.ladder) {
            unsigned* lad = reinterpret_cast<unsigned*>(work_t);
            // (range, Doppler) settings of tests/verify_processors.py:121-143 and gui_configs/processor_params.yaml:41-47
            if (a.tr == 8 && a.gr == 2 && a.td == 4 && a.gd == 1) os_ladder_cells<8, 2, 4, 1>(a, map, mask, lad, work_cell, &work_n, ladder_lvl);
            else if (a.tr == 5 && a.gr == 3 && a.td == 5 && a.gd == 2) os_ladder_cells<5, 3, 5, 2>(a, map, mask, lad, work_cell, &work_n, ladder_lvl);
            else os_ladder_cells<0, 0, 0, 0>(a, map, mask, lad, work_cell, &work_n, ladder_lvl);
        } else if (counting) {
            CfarArgs la = a;                                   // per-map row_select (a local copy only on the paths that read it)
            la.row_select = rsel;
            os_count_dispatch(la, map, mask, work_cell, work_t, &work_n);
        } else if (a.kind == RADAR_CFAR_OS) {
            const int wcells = (2 * cr + 1) * (2 * cd + 1);
            CfarArgs la = a;
            la.row_select = rsel;
            if (wcells <= 32 * 8) os_cells<8>(la, map, mask, noise_dst, npitch);
            else if (wcells <= 32 * 16) os_cells<16>(la, map, mask, noise_dst, npitch);
            else os_cells<32>(la, map, mask, noise_dst, npitch);
        } else if (a.kind == RADAR_CFAR_CA && a.ca_box) {
            ca_box_cells(a, map, mask, reinterpret_cast<double*>(work_t), noise_dst, npitch, rsel);
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
                if (noise_dst) noise_dst[r * npitch + d] = noise;
            }
        }
    }
    __syncthreads();

    if (a.tiled) {
        // the tile's valid cells plus the map's own edge rows / columns it touches; the ordered compaction runs afterwards
        const int r_lo = R0 == 0 ? 0 : cr, r_hi = R0 + a.rows == a.full_rows ? a.rows : a.rows - cr;
        const int c_lo = C0 == 0 ? 0 : cd, c_hi = C0 + a.cols == a.full_cols ? a.cols : a.cols - cd;
        uint8_t* dst = a.mask_glob + m * map_cells + (long long)R0 * a.full_cols + C0;
        for (int i = threadIdx.x; i < n_cells; i += blockDim.x) {
            const int y = (int)a.div_cols.div((uint32_t)i), x = i - y * a.cols;
            if (y >= r_lo && y < r_hi && x >= c_lo && x < c_hi) dst[(long long)y * a.full_cols + x] = mask[i];
        }
        return;
    }
    if (a.mask_out) {
        uint8_t* dst = a.mask_out + m * n_cells;
        for (int i = threadIdx.x; i < n_cells; i += blockDim.x) dst[i] = mask[i];
    }
    if (!a.dets_out && !a.count_out) return;

    // ordered compaction: every thread owns `ch` consecutive cells (whole 32-bit words of mask bytes), counts them with one
    // popcount per word, an exclusive scan over the threads (shuffles inside a warp, shared memory across warps) gives its
    // first slot, and it writes its detections in cell order - row-major (range, Doppler) like np.where.
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ch_words = ((n_cells + kCfarThreads - 1) / kCfarThreads + 3) / 4;
    const int w_begin = threadIdx.x * ch_words;
    const uint32_t* mw = reinterpret_cast<const uint32_t*>(mask);
    const int n_words = (n_cells + 3) / 4;                   // the bytes behind the last cell are zero (see the load above)
    auto mask_word = [&](int w) -> uint32_t {               // detections of cells 4w .. 4w+3 as bits 0, 8, 16, 24
        return w < n_words ? mw[w] & 0x01010101u : 0u;
    };
    int cnt = 0;
    for (int w = w_begin; w < w_begin + ch_words; ++w) cnt += __popc(mask_word(w));
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_counts[warp] = incl;
    __syncthreads();
    int offset = incl - cnt, total = 0;
    for (int w = 0; w < kCfarWarps; ++w) {
        if (w < warp) offset += warp_counts[w];
        total += warp_counts[w];
    }
    if (threadIdx.x == 0 && a.count_out) a.count_out[m] = total;
    if (!a.dets_out || cnt == 0) return;
    int32_t* dets = a.dets_out + m * (long long)a.max_dets * 2;
    for (int w = w_begin; w < w_begin + ch_words; ++w) {
        uint32_t v = mask_word(w);
        while (v) {
            const int cell = 4 * w + ((__ffs(v) - 1) >> 3);
            v &= v - 1u;
            if (offset < a.max_dets) {
                const int r = (int)a.div_cols.div((uint32_t)cell);
                dets[2 * offset] = r;
                dets[2 * offset + 1] = cell - r * a.cols;
            }
            ++offset;
        }
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

// Ordered compaction of full-size masks in global memory (tiled maps): one CTA per map, every thread a contiguous run of cells.
__global__ void __launch_bounds__(kCfarThreads) cfar_compact_kernel(const uint8_t* __restrict__ masks, long long n_cells, FastDiv div_cols,
                                                                     int cols, int32_t* dets_out, int32_t* count_out, int max_dets) {
    __shared__ int warp_counts[kCfarWarps];
    const long long m = blockIdx.x;
    const uint8_t* mask = masks + m * n_cells;
    const long long ch = (n_cells + kCfarThreads - 1) / kCfarThreads;
    const long long begin = (long long)threadIdx.x * ch, end = begin + ch < n_cells ? begin + ch : n_cells;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int cnt = 0;
    for (long long i = begin; i < end; ++i) cnt += mask[i] ? 1 : 0;
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_counts[warp] = incl;
    __syncthreads();
    int offset = incl - cnt, total = 0;
    for (int w = 0; w < kCfarWarps; ++w) {
        if (w < warp) offset += warp_counts[w];
        total += warp_counts[w];
    }
    if (threadIdx.x == 0 && count_out) count_out[m] = total;
    if (!dets_out || cnt == 0) return;
    int32_t* dets = dets_out + m * (long long)max_dets * 2;
    for (long long i = begin; i < end; ++i) {
        if (!mask[i]) continue;
        if (offset < max_dets) {
            const int r = (int)div_cols.div((uint32_t)i);
            dets[2 * offset] = r;
            dets[2 * offset + 1] = (int)(i - (long long)r * cols);
        }
        ++offset;
    }
}

// Per-shape launch state: everything that depends on the (tile) shape the CTA holds in shared memory.
static size_t configure_shape(radar_ctx* ctx, const radar_cfar_params& p, bool mask_only, bool has_row_select, int rows, int cols,
                              CfarArgs* a) {
    a->rows = rows; a->cols = cols;
    a->div_cols = make_fastdiv((uint32_t)cols);
    const int nvd = cols - 2 * (a->td + a->gd);
    a->div_vd = make_fastdiv((uint32_t)(nvd > 0 ? nvd : 1));
    const size_t n_cells = (size_t)rows * cols;
    const int nvr = rows - 2 * (a->tr + a->gr);
    // level-ladder screen: OS, mask only, every packed field wide enough (5 bits of window width, 3 of guard width, 8 of window
    // cells), at least 64 cells to sample; RADAR_B200_CFAR_LADDER=0 switches it off (tests compare the paths)
    const int wr = 2 * (a->tr + a->gr) + 1, wd = 2 * (a->td + a->gd) + 1;
    a->ladder = (p.kind == RADAR_CFAR_OS && mask_only && !has_row_select && a->alpha > 0.0 && ctx->cfar_ladder && nvr > 0 && nvd > 0 &&
                 wd <= 31 && 2 * a->gd + 1 <= 7 && wr * wd <= 255 && n_cells >= 64) ? 1 : 0;
    if (a->ladder) {
        // expected order statistic of a Rayleigh noise window relative to the map median: sqrt(ln(1/(1-k/N)) / ln 2)
        const double q = (double)a->k_rank / (double)a->n_train;
        const double vk = q < 1.0 ? std::sqrt(std::log(1.0 / (1.0 - q)) / std::log(2.0)) : 2.0;
        const double rel[4] = {0.57, 0.85, 1.28, 1.9};
        for (int j = 0; j < 4; ++j) a->lvl_mul[j] = (float)(vk * rel[j]);
        const int per_d = kCfarThreads / rows > 0 ? kCfarThreads / rows : 1;
        const bool fixed = (a->tr == 8 && a->gr == 2 && a->td == 4 && a->gd == 1) || (a->tr == 5 && a->gr == 3 && a->td == 5 && a->gd == 2);
        a->seg_d = fixed ? kLadderSeg : (nvd + per_d - 1) / per_d;   // the kernel's compile-time instantiations use kLadderSeg
        a->nseg_d = (nvd + a->seg_d - 1) / a->seg_d;
        a->div_nseg_d = make_fastdiv((uint32_t)a->nseg_d);
        const int per_r = kCfarThreads / nvd > 0 ? kCfarThreads / nvd : 1;
        a->seg_r = (nvr + per_r - 1) / per_r;
        a->nseg_r = (nvr + a->seg_r - 1) / a->seg_r;
    }
    // CA: separable box sums (RADAR_B200_CFAR_BOX=0: one window walk per cell, kept for comparison)
    a->ca_box = (p.kind == RADAR_CFAR_CA && ctx->cfar_box && nvr > 0 && nvd > 0) ? 1 : 0;
    if (a->ca_box) {
        const int per_d = kCfarThreads / rows > 0 ? kCfarThreads / rows : 1;
        a->seg_d = (nvd + per_d - 1) / per_d;
        a->nseg_d = (nvd + a->seg_d - 1) / a->seg_d;
        a->div_nseg_d = make_fastdiv((uint32_t)a->nseg_d);
        const int per_r = kCfarThreads / nvd > 0 ? kCfarThreads / nvd : 1;
        a->seg_r = (nvr + per_r - 1) / per_r;
        a->nseg_r = (nvr + a->seg_r - 1) / a->seg_r;
    }
    a->pitch = cols;
    if (nvd > 0 && !a->ladder) while ((a->pitch - nvd) % 32 != 0) ++a->pitch;
    const size_t n_work_max = (size_t)(nvr > 0 ? nvr : 0) * (size_t)(nvd > 0 ? nvd : 0);
    return arena_bytes((size_t)rows * a->pitch + kMapSpare, 4) + arena_bytes(n_cells + 32, 1) + arena_bytes(n_work_max, 4) +
           arena_bytes(a->ladder ? n_cells : (a->ca_box ? 2 * n_cells : n_work_max), 4);
}

static int launch_cfar(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const radar_cfar_params& p,
                       bool one_d, const uint8_t* row_select, double* noise_out, uint8_t* mask_out,
                       int32_t* dets_out, int32_t* count_out, int max_dets, cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return set_error(ctx, RADAR_E_INVALID, "empty map");
    if (dets_out && max_dets <= 0) return set_error(ctx, RADAR_E_INVALID, "max_dets must be positive");
    if (n_maps <= 0) return RADAR_OK;
    CfarArgs a{};
    if (int rc = resolve_params(ctx, p, one_d, &a)) return rc;
    a.maps = maps;
    a.row_select = row_select;
    a.noise_out = noise_out; a.mask_out = mask_out; a.dets_out = dets_out; a.count_out = count_out; a.max_dets = max_dets;
    const size_t limit = (size_t)ctx->max_smem_optin - 1024;        // static shared memory of the kernel
    size_t smem = configure_shape(ctx, p, noise_out == nullptr, row_select != nullptr, rows, cols, &a);
    unsigned grid = (unsigned)n_maps;
    if (smem > limit || ctx->cfar_tile > 0) {
        // the map does not fit one CTA: tiles of valid cells plus halo, masks into a full-size array, compaction afterwards
        const int cr = a.tr + a.gr, cd = a.td + a.gd;
        int cols_t = cols, rows_t = rows;
        if (ctx->cfar_tile > 0) {                                   // RADAR_B200_CFAR_TILE=n: tiles of n x n valid cells (tests)
            cols_t = cols < ctx->cfar_tile + 2 * cd ? cols : ctx->cfar_tile + 2 * cd;
            rows_t = rows < ctx->cfar_tile + 2 * cr ? rows : ctx->cfar_tile + 2 * cr;
        } else {
            for (;;) {
                // the tallest tile of this width that fits; narrower tiles when not even one window row band does
                int lo = 2 * cr + 1 < rows ? 2 * cr + 1 : rows, hi = rows;
                if (configure_shape(ctx, p, noise_out == nullptr, row_select != nullptr, lo, cols_t, &a) <= limit) {
                    while (lo < hi) {
                        const int mid = (lo + hi + 1) / 2;
                        if (configure_shape(ctx, p, noise_out == nullptr, row_select != nullptr, mid, cols_t, &a) <= limit) lo = mid; else hi = mid - 1;
                    }
                    rows_t = lo;
                    break;
                }
                if (cols_t <= 2 * cd + 1) return set_error(ctx, RADAR_E_UNSUPPORTED, "CFAR window of %d x %d cells exceeds shared memory", 2 * cr + 1, 2 * cd + 1);
                cols_t = (cols_t + 2 * cd + 1) / 2 > 2 * cd + 1 ? (cols_t + 2 * cd + 1) / 2 : 2 * cd + 1;
            }
        }
        smem = configure_shape(ctx, p, noise_out == nullptr, row_select != nullptr, rows_t, cols_t, &a);
        if (smem > limit) return set_error(ctx, RADAR_E_UNSUPPORTED, "CFAR tile of %d x %d cells exceeds shared memory", rows_t, cols_t);
        a.tiled = 1; a.full_rows = rows; a.full_cols = cols;
        a.tile_r = rows_t - 2 * cr > 0 ? rows_t - 2 * cr : 1;
        a.tile_c = cols_t - 2 * cd > 0 ? cols_t - 2 * cd : 1;
        a.n_tr = rows_t >= rows ? 1 : (rows - 2 * cr + a.tile_r - 1) / a.tile_r;
        a.n_tc = cols_t >= cols ? 1 : (cols - 2 * cd + a.tile_c - 1) / a.tile_c;
        const long long blocks = (long long)n_maps * a.n_tr * a.n_tc;
        if (blocks > 0x7fffffffLL) return set_error(ctx, RADAR_E_INVALID, "too many CFAR tiles in one call");
        grid = (unsigned)blocks;
        const size_t full = (size_t)rows * cols;
        a.mask_glob = mask_out;
        if (!a.mask_glob) {
            if (int rc = ensure_scratch(ctx, (size_t)n_maps * full)) return rc;
            a.mask_glob = reinterpret_cast<uint8_t*>(ctx->scratch);
        }
        if (noise_out) RB_CUDA_OK(ctx, cudaMemsetAsync(noise_out, 0, (size_t)n_maps * full * sizeof(double), st));
    }
    RB_CUDA_OK(ctx, cudaFuncSetAttribute(cfar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cfar_kernel<<<grid, kCfarThreads, smem, st>>>(a);
    RB_CUDA_OK(ctx, cudaGetLastError());
    if (a.tiled && (dets_out || count_out)) {
        cfar_compact_kernel<<<(unsigned)n_maps, kCfarThreads, 0, st>>>(a.mask_glob, (long long)rows * cols, make_fastdiv((uint32_t)cols), cols,
                                                                       dets_out, count_out, max_dets);
        RB_CUDA_OK(ctx, cudaGetLastError());
    }
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
