// Mixed-radix in-place FFT plan shared by host (planner, CPU self-check) and device (stage executor).
//
// Decimation-in-frequency over N = r0*r1*...: stage j works on sub-arrays of length N_j = N/(r0..r_{j-1});
// a "task" is one radix-r_j butterfly on the elements {base + M_j*i}, M_j = N_j/r_j, followed by the
// twiddles W_{N_j}^{n2*k1}.  Output bin k = k0 + r0*(k1 + r1*(...)) ends at position
// M_0*k0 + M_1*k1 + ... (digit-reversed); `pos[k]` tabulates it so consumers gather in natural order.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "fft_radix.cuh"

namespace rb {

constexpr int kMaxStages = 8;
constexpr int kMaxGenericRadix = 128;   // largest prime factor handled (O(p^2) path)

struct FastDiv {            // exact for n*d < 2^32
    uint32_t magic, d;
    RB_HD uint32_t div(uint32_t n) const {
#ifdef __CUDA_ARCH__
        return d == 1 ? n : __umulhi(n, magic);
#else
        return d == 1 ? n : (uint32_t)(((uint64_t)n * magic) >> 32);
#endif
    }
};
inline FastDiv make_fastdiv(uint32_t d) {
    FastDiv f;
    f.d = d;
    f.magic = d <= 1 ? 0u : (uint32_t)((0x100000000ull / d) + 1ull);
    return f;
}

struct FftPlan {
    int n;
    int n_stages;
    int radix[kMaxStages];
    int sub_len[kMaxStages];   // N_j
    int m[kMaxStages];         // M_j
    int tw_mul[kMaxStages];    // N / N_j : stage twiddle W_{N_j}^e == table[tw_mul*e]
    FastDiv m_div[kMaxStages];
};

inline bool is_fast_radix(int r) {
    switch (r) { case 2: case 3: case 4: case 5: case 6: case 7: case 8: case 9: case 10: case 11: case 13: case 16: return true; }
    return false;
}

// Fewest stages first, then the smallest maximum radix (balanced butterflies, bounded registers).
inline bool plan_search(int n, int depth, int* cur, int& best_depth, int& best_max, int* best) {
    static const int cand[] = {10, 9, 8, 7, 6, 5, 4, 3, 2, 16, 13, 11};
    if (n == 1) {
        int mx = 0;
        for (int i = 0; i < depth; ++i) mx = cur[i] > mx ? cur[i] : mx;
        if (depth < best_depth || (depth == best_depth && mx < best_max)) {
            best_depth = depth; best_max = mx;
            for (int i = 0; i < depth; ++i) best[i] = cur[i];
        }
        return true;
    }
    if (depth >= kMaxStages || depth >= best_depth) return false;
    bool any = false;
    for (int c : cand)
        if (n % c == 0) { cur[depth] = c; any |= plan_search(n / c, depth + 1, cur, best_depth, best_max, best); }
    if (!any) {   // leftover with a large prime factor: peel the smallest prime factor > 13 as a generic stage
        int p = 2;
        while (n % p) ++p;
        if (p > kMaxGenericRadix) return false;
        cur[depth] = p;
        return plan_search(n / p, depth + 1, cur, best_depth, best_max, best);
    }
    return any;
}

// Returns false when N has a prime factor above kMaxGenericRadix or needs too many stages.
inline bool make_plan(int n, FftPlan* plan) {
    if (n < 1) return false;
    int cur[kMaxStages], best[kMaxStages];
    int best_depth = kMaxStages + 1, best_max = 1 << 30;
    if (n == 1) { best_depth = 0; }
    else if (!plan_search(n, 0, cur, best_depth, best_max, best) || best_depth > kMaxStages) return false;
    // ascending radix order: the (twiddle-free) last stage gets the largest butterfly
    for (int i = 0; i < best_depth; ++i)
        for (int j = i + 1; j < best_depth; ++j)
            if (best[j] < best[i]) { int t = best[i]; best[i] = best[j]; best[j] = t; }
    plan->n = n;
    plan->n_stages = best_depth;
    int len = n;
    for (int j = 0; j < kMaxStages; ++j) {
        plan->radix[j] = 1; plan->sub_len[j] = 1; plan->m[j] = 1; plan->tw_mul[j] = 0;
        plan->m_div[j] = make_fastdiv(1);
    }
    for (int j = 0; j < best_depth; ++j) {
        plan->radix[j] = best[j];
        plan->sub_len[j] = len;
        plan->m[j] = len / best[j];
        plan->tw_mul[j] = n / len;
        plan->m_div[j] = make_fastdiv((uint32_t)plan->m[j]);
        len /= best[j];
    }
    return true;
}

// pos[k]: where natural-order bin k sits after the in-place transform.
inline void plan_positions(const FftPlan& p, uint16_t* pos) {
    for (int k = 0; k < p.n; ++k) {
        int rem = k, at = 0;
        for (int j = 0; j < p.n_stages; ++j) {
            at += p.m[j] * (rem % p.radix[j]);
            rem /= p.radix[j];
        }
        pos[k] = (uint16_t)at;
    }
}

// ---- stage executor (one task) ---------------------------------------------------------------
// data: element e of the transform lives at data[e*stride].  tw: N-entry table W_N^j.
template <int R>
RB_HD void stage_task(cf* data, int stride, int base, int m, int n2, int tw_mul, const cf* tw) {
    cf v[R];
#pragma unroll
    for (int i = 0; i < R; ++i) v[i] = data[(base + m * i) * stride];
    Dft<R>::run(v);
    if (m > 1) {
        const int step = tw_mul * n2;
#pragma unroll
        for (int k = 1; k < R; ++k) v[k] = cmul(v[k], tw[step * k]);
    }
#pragma unroll
    for (int k = 0; k < R; ++k) data[(base + m * k) * stride] = v[k];
}

// O(p^2) butterfly for prime factors without a hand-written kernel (23, 127, ...).
// tw_p_mul = N/p so that W_p^e == tw[tw_p_mul*(e mod p)].
#ifdef __CUDA_ARCH__
__device__ __noinline__
#else
inline
#endif
void stage_task_generic(cf* data, int stride, int base, int m, int n2, int tw_mul, const cf* tw, int p, int n_total) {
    cf loc[kMaxGenericRadix];
    for (int i = 0; i < p; ++i) loc[i] = data[(base + m * i) * stride];
    const int tw_p_mul = n_total / p;
    for (int k = 0; k < p; ++k) {
        cf acc = loc[0];
        int e = 0;
        for (int i = 1; i < p; ++i) {
            e += k; if (e >= p) e -= p;
            acc = cadd(acc, cmul(loc[i], tw[tw_p_mul * e]));
        }
        if (m > 1 && k > 0) acc = cmul(acc, tw[tw_mul * n2 * k]);
        data[(base + m * k) * stride] = acc;
    }
}

}  // namespace rb
