// Shared-memory tile FFT machinery used by the range-Doppler, range-profile, range-angle and DBS kernels.
//
// A CTA owns a 2-D complex tile in shared memory.  `fft_axis` runs every stage of the mixed-radix plan
// along one axis ("F", element stride strideF) for all O lines along the other axis ("O", stride
// strideO).  Lanes always advance along the O axis, whose stride is 1 or an odd pitch, so every 8-byte
// shared-memory access of a warp is bank-conflict free.
#pragma once
#include "fft_plan.h"

namespace rb {

template <int R>
__device__ __forceinline__ void run_stage(cf* tile, int strideF, int strideO, int O, FastDiv odiv,
                                          const FftPlan& p, int st, const cf* tw) {
    const int m = p.m[st], nj = p.sub_len[st], twm = p.tw_mul[st];
    const FastDiv mdiv = p.m_div[st];
    const int tasks = (p.n / R) * O;
    for (int t = threadIdx.x; t < tasks; t += blockDim.x) {
        const uint32_t u = odiv.div((uint32_t)t);
        const int o = t - (int)u * O;
        const uint32_t b = mdiv.div(u);
        const int n2 = (int)u - (int)b * m;
        stage_task<R>(tile + o * strideO, strideF, (int)b * nj + n2, m, n2, twm, tw);
    }
}

static __device__ __noinline__ void run_stage_generic(cf* tile, int strideF, int strideO, int O, FastDiv odiv,
                                               const FftPlan& p, int st, const cf* tw) {
    const int r = p.radix[st], m = p.m[st], nj = p.sub_len[st], twm = p.tw_mul[st];
    const FastDiv mdiv = p.m_div[st];
    const int tasks = (p.n / r) * O;
    for (int t = threadIdx.x; t < tasks; t += blockDim.x) {
        const uint32_t u = odiv.div((uint32_t)t);
        const int o = t - (int)u * O;
        const uint32_t b = mdiv.div(u);
        const int n2 = (int)u - (int)b * m;
        stage_task_generic(tile + o * strideO, strideF, (int)b * nj + n2, m, n2, twm, tw, r, p.n);
    }
}

// All stages; a __syncthreads() closes every stage (the caller syncs before the first one).
__device__ __forceinline__ void fft_axis(cf* tile, int strideF, int strideO, int O, FastDiv odiv,
                                         const FftPlan& p, const cf* tw) {
    for (int st = 0; st < p.n_stages; ++st) {
        switch (p.radix[st]) {
            case 2: run_stage<2>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 3: run_stage<3>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 4: run_stage<4>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 5: run_stage<5>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 6: run_stage<6>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 7: run_stage<7>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 8: run_stage<8>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 9: run_stage<9>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 10: run_stage<10>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 11: run_stage<11>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 13: run_stage<13>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            case 16: run_stage<16>(tile, strideF, strideO, O, odiv, p, st, tw); break;
            default: run_stage_generic(tile, strideF, strideO, O, odiv, p, st, tw); break;
        }
        __syncthreads();
    }
}

// Cooperative copy of a small table from global to shared memory.
template <typename T>
__device__ __forceinline__ void load_table(T* dst, const T* __restrict__ src, int n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

__device__ __forceinline__ int odd_pitch(int n) { return n | 1; }

// Carve a dynamic shared-memory arena (16-byte aligned pieces).
struct SmemArena {
    unsigned char* p;
    __device__ explicit SmemArena(unsigned char* base) : p(base) {}
    template <typename T> __device__ T* take(int count) {
        T* r = reinterpret_cast<T*>(p);
        p += ((size_t)count * sizeof(T) + 15) & ~(size_t)15;
        return r;
    }
};
__host__ __device__ constexpr size_t arena_bytes(size_t count, size_t elem) { return (count * elem + 15) & ~(size_t)15; }

}  // namespace rb
