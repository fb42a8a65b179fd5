// Micro-benchmark: issue rate of scalar FFMA vs packed FFMA2/FADD2 on sm_100a (decides how the FFT butterflies are written).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpk(u64 v, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
    float a[16];
    u64 p[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
#pragma unroll
    for (int i = 0; i < 8; ++i) p[i] = pk(a[2 * i], a[2 * i + 1]);
    u64 ss = pk(s, s * 0.5f), tt = pk(0.25f, 0.125f);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, 0.25f + a[(i + 1) & 15] * 0.0f);   // keep 3-reg form below
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(s), "f"(a[(i + 5) & 15]));
        } else if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(ss), "l"(p[(i + 3) & 7]));
        } else if (MODE == 3) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(p[(i + 3) & 7]));
        } else if (MODE == 4) {
#pragma unroll
            for (int i = 0; i < 16; ++i) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(a[(i + 5) & 15]));
        } else if (MODE == 5) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(tt));
        }
    }
    float acc = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc += a[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) { float x, y; unpk(p[i], x, y); acc += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE> void run(const char* name, int per_iter_instr, int flops_per_instr) {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    int iters = 20000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 8, 256>>>(out, 100, 1.0001f);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(out, iters, 1.0001f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double warp_instr = (double)148 * 8 * 8 * iters * per_iter_instr;
    printf("%-14s %8.3f ms  %7.2f Gwarp-instr/s  %6.2f TFLOP/s  (%.3f warp-instr/clk/SM at 1.965 GHz)\n", name, ms,
           warp_instr / ms / 1e6, warp_instr * 32 * flops_per_instr / ms / 1e9, warp_instr / (ms * 1e-3) / 148 / 1.965e9);
    cudaFree(out);
}
int main() {
    run<1>("FFMA 3reg", 16, 2);
    run<4>("FADD", 16, 1);
    run<2>("FFMA2", 8, 4);
    run<3>("FADD2", 8, 2);
    run<5>("FMUL2", 8, 2);
    return 0;
}
