// CPU self-check of the mixed-radix plan + butterflies (the same __host__ __device__ code the kernels run).
// Build: nvcc -O2 -I mmwave_radar_processing_b200/csrc tests/native/host_fft_check.cu -o build/host_fft_check
// Prints one line per length and exits non-zero if any relative L2 error exceeds 2e-6.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "fft_plan.h"

using namespace rb;

template <int R>
static void run_stage(std::vector<cf>& a, const FftPlan& p, int st, const std::vector<cf>& tw) {
    const int tasks = p.n / R;
    for (int u = 0; u < tasks; ++u) {
        const int b = (int)p.m_div[st].div((uint32_t)u);
        const int n2 = u - b * p.m[st];
        stage_task<R>(a.data(), 1, b * p.sub_len[st] + n2, p.m[st], n2, p.tw_mul[st], tw.data());
    }
}

static double check(int n) {
    FftPlan p;
    if (!make_plan(n, &p)) { printf("N=%d: no plan\n", n); return 1.0; }
    std::vector<cf> tw(n);
    for (int j = 0; j < n; ++j) {
        double ang = -2.0 * M_PI * (double)j / (double)n;
        tw[j] = cmake((float)cos(ang), (float)sin(ang));
    }
    std::vector<uint16_t> pos(n);
    plan_positions(p, pos.data());
    std::vector<cf> a(n);
    std::vector<double> xr(n), xi(n);
    unsigned s = 12345u + (unsigned)n;
    for (int i = 0; i < n; ++i) {
        s = s * 1664525u + 1013904223u; xr[i] = (double)((int)(s >> 16) % 2000 - 1000);
        s = s * 1664525u + 1013904223u; xi[i] = (double)((int)(s >> 16) % 2000 - 1000);
        a[i] = cmake((float)xr[i], (float)xi[i]);
    }
    for (int st = 0; st < p.n_stages; ++st) {
        switch (p.radix[st]) {
            case 2: run_stage<2>(a, p, st, tw); break;
            case 3: run_stage<3>(a, p, st, tw); break;
            case 4: run_stage<4>(a, p, st, tw); break;
            case 5: run_stage<5>(a, p, st, tw); break;
            case 6: run_stage<6>(a, p, st, tw); break;
            case 7: run_stage<7>(a, p, st, tw); break;
            case 8: run_stage<8>(a, p, st, tw); break;
            case 9: run_stage<9>(a, p, st, tw); break;
            case 10: run_stage<10>(a, p, st, tw); break;
            case 11: run_stage<11>(a, p, st, tw); break;
            case 13: run_stage<13>(a, p, st, tw); break;
            case 16: run_stage<16>(a, p, st, tw); break;
            default: {
                const int r = p.radix[st], tasks = n / r;
                for (int u = 0; u < tasks; ++u) {
                    const int b = u / p.m[st], n2 = u % p.m[st];
                    stage_task_generic(a.data(), 1, b * p.sub_len[st] + n2, p.m[st], n2, p.tw_mul[st], tw.data(), r, n);
                }
            }
        }
    }
    double num = 0, den = 0;
    for (int k = 0; k < n; ++k) {
        double sr = 0, si = 0;
        for (int i = 0; i < n; ++i) {
            double ang = -2.0 * M_PI * (double)((long long)k * i % n) / (double)n;
            sr += xr[i] * cos(ang) - xi[i] * sin(ang);
            si += xr[i] * sin(ang) + xi[i] * cos(ang);
        }
        double dr = a[pos[k]].x - sr, di = a[pos[k]].y - si;
        num += dr * dr + di * di;
        den += sr * sr + si * si;
    }
    double err = sqrt(num / den);
    printf("N=%4d stages=%d radices=", n, p.n_stages);
    for (int j = 0; j < p.n_stages; ++j) printf("%d%s", p.radix[j], j + 1 < p.n_stages ? "x" : "");
    printf(" rel_l2=%.3e\n", err);
    return err;
}

int main() {
    const int lens[] = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 23, 24, 25, 27, 30, 32, 36, 40, 48, 49, 50,
                        60, 63, 64, 70, 80, 81, 90, 96, 100, 115, 120, 121, 126, 127, 128, 130, 169, 200, 254, 256, 512, 1000, 1024};
    double worst = 0;
    for (int n : lens) { double e = check(n); if (e > worst) worst = e; }
    printf("worst rel_l2 = %.3e\n", worst);
    return worst < 2e-6 ? 0 : 1;
}
