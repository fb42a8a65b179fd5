// Register-resident small DFT butterflies (forward transform, e^{-2 pi i kn/N}, unnormalised) used by
// the shared-memory mixed-radix FFT stages of the range / Doppler / angle kernels.
//
// numpy's forward DFT convention (np.fft.fft) is what the reference relies on
// (reference: processors/range_doppler_resp.py:98-103); nothing here is taken from pocketfft - these
// are the textbook symmetric forms for prime radices plus in-register Cooley-Tukey for 6/8/9/10/16.
//
// Everything is __host__ __device__ so tests/host_fft_check.cu can run the same code on the CPU.
#pragma once
#include <cuda_runtime.h>

#ifndef RB_HD
#define RB_HD __host__ __device__ __forceinline__
#endif

namespace rb {

typedef float2 cf;

RB_HD cf cmake(float re, float im) { cf r; r.x = re; r.y = im; return r; }
#if defined(__CUDA_ARCH__)
// sm_100a packed fp32 pairs: one FADD2 / FMUL2 / FFMA2 per complex add / real scale / real-coefficient FMA.  They retire the
// same flops per clock as the scalar forms but take half the issue slots (profiles/ubench_fma_pipe.txt), which is what
// the issue-bound FFT kernels need; ptxas folds the (s,s) broadcast of a constant coefficient into an immediate.
typedef unsigned long long pk64;
__device__ __forceinline__ pk64 pk(cf a) { pk64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a.x), "f"(a.y)); return r; }
__device__ __forceinline__ cf unpk(pk64 v) { cf a; asm("mov.b64 {%0,%1}, %2;" : "=f"(a.x), "=f"(a.y) : "l"(v)); return a; }
__device__ __forceinline__ cf cadd(cf a, cf b) { pk64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(b))); return unpk(r); }
__device__ __forceinline__ cf csub(cf a, cf b) { pk64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(b))); return unpk(r); }
__device__ __forceinline__ cf cscale(cf a, float s) { pk64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(cmake(s, s)))); return unpk(r); }
// a + s*b (real s)
__device__ __forceinline__ cf cfma(float s, cf b, cf a) { pk64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pk(cmake(s, s))), "l"(pk(b)), "l"(pk(a))); return unpk(r); }
#else
inline cf cadd(cf a, cf b) { return cmake(a.x + b.x, a.y + b.y); }
inline cf csub(cf a, cf b) { return cmake(a.x - b.x, a.y - b.y); }
inline cf cscale(cf a, float s) { return cmake(a.x * s, a.y * s); }
// a + s*b (real s); fused like the device form so the host self-check rounds identically
inline cf cfma(float s, cf b, cf a) { return cmake(fmaf(s, b.x, a.x), fmaf(s, b.y, a.y)); }
#endif
RB_HD cf cmul(cf a, cf b) { return cmake(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
// multiply by -i : (x + iy)(-i) = y - ix
RB_HD cf cmul_mi(cf a) { return cmake(a.y, -a.x); }
// out = m - i*n  and  m + i*n
RB_HD cf c_m_minus_i_n(cf m, cf n) { return cmake(m.x + n.y, m.y - n.x); }
RB_HD cf c_m_plus_i_n(cf m, cf n) { return cmake(m.x - n.y, m.y + n.x); }

// -i*(a - b) = (a.y - b.y, b.x - a.x): the odd parts of the symmetric butterflies are formed already rotated, so the
// conjugate output pairs m -/+ i*n become one packed add and one packed subtract instead of four scalar adds
RB_HD cf csub_mi(cf a, cf b) { return cmake(a.y - b.y, b.x - a.x); }

template <int R> struct Dft;

template <> struct Dft<1> { RB_HD static void run(cf*) {} };

template <> struct Dft<2> {
    RB_HD static void run(cf* v) {
        cf a = v[0], b = v[1];
        v[0] = cadd(a, b);
        v[1] = csub(a, b);
    }
};

template <> struct Dft<3> {
    RB_HD static void run(cf* v) {
        const float S3 = 0.86602540378443864676f;   // sin(2pi/3)
        cf t = cadd(v[1], v[2]);
        cf u = cscale(csub_mi(v[1], v[2]), S3);       // -i * S3 * (v1 - v2)
        cf m = cfma(-0.5f, t, v[0]);
        v[0] = cadd(v[0], t);
        v[1] = cadd(m, u);
        v[2] = csub(m, u);
    }
};

template <> struct Dft<4> {
    RB_HD static void run(cf* v) {
        cf s02 = cadd(v[0], v[2]), d02 = csub(v[0], v[2]);
        cf s13 = cadd(v[1], v[3]), d13 = csub(v[1], v[3]);
        v[0] = cadd(s02, s13);
        v[2] = csub(s02, s13);
        v[1] = c_m_minus_i_n(d02, d13);
        v[3] = c_m_plus_i_n(d02, d13);
    }
};

template <> struct Dft<5> {
    RB_HD static void run(cf* v) {
        const float C1 = 0.30901699437494742410f;    // cos(2pi/5)
        const float C2 = -0.80901699437494742410f;   // cos(4pi/5)
        const float S1 = 0.95105651629515357212f;    // sin(2pi/5)
        const float S2 = 0.58778525229247312917f;    // sin(4pi/5)
        cf t1 = cadd(v[1], v[4]), t2 = cadd(v[2], v[3]);
        cf u1 = csub_mi(v[1], v[4]), u2 = csub_mi(v[2], v[3]);      // already multiplied by -i
        cf m1 = cfma(C2, t2, cfma(C1, t1, v[0]));
        cf m2 = cfma(C1, t2, cfma(C2, t1, v[0]));
        cf n1 = cfma(S2, u2, cscale(u1, S1));
        cf n2 = cfma(-S1, u2, cscale(u1, S2));
        v[0] = cadd(v[0], cadd(t1, t2));
        v[1] = cadd(m1, n1);
        v[4] = csub(m1, n1);
        v[2] = cadd(m2, n2);
        v[3] = csub(m2, n2);
    }
};

template <> struct Dft<7> {
    RB_HD static void run(cf* v) {
        const float C1 = 0.62348980185873353053f;    // cos(2pi/7)
        const float C2 = -0.22252093395631440429f;   // cos(4pi/7)
        const float C3 = -0.90096886790241912624f;   // cos(6pi/7)
        const float S1 = 0.78183148246802980871f;    // sin(2pi/7)
        const float S2 = 0.97492791218182360702f;    // sin(4pi/7)
        const float S3 = 0.43388373911755812048f;    // sin(6pi/7)
        cf t1 = cadd(v[1], v[6]), t2 = cadd(v[2], v[5]), t3 = cadd(v[3], v[4]);
        cf u1 = csub_mi(v[1], v[6]), u2 = csub_mi(v[2], v[5]), u3 = csub_mi(v[3], v[4]);   // already multiplied by -i
        // cos(2pi jk/7): k=1: C1 C2 C3 ; k=2: C2 C3 C1 ; k=3: C3 C1 C2
        cf m1 = cfma(C3, t3, cfma(C2, t2, cfma(C1, t1, v[0])));
        cf m2 = cfma(C1, t3, cfma(C3, t2, cfma(C2, t1, v[0])));
        cf m3 = cfma(C2, t3, cfma(C1, t2, cfma(C3, t1, v[0])));
        // sin(2pi jk/7): k=1: S1 S2 S3 ; k=2: S2 -S3 -S1 ; k=3: S3 -S1 S2
        cf n1 = cfma(S3, u3, cfma(S2, u2, cscale(u1, S1)));
        cf n2 = cfma(-S1, u3, cfma(-S3, u2, cscale(u1, S2)));
        cf n3 = cfma(S2, u3, cfma(-S1, u2, cscale(u1, S3)));
        v[0] = cadd(v[0], cadd(t1, cadd(t2, t3)));
        v[1] = cadd(m1, n1);
        v[6] = csub(m1, n1);
        v[2] = cadd(m2, n2);
        v[5] = csub(m2, n2);
        v[3] = cadd(m3, n3);
        v[4] = csub(m3, n3);
    }
};

// Odd primes 11 and 13: the same symmetric form, driven by small cos/sin tables.
template <int P> struct OddPrimeTables;
template <> struct OddPrimeTables<11> {
    RB_HD static float c(int i) {
        const float t[11] = {1.0f, 0.84125353283118116886f, 0.41541501300188642553f, -0.14231483827328514044f,
                             -0.65486073394528506406f, -0.95949297361449738989f, -0.95949297361449738989f,
                             -0.65486073394528506406f, -0.14231483827328514044f, 0.41541501300188642553f,
                             0.84125353283118116886f};
        return t[i];
    }
    RB_HD static float s(int i) {
        const float t[11] = {0.0f, 0.54064081745559758211f, 0.90963199535451837141f, 0.98982144188093273238f,
                             0.75574957435425828377f, 0.28173255684142969771f, -0.28173255684142969771f,
                             -0.75574957435425828377f, -0.98982144188093273238f, -0.90963199535451837141f,
                             -0.54064081745559758211f};
        return t[i];
    }
};
template <> struct OddPrimeTables<13> {
    RB_HD static float c(int i) {
        const float t[13] = {1.0f, 0.88545602565320989590f, 0.56806474673115580251f, 0.12053668025532305335f,
                             -0.35460488704253562597f, -0.74851074817110109863f, -0.97094181742605202716f,
                             -0.97094181742605202716f, -0.74851074817110109863f, -0.35460488704253562597f,
                             0.12053668025532305335f, 0.56806474673115580251f, 0.88545602565320989590f};
        return t[i];
    }
    RB_HD static float s(int i) {
        const float t[13] = {0.0f, 0.46472317204376854566f, 0.82298386589365639458f, 0.99270887409805399280f,
                             0.93501624268541482344f, 0.66312265824079520238f, 0.23931566428755776715f,
                             -0.23931566428755776715f, -0.66312265824079520238f, -0.93501624268541482344f,
                             -0.99270887409805399280f, -0.82298386589365639458f, -0.46472317204376854566f};
        return t[i];
    }
};

template <int P> struct DftOddPrime {
    RB_HD static void run(cf* v) {
        constexpr int H = (P - 1) / 2;
        cf t[H], u[H];
#pragma unroll
        for (int j = 0; j < H; ++j) {
            t[j] = cadd(v[j + 1], v[P - 1 - j]);
            u[j] = csub_mi(v[j + 1], v[P - 1 - j]);      // already multiplied by -i
        }
        cf x0 = v[0];
        cf sum = x0;
#pragma unroll
        for (int j = 0; j < H; ++j) sum = cadd(sum, t[j]);
        v[0] = sum;
#pragma unroll
        for (int k = 1; k <= H; ++k) {
            cf m = x0, n = cmake(0.f, 0.f);
#pragma unroll
            for (int j = 1; j <= H; ++j) {
                const int idx = (j * k) % P;
                m = cfma(OddPrimeTables<P>::c(idx), t[j - 1], m);
                n = cfma(OddPrimeTables<P>::s(idx), u[j - 1], n);
            }
            v[k] = cadd(m, n);
            v[P - k] = csub(m, n);
        }
    }
};
template <> struct Dft<11> : DftOddPrime<11> {};
template <> struct Dft<13> : DftOddPrime<13> {};

// ---- composites via in-register Cooley-Tukey ------------------------------------------------

// Coprime composites by the prime-factor (Good-Thomas) map: no internal twiddles.  With N = N1*N2,
// input n = (N2*n1 + N1*n2) mod N and output k = (E1*k1 + E2*k2) mod N where E1 = 1 (mod N1), 0 (mod N2) and
// E2 = 0 (mod N1), 1 (mod N2); every index is a compile-time constant after unrolling, so the permutations cost nothing.
constexpr int pfa_inv(int a, int m) { for (int x = 1; x < m; ++x) if ((a * x) % m == 1) return x; return 1; }
template <int N1, int N2> struct DftPfa {
    RB_HD static void run(cf* v) {
        constexpr int N = N1 * N2;
        constexpr int E1 = N2 * pfa_inv(N2 % N1, N1), E2 = N1 * pfa_inv(N1 % N2, N2);
        cf a[N2][N1];
#pragma unroll
        for (int n2 = 0; n2 < N2; ++n2) {
#pragma unroll
            for (int n1 = 0; n1 < N1; ++n1) a[n2][n1] = v[(N2 * n1 + N1 * n2) % N];
            Dft<N1>::run(a[n2]);
        }
#pragma unroll
        for (int k1 = 0; k1 < N1; ++k1) {
            cf b[N2];
#pragma unroll
            for (int n2 = 0; n2 < N2; ++n2) b[n2] = a[n2][k1];
            Dft<N2>::run(b);
#pragma unroll
            for (int k2 = 0; k2 < N2; ++k2) v[(E1 * k1 + E2 * k2) % N] = b[k2];
        }
    }
};
template <> struct Dft<6> : DftPfa<3, 2> {};

template <> struct Dft<8> {
    RB_HD static void run(cf* v) {
        const float Q = 0.70710678118654752440f;
        cf e[4] = {v[0], v[2], v[4], v[6]};
        cf o[4] = {v[1], v[3], v[5], v[7]};
        Dft<4>::run(e);
        Dft<4>::run(o);
        cf o1 = cmake(Q * (o[1].x + o[1].y), Q * (o[1].y - o[1].x));     // * (1 - i)/sqrt2
        cf o2 = cmul_mi(o[2]);                                           // * -i
        cf o3 = cmake(Q * (o[3].y - o[3].x), -Q * (o[3].x + o[3].y));    // * (-1 - i)/sqrt2
        v[0] = cadd(e[0], o[0]);  v[4] = csub(e[0], o[0]);
        v[1] = cadd(e[1], o1);    v[5] = csub(e[1], o1);
        v[2] = cadd(e[2], o2);    v[6] = csub(e[2], o2);
        v[3] = cadd(e[3], o3);    v[7] = csub(e[3], o3);
    }
};

template <> struct Dft<9> {
    RB_HD static void run(cf* v) {
        // n = 3*n1 + n2 ; k = k1 + 3*k2
        const cf W1 = cmake(0.76604444311897803520f, -0.64278760968653932632f);   // W9^1
        const cf W2 = cmake(0.17364817766693034885f, -0.98480775301220805937f);   // W9^2
        const cf W4 = cmake(-0.93969262078590838405f, -0.34202014332566873304f);  // W9^4
        cf a[3] = {v[0], v[3], v[6]};
        cf b[3] = {v[1], v[4], v[7]};
        cf c[3] = {v[2], v[5], v[8]};
        Dft<3>::run(a);
        Dft<3>::run(b);
        Dft<3>::run(c);
        b[1] = cmul(b[1], W1);  b[2] = cmul(b[2], W2);
        c[1] = cmul(c[1], W2);  c[2] = cmul(c[2], W4);
#pragma unroll
        for (int k1 = 0; k1 < 3; ++k1) {
            cf y[3] = {a[k1], b[k1], c[k1]};
            Dft<3>::run(y);
            v[k1] = y[0];
            v[k1 + 3] = y[1];
            v[k1 + 6] = y[2];
        }
    }
};

template <> struct Dft<10> : DftPfa<5, 2> {};
template <> struct Dft<12> : DftPfa<3, 4> {};
template <> struct Dft<14> : DftPfa<7, 2> {};
template <> struct Dft<15> : DftPfa<5, 3> {};
template <> struct Dft<20> : DftPfa<5, 4> {};
template <> struct Dft<30> : DftPfa<5, 6> {};
// whole range / Doppler axes of the register-resident kernels (rd_reg.cuh, sar_range_fft_reg_kernel)
template <> struct Dft<63> : DftPfa<7, 9> {};
template <> struct Dft<70> : DftPfa<7, 10> {};
template <> struct Dft<40> : DftPfa<5, 8> {};
template <> struct Dft<35> : DftPfa<5, 7> {};
template <> struct Dft<45> : DftPfa<5, 9> {};

template <> struct Dft<16> {
    RB_HD static void run(cf* v) {
        // n = 4*n1 + n2 ; k = k1 + 4*k2
        const float C1 = 0.92387953251128675613f, S1 = 0.38268343236508977173f;   // cos/sin(pi/8)
        const float Q = 0.70710678118654752440f;
        cf g[4][4];
#pragma unroll
        for (int n2 = 0; n2 < 4; ++n2) {
#pragma unroll
            for (int n1 = 0; n1 < 4; ++n1) g[n2][n1] = v[4 * n1 + n2];
            Dft<4>::run(g[n2]);
        }
        // twiddles W16^(n2*k1)
        g[1][1] = cmul(g[1][1], cmake(C1, -S1));
        g[1][2] = cmul(g[1][2], cmake(Q, -Q));
        g[1][3] = cmul(g[1][3], cmake(S1, -C1));
        g[2][1] = cmul(g[2][1], cmake(Q, -Q));
        g[2][2] = cmul_mi(g[2][2]);
        g[2][3] = cmul(g[2][3], cmake(-Q, -Q));
        g[3][1] = cmul(g[3][1], cmake(S1, -C1));
        g[3][2] = cmul(g[3][2], cmake(-Q, -Q));
        g[3][3] = cmul(g[3][3], cmake(-C1, S1));
#pragma unroll
        for (int k1 = 0; k1 < 4; ++k1) {
            cf y[4] = {g[0][k1], g[1][k1], g[2][k1], g[3][k1]};
            Dft<4>::run(y);
#pragma unroll
            for (int k2 = 0; k2 < 4; ++k2) v[k1 + 4 * k2] = y[k2];
        }
    }
};

// 64 = 8 x 8 by Cooley-Tukey in registers: n = 8*n1 + n2, k = k1 + 8*k2, twiddles W64^(n2*k1) as literals
template <> struct Dft<64> {
    RB_HD static void run(cf* v) {
        cf a[8][8];
#pragma unroll
        for (int n2 = 0; n2 < 8; ++n2) {
#pragma unroll
            for (int n1 = 0; n1 < 8; ++n1) a[n2][n1] = v[8 * n1 + n2];
            Dft<8>::run(a[n2]);
        }
        a[1][1] = cmul(a[1][1], cmake(0.99518472667219692873f, -0.09801714032956060363f));   // W64^1
        a[1][2] = cmul(a[1][2], cmake(0.98078528040323043058f, -0.19509032201612824808f));   // W64^2
        a[1][3] = cmul(a[1][3], cmake(0.95694033573220882438f, -0.29028467725446233105f));   // W64^3
        a[1][4] = cmul(a[1][4], cmake(0.92387953251128673848f, -0.38268343236508978178f));   // W64^4
        a[1][5] = cmul(a[1][5], cmake(0.88192126434835504956f, -0.47139673682599764204f));   // W64^5
        a[1][6] = cmul(a[1][6], cmake(0.83146961230254523567f, -0.55557023301960217765f));   // W64^6
        a[1][7] = cmul(a[1][7], cmake(0.77301045336273699338f, -0.63439328416364548779f));   // W64^7
        a[2][1] = cmul(a[2][1], cmake(0.98078528040323043058f, -0.19509032201612824808f));   // W64^2
        a[2][2] = cmul(a[2][2], cmake(0.92387953251128673848f, -0.38268343236508978178f));   // W64^4
        a[2][3] = cmul(a[2][3], cmake(0.83146961230254523567f, -0.55557023301960217765f));   // W64^6
        a[2][4] = cmul(a[2][4], cmake(0.70710678118654757274f, -0.70710678118654746172f));   // W64^8
        a[2][5] = cmul(a[2][5], cmake(0.55557023301960228867f, -0.83146961230254523567f));   // W64^10
        a[2][6] = cmul(a[2][6], cmake(0.38268343236508983729f, -0.92387953251128673848f));   // W64^12
        a[2][7] = cmul(a[2][7], cmake(0.19509032201612833135f, -0.98078528040323043058f));   // W64^14
        a[3][1] = cmul(a[3][1], cmake(0.95694033573220882438f, -0.29028467725446233105f));   // W64^3
        a[3][2] = cmul(a[3][2], cmake(0.83146961230254523567f, -0.55557023301960217765f));   // W64^6
        a[3][3] = cmul(a[3][3], cmake(0.63439328416364548779f, -0.77301045336273699338f));   // W64^9
        a[3][4] = cmul(a[3][4], cmake(0.38268343236508983729f, -0.92387953251128673848f));   // W64^12
        a[3][5] = cmul(a[3][5], cmake(0.09801714032956077016f, -0.99518472667219681771f));   // W64^15
        a[3][6] = cmul(a[3][6], cmake(-0.19509032201612819257f, -0.98078528040323043058f));   // W64^18
        a[3][7] = cmul(a[3][7], cmake(-0.47139673682599769755f, -0.88192126434835504956f));   // W64^21
        a[4][1] = cmul(a[4][1], cmake(0.92387953251128673848f, -0.38268343236508978178f));   // W64^4
        a[4][2] = cmul(a[4][2], cmake(0.70710678118654757274f, -0.70710678118654746172f));   // W64^8
        a[4][3] = cmul(a[4][3], cmake(0.38268343236508983729f, -0.92387953251128673848f));   // W64^12
        a[4][4] = cmul_mi(a[4][4]);
        a[4][5] = cmul(a[4][5], cmake(-0.38268343236508972627f, -0.92387953251128673848f));   // W64^20
        a[4][6] = cmul(a[4][6], cmake(-0.70710678118654746172f, -0.70710678118654757274f));   // W64^24
        a[4][7] = cmul(a[4][7], cmake(-0.92387953251128673848f, -0.38268343236508989280f));   // W64^28
        a[5][1] = cmul(a[5][1], cmake(0.88192126434835504956f, -0.47139673682599764204f));   // W64^5
        a[5][2] = cmul(a[5][2], cmake(0.55557023301960228867f, -0.83146961230254523567f));   // W64^10
        a[5][3] = cmul(a[5][3], cmake(0.09801714032956077016f, -0.99518472667219681771f));   // W64^15
        a[5][4] = cmul(a[5][4], cmake(-0.38268343236508972627f, -0.92387953251128673848f));   // W64^20
        a[5][5] = cmul(a[5][5], cmake(-0.77301045336273699338f, -0.63439328416364548779f));   // W64^25
        a[5][6] = cmul(a[5][6], cmake(-0.98078528040323043058f, -0.19509032201612860891f));   // W64^30
        a[5][7] = cmul(a[5][7], cmake(-0.95694033573220893540f, 0.29028467725446210901f));   // W64^35
        a[6][1] = cmul(a[6][1], cmake(0.83146961230254523567f, -0.55557023301960217765f));   // W64^6
        a[6][2] = cmul(a[6][2], cmake(0.38268343236508983729f, -0.92387953251128673848f));   // W64^12
        a[6][3] = cmul(a[6][3], cmake(-0.19509032201612819257f, -0.98078528040323043058f));   // W64^18
        a[6][4] = cmul(a[6][4], cmake(-0.70710678118654746172f, -0.70710678118654757274f));   // W64^24
        a[6][5] = cmul(a[6][5], cmake(-0.98078528040323043058f, -0.19509032201612860891f));   // W64^30
        a[6][6] = cmul(a[6][6], cmake(-0.92387953251128684951f, 0.38268343236508967076f));   // W64^36
        a[6][7] = cmul(a[6][7], cmake(-0.55557023301960217765f, 0.83146961230254523567f));   // W64^42
        a[7][1] = cmul(a[7][1], cmake(0.77301045336273699338f, -0.63439328416364548779f));   // W64^7
        a[7][2] = cmul(a[7][2], cmake(0.19509032201612833135f, -0.98078528040323043058f));   // W64^14
        a[7][3] = cmul(a[7][3], cmake(-0.47139673682599769755f, -0.88192126434835504956f));   // W64^21
        a[7][4] = cmul(a[7][4], cmake(-0.92387953251128673848f, -0.38268343236508989280f));   // W64^28
        a[7][5] = cmul(a[7][5], cmake(-0.95694033573220893540f, 0.29028467725446210901f));   // W64^35
        a[7][6] = cmul(a[7][6], cmake(-0.55557023301960217765f, 0.83146961230254523567f));   // W64^42
        a[7][7] = cmul(a[7][7], cmake(0.09801714032956009015f, 0.99518472667219692873f));   // W64^49
#pragma unroll
        for (int k1 = 0; k1 < 8; ++k1) {
            cf y[8] = {a[0][k1], a[1][k1], a[2][k1], a[3][k1], a[4][k1], a[5][k1], a[6][k1], a[7][k1]};
            Dft<8>::run(y);
#pragma unroll
            for (int k2 = 0; k2 < 8; ++k2) v[k1 + 8 * k2] = y[k2];
        }
    }
};

// 32 = 2 x 16: decimation in time over two register-resident 16-point transforms, twiddles W32^k as literals
template <> struct Dft<32> {
    RB_HD static void run(cf* v) {
        cf e[16], o[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { e[i] = v[2 * i]; o[i] = v[2 * i + 1]; }
        Dft<16>::run(e);
        Dft<16>::run(o);
        o[1] = cmul(o[1], cmake(0.98078528040323043058f, -0.19509032201612824808f));
        o[2] = cmul(o[2], cmake(0.92387953251128673848f, -0.38268343236508978178f));
        o[3] = cmul(o[3], cmake(0.83146961230254523567f, -0.55557023301960217765f));
        o[4] = cmul(o[4], cmake(0.70710678118654757274f, -0.70710678118654746172f));
        o[5] = cmul(o[5], cmake(0.55557023301960228867f, -0.83146961230254523567f));
        o[6] = cmul(o[6], cmake(0.38268343236508983729f, -0.92387953251128673848f));
        o[7] = cmul(o[7], cmake(0.19509032201612833135f, -0.98078528040323043058f));
        o[8] = cmul_mi(o[8]);
        o[9] = cmul(o[9], cmake(-0.19509032201612819257f, -0.98078528040323043058f));
        o[10] = cmul(o[10], cmake(-0.38268343236508972627f, -0.92387953251128673848f));
        o[11] = cmul(o[11], cmake(-0.55557023301960195560f, -0.83146961230254545772f));
        o[12] = cmul(o[12], cmake(-0.70710678118654746172f, -0.70710678118654757274f));
        o[13] = cmul(o[13], cmake(-0.83146961230254534669f, -0.55557023301960217765f));
        o[14] = cmul(o[14], cmake(-0.92387953251128673848f, -0.38268343236508989280f));
        o[15] = cmul(o[15], cmake(-0.98078528040323043058f, -0.19509032201612860891f));
#pragma unroll
        for (int k = 0; k < 16; ++k) { v[k] = cadd(e[k], o[k]); v[k + 16] = csub(e[k], o[k]); }
    }
};

}  // namespace rb
