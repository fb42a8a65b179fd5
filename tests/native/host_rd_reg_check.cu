// CPU emulation of the register-resident range-Doppler kernel (rd_reg.cuh): the same __host__ __device__ phases the device
// runs - raw ring stage filled the way the TMA warp fills it (16-byte aligned window around the item's run), phase A per
// (column, residue class), phase B per (antenna, Doppler position) - checked against a direct fp64 DFT of the windowed cube.
// Build: nvcc -O2 -std=c++17 --expt-relaxed-constexpr -I mmwave_radar_processing_b200/csrc tests/native/host_rd_reg_check.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "rd_reg.cuh"

using namespace rb;

template <class C>
static double check(int n_frames) {
    constexpr int S = C::S, L = C::L, V = C::CR, IPF = C::CR / C::G;
    const size_t frame_words = (size_t)C::ROW * L;
    std::vector<uint32_t> iq(n_frames * frame_words + 8);
    unsigned s = 12345u;
    for (auto& w : iq) {
        s = s * 1664525u + 1013904223u; int re = (int)((s >> 16) % 4001) - 2000;
        s = s * 1664525u + 1013904223u; int im = (int)((s >> 16) % 4001) - 2000;
        w = ((uint32_t)(uint16_t)(int16_t)re) | ((uint32_t)(uint16_t)(int16_t)im << 16);
    }
    RdRegParams<C> a;
    std::vector<double> wl(L), ws(S);
    for (int k = 0; k < L; ++k) wl[k] = 0.5 - 0.5 * cos(2.0 * M_PI * k / (L - 1));
    for (int k = 0; k < S; ++k) ws[k] = 0.5 - 0.5 * cos(2.0 * M_PI * k / (S - 1));
    for (int h = 0; h < C::DF; ++h)
        for (int m = 0; m < C::LH; ++m) { const float w = (float)wl[rd_reg_doppler_input<C>(h, m)]; a.winL[h * C::LH + m] = make_float2(w, w); }
    if (C::TW) for (int k = 0; k < L; ++k) a.twL[k] = cmake((float)cos(-2.0 * M_PI * k / L), (float)sin(-2.0 * M_PI * k / L));
    for (int k = 0; k < S; ++k) a.winS[k] = (float)ws[k];
    std::vector<cf> rd((size_t)n_frames * V * S * L);
    std::vector<float> mag((size_t)n_frames * S * L, -1.0f);
    std::vector<uint32_t> stage(C::RAW_STAGE_W);
    std::vector<cf> tile(C::TILE);
    for (int item = 0; item < n_frames * IPF; ++item) {
        const int frame = item / IPF, grp = item % IPF;
        const unsigned off = grp * (C::G * S * 4), a0 = off & ~15u, bytes = ((off - a0) + C::G * S * 4 + 15u) & ~15u;
        const int shift = (int)(off & 15u) >> 2;
        for (int l = 0; l < L; ++l)
            memcpy(stage.data() + l * C::RAW_PITCH_W, (const unsigned char*)iq.data() + ((size_t)frame * frame_words + (size_t)l * C::ROW) * 4 + a0, bytes);
        for (int h = 0; h < C::DF; ++h)
            for (int col = 0; col < C::COLS_A; ++col) {
                cf v[C::LH];
                rd_reg_doppler_load<C>(a, stage.data(), shift + col, h, v);
                rd_reg_doppler_finish<C>(a.winS[col % S], v);
                rd_reg_doppler_twiddle<C>(a, h, v);
                rd_reg_store_row<C>(v, tile.data() + (size_t)col * C::P, h);
            }
        for (int cls = 0; cls < C::BF; ++cls)
            for (int col = 0; col < C::COLS_B; ++col) {
                const int g = col / L, d = col % L, ant = grp * C::G + g;
                const int ca = C::BF == 4 ? cls >> 1 : (C::BF == 2 ? cls : 0), odd = C::BF == 4 ? cls & 1 : 0;
                const long long half = (long long)ca * (S / 2) * L;
                cf* out = rd.data() + (((size_t)frame * V + ant) * S) * L + d;
                float* mg = ant == 0 ? mag.data() + (size_t)frame * S * L + d : nullptr;
                cf v[C::SH];
                if (odd) {
                    rd_reg_range_load<C, 1>(a, tile.data(), g, d, ca ? -1.0f : 1.0f, v);
                    rd_reg_range_store<C, 1>(v, out + half, out - half, mg ? mg + half : nullptr, mg ? mg - half : nullptr);
                } else {
                    rd_reg_range_load<C, 0>(a, tile.data(), g, d, ca ? -1.0f : 1.0f, v);
                    rd_reg_range_store<C, 0>(v, out + half, out - half, mg ? mg + half : nullptr, mg ? mg - half : nullptr);
                }
            }
    }
    // direct fp64 transform of (frame 0 .. n_frames-1, a few antennas)
    double worst = 0.0;
    for (int frame = 0; frame < n_frames; ++frame)
        for (int ant : {0, V / 2, V - 1}) {
            double num = 0.0, den = 0.0;
            std::vector<double> xr((size_t)S * L), xi((size_t)S * L);
            for (int l = 0; l < L; ++l)
                for (int sidx = 0; sidx < S; ++sidx) {
                    const uint32_t w = iq[(size_t)frame * frame_words + (size_t)l * C::ROW + (size_t)ant * S + sidx];
                    xr[(size_t)sidx * L + l] = (double)(short)(w & 0xffffu) * wl[l] * ws[sidx];
                    xi[(size_t)sidx * L + l] = (double)(short)(w >> 16) * wl[l] * ws[sidx];
                }
            for (int k = 0; k < S; k += 7)
                for (int d = 0; d < L; d += 3) {
                    const int kd = (d + L - L / 2) % L;
                    double ar = 0.0, ai = 0.0;
                    for (int sidx = 0; sidx < S; ++sidx)
                        for (int l = 0; l < L; ++l) {
                            const double ang = -2.0 * M_PI * ((double)((long long)k * sidx % S) / S + (double)((long long)kd * l % L) / L);
                            const double c = cos(ang), sn = sin(ang);
                            ar += xr[(size_t)sidx * L + l] * c - xi[(size_t)sidx * L + l] * sn;
                            ai += xr[(size_t)sidx * L + l] * sn + xi[(size_t)sidx * L + l] * c;
                        }
                    const cf got = rd[(((size_t)frame * V + ant) * S + k) * L + d];
                    num += (got.x - ar) * (got.x - ar) + (got.y - ai) * (got.y - ai);
                    den += ar * ar + ai * ai;
                    if (ant == 0) {
                        const double m = sqrt(ar * ar + ai * ai), gm = mag[((size_t)frame * S + k) * L + d];
                        if (fabs(gm - m) > 1e-5 * (m + 1.0)) { printf("magnitude mismatch at (%d,%d,%d): %g vs %g\n", frame, k, d, gm, m); exit(1); }
                    }
                }
            const double rel = sqrt(num / den);
            if (rel > worst) worst = rel;
        }
    return worst;
}

int main() {
    double worst = 0.0, r;
    r = check<RdRegShape<63, 70, 12, 2, 1, 3, 1>>(2); printf("63x70 G=2 DF=1: rel_l2 = %.3e\n", r); worst = fmax(worst, r);
    r = check<RdRegShape<63, 70, 12, 2, 2, 3, 1>>(2); printf("63x70 G=2 DF=2: rel_l2 = %.3e\n", r); worst = fmax(worst, r);
    r = check<RdRegShape<100, 30, 8, 2, 1, 3, 1, 4>>(2); printf("100x30 G=2 BF=4: rel_l2 = %.3e\n", r); worst = fmax(worst, r);
    r = check<RdRegShape<64, 40, 4, 2, 1, 3, 1, 1>>(2); printf("64x40 G=2: rel_l2 = %.3e\n", r); worst = fmax(worst, r);
    r = check<RdRegShape<90, 100, 8, 1, 2, 3, 1, 2>>(2); printf("90x100 G=1 DF=2 (twiddle) BF=2: rel_l2 = %.3e\n", r); worst = fmax(worst, r);
    printf("worst rel_l2 = %.3e\n", worst);
    return worst < 2e-6 ? 0 : 1;
}
