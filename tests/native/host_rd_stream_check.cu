// CPU emulation of the streaming range-Doppler kernel (rd_stream.cuh): the same __host__ __device__ phases the
// device runs, threads emulated by a loop, __syncthreads() by the phase boundaries.  Checks index maps, the
// prime-factor butterflies and the fftshift against a direct fp64 DFT of the windowed cube.
// Build: nvcc -O2 -I mmwave_radar_processing_b200/csrc tests/native/host_rd_stream_check.cu -o build/host_rd_stream_check
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "rd_stream.cuh"

using namespace rb;

static std::vector<float> hann(int n) {
    std::vector<float> w(n, 1.0f);
    if (n > 1) for (int k = 0; k < n; ++k) w[k] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * k / (n - 1)));
    return w;
}

template <class C>
static double check(int R, int cpl, int n_frames, int mag_antenna) {
    constexpr int S = C::S, L = C::L;
    const int V = R * cpl, Cc = cpl * L;
    std::vector<uint32_t> iq((size_t)n_frames * Cc * R * S);
    unsigned s = 99u;
    for (auto& w : iq) {
        s = s * 1664525u + 1013904223u; int re = (int)((s >> 16) % 4001) - 2000;
        s = s * 1664525u + 1013904223u; int im = (int)((s >> 16) % 4001) - 2000;
        w = ((uint32_t)(uint16_t)(int16_t)re) | ((uint32_t)(uint16_t)(int16_t)im << 16);
    }
    RdStreamParams<C> a;
    std::vector<cf> rd((size_t)n_frames * V * S * L);
    std::vector<float> mag((size_t)n_frames * (mag_antenna < 0 ? V : 1) * S * L, -1.0f);
    a.iq = iq.data(); a.rd_out = rd.data(); a.mag_out = mag.data(); a.mag_antenna = mag_antenna;
    a.R = R; a.cpl = cpl; a.V = V; a.items_per_frame = V / C::G; a.n_items = (long long)n_frames * a.items_per_frame;
    auto wl = hann(L), ws = hann(S);
    for (int i = 0; i < L; ++i) a.winL[i] = wl[i];
    for (int i = 0; i < S; ++i) a.winS[i] = ws[i];
    for (int n2 = 0; n2 < C::RS2; ++n2) for (int k = 0; k < C::RS1; ++k) {
        double ang = -2.0 * M_PI * ((n2 * k) % S) / S; a.twS[n2 * C::RS1 + k] = cmake((float)cos(ang), (float)sin(ang)); }
    for (int n2 = 0; n2 < C::DL2; ++n2) for (int k = 0; k < C::DL1; ++k) {
        double ang = -2.0 * M_PI * ((n2 * k) % L) / L; a.twD[n2 * C::DL1 + k] = cmake((float)cos(ang), (float)sin(ang)); }
    std::vector<cf> tiles((size_t)C::NBUF * C::G * C::TILE);
    const long long frame_words = (long long)C::ROW * L;
    struct Regs { uint32_t w[C::RAW_WORDS]; };
    std::vector<Regs> regs(C::NT);
    auto item_src = [&](long long item) {
        const long long frame = item / a.items_per_frame;
        const int v0 = (int)(item - frame * a.items_per_frame) * C::G;
        const int tx = v0 / R, rx0 = v0 - tx * R;
        return a.iq + frame * frame_words + ((long long)tx * R + rx0) * S;
    };
    int buf = 0;
    struct Halves { uint32_t w0[C::DL1], w1[C::DL1]; };
    std::vector<Halves> hv(C::NT);
    if constexpr (C::DF > 1) { for (int tid = 0; tid < C::NT; ++tid) rd_stream_load_half<C, 0>(tid, item_src(0), hv[tid].w0); }
    else { for (int tid = 0; tid < C::NT; ++tid) rd_stream_load_raw<C>(tid, item_src(0), regs[tid].w); }
    for (long long item = 0; item < a.n_items; ++item) {
        cf* tile = tiles.data() + (size_t)buf * C::G * C::TILE;
        const long long frame = item / a.items_per_frame;
        const int v0 = (int)(item - frame * a.items_per_frame) * C::G;
        if constexpr (C::DF > 1) {
            for (int tid = 0; tid < C::NT; ++tid) {
                rd_stream_load_half<C, 1>(tid, item_src(item), hv[tid].w1);
                rd_stream_doppler_half<C, 0>(a, tid, hv[tid].w0, tile, ws.data());
                rd_stream_doppler_half<C, 1>(a, tid, hv[tid].w1, tile, ws.data());
                if (item + 1 < a.n_items) rd_stream_load_half<C, 0>(tid, item_src(item + 1), hv[tid].w0);
            }
            for (int tid = 0; tid < C::NT; ++tid) rd_stream_range_s0_fused<C>(a, tid, tile);
        } else {
            for (int tid = 0; tid < C::NT; ++tid) rd_stream_doppler_s0<C>(a, tid, regs[tid].w, tile, ws.data());
            if (item + 1 < a.n_items) for (int tid = 0; tid < C::NT; ++tid) rd_stream_load_raw<C>(tid, item_src(item + 1), regs[tid].w);
            if (C::DL2 > 1) for (int tid = 0; tid < C::NT; ++tid) rd_stream_doppler_s1<C>(tid, tile, ws.data());
            for (int tid = 0; tid < C::NT; ++tid) rd_stream_range_s0<C>(a, tid, tile);
        }
        const long long out_base = (frame * V + v0) * (long long)(S * L);
        float* m = a.mag_out; long long mag_base = out_base; int mag_g = -1;
        if (m && mag_antenna >= 0) {
            mag_g = mag_antenna - v0;
            if (mag_g < 0 || mag_g >= C::G) m = nullptr;
            mag_base = frame * (long long)(S * L);
        }
        for (int tid = 0; tid < C::NT; ++tid) rd_stream_range_s1<C>(tid, tile, a.rd_out, out_base, m, mag_base, mag_g);
        if (C::NBUF == 2) buf ^= 1;
    }
    // direct fp64 reference
    double num = 0, den = 0, mnum = 0, mden = 0;
    std::vector<double> xr((size_t)S * L), xi((size_t)S * L), tr((size_t)S * L), ti((size_t)S * L);
    for (int f = 0; f < n_frames; ++f) for (int v = 0; v < V; ++v) {
        const int tx = v / R, rx = v % R;
        for (int sm = 0; sm < S; ++sm) for (int l = 0; l < L; ++l) {
            uint32_t w = iq[(((size_t)f * Cc + (size_t)l * cpl + tx) * R + rx) * S + sm];
            double wgt = (double)ws[sm] * (double)wl[l];
            xr[sm * L + l] = (double)(int16_t)(w & 0xffff) * wgt; xi[sm * L + l] = (double)(int16_t)(w >> 16) * wgt;
        }
        for (int k = 0; k < S; ++k) for (int l = 0; l < L; ++l) {            // range DFT
            double sr = 0, si = 0;
            for (int sm = 0; sm < S; ++sm) { double ang = -2.0 * M_PI * ((k * sm) % S) / S; double c = cos(ang), sn = sin(ang);
                sr += xr[sm * L + l] * c - xi[sm * L + l] * sn; si += xr[sm * L + l] * sn + xi[sm * L + l] * c; }
            tr[k * L + l] = sr; ti[k * L + l] = si;
        }
        for (int k = 0; k < S; ++k) for (int d = 0; d < L; ++d) {            // Doppler DFT + shift
            const int fb = ((d - L / 2) % L + L) % L;
            double sr = 0, si = 0;
            for (int l = 0; l < L; ++l) { double ang = -2.0 * M_PI * ((fb * l) % L) / L; double c = cos(ang), sn = sin(ang);
                sr += tr[k * L + l] * c - ti[k * L + l] * sn; si += tr[k * L + l] * sn + ti[k * L + l] * c; }
            const cf got = rd[(((size_t)f * V + v) * S + k) * L + d];
            num += (got.x - sr) * (got.x - sr) + (got.y - si) * (got.y - si); den += sr * sr + si * si;
            if (mag_antenna < 0 || mag_antenna == v) {
                const float gm = mag_antenna < 0 ? mag[(((size_t)f * V + v) * S + k) * L + d] : mag[((size_t)f * S + k) * L + d];
                const double rm = sqrt(sr * sr + si * si);
                mnum += (gm - rm) * (gm - rm); mden += rm * rm;
            }
        }
    }
    const double err = sqrt(num / den), merr = sqrt(mnum / mden);
    printf("S=%d L=%d split %dx%d / %dx%d(x%d) G=%d NT=%d R=%d cpl=%d mag_ant=%d: rd rel_l2=%.3e mag rel_l2=%.3e\n", S, L, C::RS1, C::RS2,
           C::DL1, C::DL2, C::DF, C::G, C::NT, R, cpl, mag_antenna, err, merr);
    return err > merr ? err : merr;
}

int main() {
    double worst = 0;
    auto upd = [&](double e) { if (e > worst) worst = e; };
#define RB_STREAM(var, s, l, cr, rs1, rs2, dl1, dl2, g, nt, minb, df) \
    printf("variant %d: ", var); upd(check<RdStreamShape<s, l, cr, rs1, rs2, dl1, dl2, g, nt, minb, df>>(4, cr / 4, 2, 0)); \
    upd(check<RdStreamShape<s, l, cr, rs1, rs2, dl1, dl2, g, nt, minb, df>>(4, cr / 4, 1, -1)); \
    upd(check<RdStreamShape<s, l, cr, rs1, rs2, dl1, dl2, g, nt, minb, df>>(4, cr / 4, 1, 5));
#include "rd_stream_shapes.inc"
#undef RB_STREAM
    printf("worst rel_l2 = %.3e\n", worst);
    return worst < 2e-6 ? 0 : 1;
}
