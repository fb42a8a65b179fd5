// C-ABI layer of libradar_b200.so: context life cycle, host-built tables, argument checks, and the
// extern "C" entry points declared in include/radar_b200.h.  No torch types cross this boundary.
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <new>
#include "radar_internal.h"

namespace rb {

static thread_local char g_create_error[512] = {0};

int set_error(radar_ctx* ctx, int code, const char* fmt, ...) {
    char* dst = ctx ? ctx->err : g_create_error;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
    return code;
}

template <typename T>
static int upload(radar_ctx* ctx, const std::vector<T>& host, T** dev) {
    RB_CUDA_OK(ctx, cudaMalloc(reinterpret_cast<void**>(dev), host.size() * sizeof(T)));
    RB_CUDA_OK(ctx, cudaMemcpy(*dev, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice));
    return RADAR_OK;
}

static std::vector<float> hann_window(int n) {
    // np.hanning: 0.5 - 0.5 cos(2 pi k/(n-1)), zero end points, hanning(1) = [1]
    std::vector<float> w((size_t)n);
    for (int k = 0; k < n; ++k)
        w[(size_t)k] = n == 1 ? 1.0f : (float)(0.5 - 0.5 * std::cos(2.0 * M_PI * (double)k / (double)(n - 1)));
    return w;
}

static int build_fft_tables(radar_ctx* ctx, int n, FftTables* t) {
    if (n > 65535) return set_error(ctx, RADAR_E_UNSUPPORTED, "FFT length %d exceeds 65535", n);
    if (!make_plan(n, &t->plan))
        return set_error(ctx, RADAR_E_UNSUPPORTED, "FFT length %d has a prime factor above %d", n, kMaxGenericRadix);
    std::vector<cf> tw((size_t)n);
    for (int j = 0; j < n; ++j) {
        const double ang = -2.0 * M_PI * (double)j / (double)n;
        tw[(size_t)j] = cmake((float)std::cos(ang), (float)std::sin(ang));
    }
    std::vector<uint16_t> pos((size_t)n);
    plan_positions(t->plan, pos.data());
    std::vector<float> ones((size_t)n, 1.0f);
    if (int rc = upload(ctx, tw, &t->tw)) return rc;
    if (int rc = upload(ctx, pos, &t->pos)) return rc;
    if (int rc = upload(ctx, hann_window(n), &t->hann)) return rc;
    if (int rc = upload(ctx, ones, &t->ones)) return rc;
    return RADAR_OK;
}

static void free_fft_tables(FftTables* t) {
    cudaFree(t->tw); cudaFree(t->pos); cudaFree(t->hann); cudaFree(t->ones);
    *t = FftTables{};
}

int ensure_scratch(radar_ctx* ctx, size_t bytes) {
    if (bytes <= ctx->scratch_bytes) return RADAR_OK;
    if (ctx->scratch) { RB_CUDA_OK(ctx, cudaFree(ctx->scratch)); ctx->scratch = nullptr; ctx->scratch_bytes = 0; }
    RB_CUDA_OK(ctx, cudaMalloc(&ctx->scratch, bytes));
    ctx->scratch_bytes = bytes;
    return RADAR_OK;
}

int get_angle_tables(radar_ctx* ctx, int num_angle_bins, AngleTables* out) {
    auto it = ctx->angle_tabs.find(num_angle_bins);
    if (it != ctx->angle_tabs.end()) { *out = it->second; return RADAR_OK; }
    const int n = num_angle_bins;
    std::vector<cf> tw((size_t)n);
    std::vector<double2> tw64((size_t)n);
    for (int j = 0; j < n; ++j) {
        const double ang = -2.0 * M_PI * (double)j / (double)n;
        tw64[(size_t)j] = make_double2(std::cos(ang), std::sin(ang));
        tw[(size_t)j] = cmake((float)std::cos(ang), (float)std::sin(ang));
    }
    AngleTables t;
    if (int rc = upload(ctx, tw, &t.tw)) return rc;
    if (int rc = upload(ctx, tw64, &t.tw64)) return rc;
    ctx->angle_tabs[n] = t;
    *out = t;
    return RADAR_OK;
}

// n > 0: np.hanning(n);  n < 0: all-ones table of length -n
int get_hann(radar_ctx* ctx, int n, float** out) {
    auto it = ctx->hann_by_len.find(n);
    if (it != ctx->hann_by_len.end()) { *out = it->second; return RADAR_OK; }
    std::vector<float> w = n > 0 ? hann_window(n) : std::vector<float>((size_t)(-n), 1.0f);
    float* dev = nullptr;
    if (int rc = upload(ctx, w, &dev)) return rc;
    ctx->hann_by_len[n] = dev;
    *out = dev;
    return RADAR_OK;
}

static int upload_doubles(radar_ctx* ctx, const double* h, int n, double** dev, int* n_dev) {
    if (!h || n <= 0) return set_error(ctx, RADAR_E_INVALID, "empty axis");
    if (*dev) { RB_CUDA_OK(ctx, cudaFree(*dev)); *dev = nullptr; }
    RB_CUDA_OK(ctx, cudaMalloc(reinterpret_cast<void**>(dev), (size_t)n * sizeof(double)));
    RB_CUDA_OK(ctx, cudaMemcpy(*dev, h, (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
    *n_dev = n;
    return RADAR_OK;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace rb

using namespace rb;

#define RB_ENTER(ctx)                                                          \
    if (!(ctx)) return rb::set_error(nullptr, RADAR_E_INVALID, "null ctx");    \
    rb::DeviceGuard _guard((ctx)->device)

extern "C" {

int radar_abi_version(void) { return RADAR_B200_ABI_VERSION; }

const char* radar_last_error(const radar_ctx* ctx) { return ctx ? ctx->err : g_create_error; }
const char* radar_last_rd_kernel(const radar_ctx* ctx) { return ctx ? ctx->last_rd_kernel : ""; }

double radar_cfar_ca_alpha(int n_train_cells, double pfa) {
    return (double)n_train_cells * (std::pow(pfa, -1.0 / (double)n_train_cells) - 1.0);
}

int radar_ctx_create(const radar_shape* shape, int device, radar_ctx** out_ctx) {
    if (!shape || !out_ctx) return set_error(nullptr, RADAR_E_INVALID, "null argument");
    *out_ctx = nullptr;
    if (shape->num_rx < 1 || shape->num_samples < 1 || shape->chirps_per_loop < 1 || shape->num_loops < 1)
        return set_error(nullptr, RADAR_E_INVALID, "radar_shape fields must be >= 1");
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
        return set_error(nullptr, RADAR_E_CUDA, "no CUDA device available: %s", cudaGetErrorString(e));
    if (device < 0 || device >= n_dev) return set_error(nullptr, RADAR_E_INVALID, "device %d out of range (0..%d)", device, n_dev - 1);
    radar_ctx* ctx = new (std::nothrow) radar_ctx();
    if (!ctx) return set_error(nullptr, RADAR_E_NOMEM, "out of host memory");
    ctx->device = device;
    ctx->shape = *shape;
    ctx->V = shape->num_rx * shape->chirps_per_loop;
    ctx->C = shape->chirps_per_loop * shape->num_loops;
    { const char* db = std::getenv("RADAR_B200_SAR_DBG"); ctx->sar_dbg = db ? std::atoi(db) : 0; }
    { const char* tp = std::getenv("RADAR_B200_SAR_TPR"); ctx->sar_tpr = tp ? std::atoi(tp) : 0; }
    { const char* sg = std::getenv("RADAR_B200_SAR_SEG"); ctx->sar_seg_chunks = sg ? std::atoi(sg) : 0; }
    { const char* r4 = std::getenv("RADAR_B200_RA_FFT4"); ctx->ra_fft4 = r4 ? std::atoi(r4) : 1; }
    { const char* pa = std::getenv("RADAR_B200_SAR_PARTS"); ctx->sar_parts = pa ? std::atoi(pa) : 0; }
    { const char* ks = std::getenv("RADAR_B200_SAR_KSPLIT"); ctx->sar_ksplit = ks ? std::atoi(ks) : 0; }
    { const char* sm = std::getenv("RADAR_B200_SAR"); ctx->sar_mode = sm ? std::atoi(sm) : 0; }
    { const char* sv = std::getenv("RADAR_B200_RD_REG"); ctx->rd_reg = sv ? std::atoi(sv) : 1; }
    { const char* sv = std::getenv("RADAR_B200_STREAM_VARIANT"); ctx->stream_variant = sv ? std::atoi(sv) : 0; }
    { const char* sv = std::getenv("RADAR_B200_RA_FC"); ctx->ra_fc = sv ? std::atoi(sv) : 0; }
    { const char* sv = std::getenv("RADAR_B200_PC"); ctx->pc_mode = sv ? std::atoi(sv) : 0; }
    { const char* sv = std::getenv("RADAR_B200_CFAR_BOX"); ctx->cfar_box = sv ? std::atoi(sv) : 1; }
    { const char* sv = std::getenv("RADAR_B200_CFAR_TILE"); ctx->cfar_tile = sv ? std::atoi(sv) : 0; }
    { const char* sv = std::getenv("RADAR_B200_CFAR_LADDER"); ctx->cfar_ladder = sv ? std::atoi(sv) : 1; }
    { const char* sv = std::getenv("RADAR_B200_DBS"); ctx->dbs_mode = sv ? std::atoi(sv) : 0; }
    { const char* sv = std::getenv("RADAR_B200_DBS_CHUNK"); ctx->dbs_chunk = sv ? std::atoi(sv) : 0; }
    { const char* sv = std::getenv("RADAR_B200_CFAR_PH1"); ctx->cfar_ph1 = sv ? std::atoi(sv) : -1; }
    { const char* g = std::getenv("RADAR_B200_GENERIC"); ctx->force_generic = (g && g[0] == '1') ? 1 : ((g && g[0] == '2') ? 2 : 0); }
    DeviceGuard guard(device);
    int rc = RADAR_OK;
    do {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { rc = set_error(ctx, RADAR_E_CUDA, "cudaGetDeviceProperties failed"); break; }
        ctx->sm_count = prop.multiProcessorCount;
        ctx->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
        if ((rc = build_fft_tables(ctx, shape->num_samples, &ctx->fftS))) break;
        if ((rc = build_fft_tables(ctx, shape->num_loops, &ctx->fftL))) break;
        if ((rc = get_hann(ctx, ctx->V, &ctx->hannV))) break;
    } while (0);
    if (rc != RADAR_OK) {
        std::memcpy(g_create_error, ctx->err, sizeof(g_create_error));
        radar_ctx_destroy(ctx);
        return rc;
    }
    *out_ctx = ctx;
    return RADAR_OK;
}

int radar_ctx_destroy(radar_ctx* ctx) {
    if (!ctx) return RADAR_OK;
    DeviceGuard guard(ctx->device);
    free_fft_tables(&ctx->fftS);
    free_fft_tables(&ctx->fftL);
    for (auto& kv : ctx->hann_by_len) cudaFree(kv.second);
    for (auto& kv : ctx->angle_tabs) { cudaFree(kv.second.tw); cudaFree(kv.second.tw64); }
    cudaFree(ctx->range_bins); cudaFree(ctx->vel_bins); cudaFree(ctx->angle_bins); cudaFree(ctx->angle_cos); cudaFree(ctx->angle_sin); cudaFree(ctx->scratch);
    delete ctx;
    return RADAR_OK;
}

int radar_ctx_set_axes(radar_ctx* ctx, const double* h_range_bins, int n_range, const double* h_vel_bins, int n_vel) {
    RB_ENTER(ctx);
    if (n_range < ctx->shape.num_samples || n_vel < ctx->shape.num_loops)
        return set_error(ctx, RADAR_E_INVALID, "axes shorter than the map (%d x %d)", ctx->shape.num_samples, ctx->shape.num_loops);
    if (int rc = upload_doubles(ctx, h_range_bins, n_range, &ctx->range_bins, &ctx->n_range_bins)) return rc;
    return upload_doubles(ctx, h_vel_bins, n_vel, &ctx->vel_bins, &ctx->n_vel_bins);
}

int radar_ctx_set_angle_bins(radar_ctx* ctx, const double* h_angle_bins, int n_bins) {
    RB_ENTER(ctx);
    if (int rc = upload_doubles(ctx, h_angle_bins, n_bins, &ctx->angle_bins, &ctx->n_angle_bins)) return rc;
    // x = r cos(el) cos(az), y = r cos(el) sin(az), z = r sin(el) (point_cloud_generator.py:216-248): the trigonometry of the
    // bin centres is tabulated once in fp64 on the host instead of per detection on the device
    std::vector<double> c((size_t)n_bins), s((size_t)n_bins);
    for (int i = 0; i < n_bins; ++i) { c[(size_t)i] = std::cos(h_angle_bins[i]); s[(size_t)i] = std::sin(h_angle_bins[i]); }
    int n_tmp = 0;
    if (int rc = upload_doubles(ctx, c.data(), n_bins, &ctx->angle_cos, &n_tmp)) return rc;
    return upload_doubles(ctx, s.data(), n_bins, &ctx->angle_sin, &n_tmp);
}

int radar_ctx_set_point_mirrors(radar_ctx* ctx, const uint64_t* h_mirrors, int n_mirrors, int cap_rows) {
    RB_ENTER(ctx);
    if (n_mirrors < 0 || n_mirrors > 16) return set_error(ctx, RADAR_E_INVALID, "n_mirrors %d outside 0..16", n_mirrors);
    if (n_mirrors > 0 && (!h_mirrors || cap_rows <= 0)) return set_error(ctx, RADAR_E_INVALID, "mirrors need pointers and a positive capacity");
    for (int i = 0; i < n_mirrors; ++i) {
        if (!h_mirrors[i] || (h_mirrors[i] & 31u)) return set_error(ctx, RADAR_E_INVALID, "mirror %d is null or not 32-byte aligned", i);
        ctx->point_mirrors[i] = reinterpret_cast<double*>(h_mirrors[i]);
    }
    ctx->n_point_mirrors = n_mirrors;
    ctx->point_mirror_cap = n_mirrors > 0 ? cap_rows : 0;
    return RADAR_OK;
}

int radar_ctx_set_image_mirrors(radar_ctx* ctx, const uint64_t* h_mirrors, int n_mirrors, int64_t row_stride, int64_t col_offset) {
    RB_ENTER(ctx);
    if (n_mirrors < 0 || n_mirrors > 16) return set_error(ctx, RADAR_E_INVALID, "n_mirrors %d outside 0..16", n_mirrors);
    if (n_mirrors > 0 && (!h_mirrors || row_stride <= 0 || col_offset < 0)) return set_error(ctx, RADAR_E_INVALID, "mirrors need pointers, a positive row stride and a non-negative column offset");
    for (int i = 0; i < n_mirrors; ++i) {
        if (!h_mirrors[i] || (h_mirrors[i] & 7u)) return set_error(ctx, RADAR_E_INVALID, "mirror %d is null or not 8-byte aligned", i);
        ctx->image_mirrors[i] = reinterpret_cast<cf*>(h_mirrors[i]);
    }
    ctx->n_image_mirrors = n_mirrors;
    ctx->image_row_stride = n_mirrors > 0 ? row_stride : 0;
    ctx->image_col0 = n_mirrors > 0 ? col_offset : 0;
    return RADAR_OK;
}

int radar_range_doppler(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, const int32_t* ant_list, int n_ant,
                        int window_flags, void* rd_out, float* mag_out, int mag_antenna, void* stream) {
    RB_ENTER(ctx);
    if (!adc || n_frames < 0) return set_error(ctx, RADAR_E_INVALID, "null adc or negative frame count");
    if (!rd_out && !mag_out) return set_error(ctx, RADAR_E_INVALID, "no output requested");
    return launch_range_doppler(ctx, adc, in_kind, n_frames, ant_list, n_ant, nullptr, window_flags,
                                reinterpret_cast<cf*>(rd_out), mag_out, mag_antenna, (cudaStream_t)stream);
}

int radar_range_profile(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx, float* profile_out, void* stream) {
    RB_ENTER(ctx);
    if (!adc || !profile_out || n_frames < 0) return set_error(ctx, RADAR_E_INVALID, "null pointer or negative frame count");
    return launch_range_profile(ctx, adc, in_kind, n_frames, loop_idx, profile_out, (cudaStream_t)stream);
}

int radar_range_zoom(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx, const double* start_m,
                     const double* stop_m, double range_max_m, float* zoom_out, void* stream) {
    RB_ENTER(ctx);
    if (n_frames > 0 && (!adc || !start_m || !stop_m || !zoom_out)) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_range_zoom(ctx, adc, in_kind, n_frames, loop_idx, start_m, stop_m, range_max_m, zoom_out, (cudaStream_t)stream);
}

int radar_find_peaks(radar_ctx* ctx, const float* profiles, int n_profiles, int n, int in_db, double prominence_db,
                     double threshold_db, int max_peaks, int32_t* peak_idx_out, double* peak_db_out, int32_t* count_out,
                     void* stream) {
    RB_ENTER(ctx);
    if (n_profiles > 0 && !profiles) return set_error(ctx, RADAR_E_INVALID, "null profiles");
    return launch_find_peaks(ctx, profiles, n_profiles, n, in_db, prominence_db, threshold_db, max_peaks, peak_idx_out,
                             peak_db_out, count_out, (cudaStream_t)stream);
}

int radar_cfar_2d(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const radar_cfar_params* params,
                  double* noise_out, uint8_t* mask_out, int32_t* dets_out, int32_t* count_out, int max_dets, void* stream) {
    RB_ENTER(ctx);
    if (!maps || !params || n_maps < 0) return set_error(ctx, RADAR_E_INVALID, "null pointer or negative map count");
    return launch_cfar_2d(ctx, maps, n_maps, rows, cols, *params, noise_out, mask_out, dets_out, count_out, max_dets, (cudaStream_t)stream);
}

int radar_cfar_1d_rows(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const radar_cfar_params* params,
                       const uint8_t* row_select, double* noise_out, uint8_t* mask_out, int32_t* dets_out, int32_t* count_out,
                       int max_dets, void* stream) {
    RB_ENTER(ctx);
    if (!maps || !params || n_maps < 0) return set_error(ctx, RADAR_E_INVALID, "null pointer or negative map count");
    return launch_cfar_1d_rows(ctx, maps, n_maps, rows, cols, *params, row_select, noise_out, mask_out, dets_out, count_out, max_dets, (cudaStream_t)stream);
}

int radar_point_cloud(radar_ctx* ctx, const void* rd, int n_frames, const int32_t* dets, const int32_t* counts, int max_dets,
                      const int32_t* az_idx, int n_az, const int32_t* el_idx, int n_el, int shift_az, int shift_el,
                      int num_angle_bins, const int32_t* offsets, double* points_out, int32_t* az_bin_out, int32_t* el_bin_out,
                      void* stream) {
    RB_ENTER(ctx);
    if (!rd || !dets || !counts || !points_out || n_frames < 0) return set_error(ctx, RADAR_E_INVALID, "null pointer or negative frame count");
    if ((n_az > 0 && !az_idx) || (n_el > 0 && !el_idx)) return set_error(ctx, RADAR_E_INVALID, "antenna list missing");
    return launch_point_cloud(ctx, reinterpret_cast<const cf*>(rd), n_frames, dets, counts, max_dets, az_idx, n_az, el_idx, n_el,
                              shift_az, shift_el, num_angle_bins, offsets, points_out, az_bin_out, el_bin_out, (cudaStream_t)stream);
}

int radar_det_offsets(radar_ctx* ctx, const int32_t* counts, int n_frames, int max_dets, int32_t* offsets_out, void* stream) {
    RB_ENTER(ctx);
    if (!counts || !offsets_out || n_frames < 0) return set_error(ctx, RADAR_E_INVALID, "null pointer or negative frame count");
    return launch_det_offsets(ctx, counts, n_frames, max_dets, offsets_out, (cudaStream_t)stream);
}

int radar_range_angle(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx, const int32_t* ant_list, int n_ant,
                      int num_angle_bins, int window_flags, float* out, void* stream) {
    RB_ENTER(ctx);
    if (!adc || !out || n_frames < 0) return set_error(ctx, RADAR_E_INVALID, "null pointer or negative frame count");
    return launch_range_angle(ctx, adc, in_kind, n_frames, loop_idx, ant_list, n_ant, num_angle_bins, window_flags, out, (cudaStream_t)stream);
}

int radar_dbs(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, const int32_t* ant_list, int n_ant, int num_angle_bins,
              const int32_t* vel_bin, const int32_t* ang_bin, int n_out, float* out, void* stream) {
    RB_ENTER(ctx);
    if (!adc || !vel_bin || !ang_bin || !out || n_frames < 0) return set_error(ctx, RADAR_E_INVALID, "null pointer or negative frame count");
    return launch_dbs(ctx, adc, in_kind, n_frames, ant_list, n_ant, num_angle_bins, vel_bin, ang_bin, n_out, out, (cudaStream_t)stream);
}

int radar_sar_image(radar_ctx* ctx, const void* aperture, int num_samples, int n_elems, const double* positions, const double* dirs,
                    int n_dirs, double lambda_m, int g_begin, int g_end, void* out, void* stream) {
    RB_ENTER(ctx);
    if (!aperture || !positions || !dirs || !out) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_sar_image(ctx, reinterpret_cast<const cf*>(aperture), num_samples, n_elems, positions, dirs, n_dirs, lambda_m,
                            g_begin, g_end, reinterpret_cast<cf*>(out), (cudaStream_t)stream);
}

int radar_sar_calibrate(radar_ctx* ctx, const void* aperture, int num_samples, int n_elems, const double* window,
                        const double* positions, const void* image, int n_dirs, int n_elev, const double* dirs, double lambda_m,
                        double* positions_out, int32_t* targets_out, void* stream) {
    RB_ENTER(ctx);
    if (!aperture || !window || !positions || !image || !dirs || !positions_out || !targets_out)
        return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_sar_calibrate(ctx, reinterpret_cast<const cf*>(aperture), num_samples, n_elems, window, positions,
                                reinterpret_cast<const cf*>(image), n_dirs, n_elev, dirs, lambda_m, positions_out, targets_out,
                                (cudaStream_t)stream);
}

int radar_polar_to_cart(radar_ctx* ctx, const void* image, int num_samples, int n_dirs, int n_elev, const double* range_bins,
                        const double* az_bins, const int32_t* az_order, int n_az, double x0, double dx, int nx, double y0, double dy,
                        int ny, int diagonal, void* out, uint8_t* inside_out, void* stream) {
    RB_ENTER(ctx);
    if (!image || !range_bins || !az_bins || (!out && nx > 0 && ny > 0)) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_polar_to_cart(ctx, reinterpret_cast<const cf*>(image), num_samples, n_dirs, n_elev, range_bins, az_bins, az_order,
                                n_az, x0, dx, nx, y0, dy, ny, diagonal, reinterpret_cast<cf*>(out), inside_out, (cudaStream_t)stream);
}

int radar_doppler_azimuth(radar_ctx* ctx, const void* cube, int n_frames, int n_ant, int rows, int cols, int k_begin, int k_end,
                          const float* ant_weight, int num_angle_bins, int shift_angle, float* out, void* stream) {
    RB_ENTER(ctx);
    if (n_frames > 0 && (!cube || !out)) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_doppler_azimuth(ctx, reinterpret_cast<const cf*>(cube), n_frames, n_ant, rows, cols, k_begin, k_end, ant_weight,
                                  num_angle_bins, shift_angle, out, (cudaStream_t)stream);
}

int radar_chirp_zoom(radar_ctx* ctx, const void* range_cube, int n_frames, int n_ant, int k_begin, int k_end, const double* turns,
                     int n_out, void* zoom_out, void* stream) {
    RB_ENTER(ctx);
    if (n_frames > 0 && (!range_cube || !turns || !zoom_out)) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_chirp_zoom(ctx, reinterpret_cast<const cf*>(range_cube), n_frames, n_ant, k_begin, k_end, turns, n_out,
                             reinterpret_cast<cf*>(zoom_out), (cudaStream_t)stream);
}

int radar_chirp_zoom_static(radar_ctx* ctx, const void* range_cube, int n_frames, int n_ant, int k_begin, int k_end, const double* turns,
                            int n_out, void* zoom_out, void* stream) {
    RB_ENTER(ctx);
    if (n_frames > 0 && (!range_cube || !turns || !zoom_out)) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_chirp_zoom_static(ctx, reinterpret_cast<const cf*>(range_cube), n_frames, n_ant, k_begin, k_end, turns, n_out,
                                    reinterpret_cast<cf*>(zoom_out), (cudaStream_t)stream);
}

int radar_chirp_zoom_static_capacity(void) { return chirp_zoom_static_capacity(); }

int radar_row_reduce(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const uint8_t* row_select, int op, float* out,
                     void* stream) {
    RB_ENTER(ctx);
    if (n_maps > 0 && (!maps || !out)) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_row_reduce(ctx, maps, n_maps, rows, cols, row_select, op, out, (cudaStream_t)stream);
}

int radar_ego_velocity(radar_ctx* ctx, const double* points, const int32_t* offsets, int n_frames, int dims, const int32_t* subsets,
                       const int32_t* subset_index, int table_max_n, int max_trials, int min_samples, double residual_threshold, double stop_probability,
                       double* vel_out, double* r2_out, double* inlier_frac_out, int32_t* status_out, void* stream) {
    RB_ENTER(ctx);
    if (n_frames > 0 && (!points || !offsets || !subsets || !subset_index || !vel_out)) return set_error(ctx, RADAR_E_INVALID, "null pointer");
    return launch_ego_velocity(ctx, points, offsets, n_frames, dims, subsets, subset_index, table_max_n, max_trials, min_samples, residual_threshold,
                               stop_probability, vel_out, r2_out, inlier_frac_out, status_out, (cudaStream_t)stream);
}

}  // extern "C"
