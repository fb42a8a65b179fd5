// Internal declarations shared by the kernel translation units and the C-ABI layer.
#pragma once
#include <cstdint>
#include <cstdio>
#include <map>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/radar_b200.h"
#include "fft_plan.h"

namespace rb {

// Device-resident FFT tables for one transform length.
struct FftTables {
    FftPlan plan{};
    cf* tw = nullptr;          // [N] W_N^j
    uint16_t* pos = nullptr;   // [N] natural bin -> in-place position
    float* hann = nullptr;     // [N] np.hanning(N)
    float* ones = nullptr;     // [N] all-ones window (window disabled)
};

struct AngleTables {           // zero-padded angle DFT of length A
    cf* tw = nullptr;          // [A] W_A^j as float
    double2* tw64 = nullptr;   // [A] W_A^j in fp64 (per-detection angle estimate)
};

}  // namespace rb

struct radar_ctx {
    int device = 0;
    radar_shape shape{};
    int V = 0, C = 0;
    int sm_count = 0;
    int max_smem_optin = 0;
    rb::FftTables fftS, fftL;
    float* hannV = nullptr;                       // [V] np.hanning(V)
    std::map<int, float*> hann_by_len;            // extra Hann windows (DBS subset length)
    std::map<int, rb::AngleTables> angle_tabs;    // per num_angle_bins
    double* range_bins = nullptr; int n_range_bins = 0;
    double* vel_bins = nullptr;   int n_vel_bins = 0;
    double* angle_bins = nullptr; int n_angle_bins = 0;
    double* angle_cos = nullptr; double* angle_sin = nullptr;   // cos / sin of angle_bins (host fp64)
    void* scratch = nullptr; size_t scratch_bytes = 0;
    int force_generic = 0;                        // RADAR_B200_GENERIC=1: generic range-Doppler kernel only, =2: skip the streaming kernel (tests)
    int sar_mode = 0;                             // RADAR_B200_SAR: 0 auto (tcgen05 when the shape fits), 1 FP32-pipe kernel, 2/3 tcgen05 with 1/2 tiles per CTA
    int sar_dbg = 0;                              // RADAR_B200_SAR_DBG: timing experiments of the tensor-core SAR kernel (results invalid)
    int sar_tpr = 0;                              // RADAR_B200_SAR_TPR: threads per direction row of the tensor-core SAR kernel (0 = default)
    int sar_parts = 0;                            // RADAR_B200_SAR_PARTS=1: per-segment arrays (plain stores) even without a split
    int sar_ksplit = 0;                           // RADAR_B200_SAR_KSPLIT: CTA rows the aperture is cut into (0 = fill the SMs)
    int sar_seg_chunks = 0;                       // RADAR_B200_SAR_SEG: chunks (16 elements) per accumulation segment, 0 = automatic
    double* point_mirrors[16] = {nullptr};       // radar_ctx_set_point_mirrors: extra destinations of every packed point row
    int n_point_mirrors = 0, point_mirror_cap = 0;
    rb::cf* image_mirrors[16] = {nullptr};       // radar_ctx_set_image_mirrors: extra destinations of every cell radar_sar_image writes
    int n_image_mirrors = 0;
    long long image_row_stride = 0, image_col0 = 0;
    int cfar_ph1 = -1;                            // RADAR_B200_CFAR_PH1: extra full rows of the OS-CFAR first counting phase (-1 = default)
    int cfar_box = 1;                             // RADAR_B200_CFAR_BOX=0: CA-CFAR by one window walk per cell (comparison)
    int cfar_tile = 0;                            // RADAR_B200_CFAR_TILE=n: force n x n tiles in the CFAR kernel (tests of the large-map path)
    int cfar_ladder = 1;                          // RADAR_B200_CFAR_LADDER=0: OS-CFAR without the level-ladder screen (tests / comparison)
    int pc_mode = 0;                              // RADAR_B200_PC=1: warp-per-detection point-cloud kernel only (tests / comparison)
    int ra_fft4 = 1;                              // RADAR_B200_RA_FFT4=0: one residue per lane in the range-angle stage (comparison)
    int ra_fc = 0;                                // RADAR_B200_RA_FC: frames per pass of the range-angle kernel (0 = cost model)
    int dbs_mode = 0;                             // RADAR_B200_DBS=1: one-kernel DBS (dbs_kernel) only, =2: two passes over the range-Doppler cube (tests / comparison)
    int dbs_chunk = 0;                            // RADAR_B200_DBS_CHUNK: frames per block of the two-pass DBS (0 = L2-sized)
    int rd_reg = 1;                               // RADAR_B200_RD_REG=0: skip the register-resident range-Doppler kernel (tests / comparison)
    int stream_variant = 0;                       // RADAR_B200_STREAM_VARIANT: tuning candidate of the streaming kernel
    const char* last_rd_kernel = "";              // name of the range-Doppler kernel the last radar_range_doppler call launched
    char err[512] = {0};
};

namespace rb {

int set_error(radar_ctx* ctx, int code, const char* fmt, ...);
int ensure_scratch(radar_ctx* ctx, size_t bytes);
int get_angle_tables(radar_ctx* ctx, int num_angle_bins, AngleTables* out);
int get_hann(radar_ctx* ctx, int n, float** out);

#define RB_CUDA_OK(ctx, expr)                                                                   \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess)                                                                  \
            return rb::set_error((ctx), RADAR_E_CUDA, "%s failed: %s (%s:%d)", #expr,           \
                                 cudaGetErrorString(_e), __FILE__, __LINE__);                   \
    } while (0)

// ---- launchers (kernels_*.cu) ------------------------------------------------------------------
int launch_range_doppler(radar_ctx* ctx, const void* adc, int in_kind, int n_frames,
                         const int32_t* ant_list, int n_ant, const float* ant_weight, int window_flags,
                         cf* rd_out, float* mag_out, int mag_antenna, cudaStream_t st);
int launch_range_doppler_spec(radar_ctx* ctx, const void* adc, int n_frames, const int32_t* ant_list, int n_ant,
                              int window_flags, cf* rd_out, float* mag_out, int mag_antenna, cudaStream_t st);
int launch_range_doppler_stream(radar_ctx* ctx, const void* adc, int n_frames, int window_flags, cf* rd_out,
                                float* mag_out, int mag_antenna, cudaStream_t st);
int launch_range_doppler_reg(radar_ctx* ctx, const void* adc, int n_frames, int window_flags, cf* rd_out,
                             float* mag_out, int mag_antenna, cudaStream_t st);
int launch_dbs_reg(radar_ctx* ctx, const void* adc, int n_frames, const int32_t* ant_list, int N, int num_angle_bins,
                   const int32_t* vel_bin, const int32_t* ang_bin, int n_out, const cf* twA, const float* ant_weight, float* out,
                   cudaStream_t st);
int launch_range_profile(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx,
                         float* profile_out, cudaStream_t st);
int launch_range_zoom(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx, const double* start_m,
                      const double* stop_m, double range_max_m, float* out, cudaStream_t st);
int launch_find_peaks(radar_ctx* ctx, const float* profiles, int n_profiles, int n, int in_db, double prominence,
                      double threshold_db, int max_peaks, int32_t* idx_out, double* db_out, int32_t* count_out,
                      cudaStream_t st);
int launch_range_angle(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx,
                       const int32_t* ant_list, int n_ant, int num_angle_bins, int window_flags,
                       float* out, cudaStream_t st);
int launch_dbs(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, const int32_t* ant_list, int n_ant,
               int num_angle_bins, const int32_t* vel_bin, const int32_t* ang_bin, int n_out, float* out,
               cudaStream_t st);
int launch_cfar_2d(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols,
                   const radar_cfar_params& p, double* noise_out, uint8_t* mask_out, int32_t* dets_out,
                   int32_t* count_out, int max_dets, cudaStream_t st);
int launch_cfar_1d_rows(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols,
                        const radar_cfar_params& p, const uint8_t* row_select, double* noise_out,
                        uint8_t* mask_out, int32_t* dets_out, int32_t* count_out, int max_dets, cudaStream_t st);
int launch_point_cloud(radar_ctx* ctx, const cf* rd, int n_frames, const int32_t* dets, const int32_t* counts,
                       int max_dets, const int32_t* az_idx, int n_az, const int32_t* el_idx, int n_el,
                       int shift_az, int shift_el, int num_angle_bins, const int32_t* offsets, double* points_out,
                       int32_t* az_bin_out, int32_t* el_bin_out, cudaStream_t st);
int launch_det_offsets(radar_ctx* ctx, const int32_t* counts, int n_frames, int max_dets, int32_t* offsets_out,
                       cudaStream_t st);
int launch_doppler_azimuth(radar_ctx* ctx, const cf* cube, int n_frames, int n_ant, int rows, int cols, int k_begin,
                           int k_end, const float* weight, int num_angle_bins, int shift_angle, float* out,
                           cudaStream_t st);
int launch_chirp_zoom(radar_ctx* ctx, const cf* range_cube, int n_frames, int n_ant, int k_begin, int k_end,
                      const double* turns, int n_out, cf* out, cudaStream_t st);
int launch_chirp_zoom_static(radar_ctx* ctx, const cf* range_cube, int n_frames, int n_ant, int k_begin, int k_end,
                             const double* turns, int n_out, cf* out, cudaStream_t st);
int chirp_zoom_static_capacity();
int launch_row_reduce(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const uint8_t* row_select,
                      int op, float* out, cudaStream_t st);
int launch_ego_velocity(radar_ctx* ctx, const double* points, const int32_t* offsets, int n_frames, int dims,
                        const int32_t* subsets, const int32_t* subset_index, int table_max_n, int max_trials,
                        int min_samples, double residual_threshold, double stop_probability, double* vel_out,
                        double* r2_out, double* frac_out, int32_t* status_out, cudaStream_t st);
int launch_sar_image_tc(radar_ctx* ctx, const cf* aperture, int S, int N, const double* positions, const double* dirs,
                        int G, double lambda_m, int g_begin, int g_end, cf* out, cudaStream_t st);
int launch_sar_calibrate(radar_ctx* ctx, const cf* aperture, int S, int N, const double* window, const double* positions,
                         const cf* image, int n_dirs, int n_elev, const double* dirs, double lambda_m, double* positions_out,
                         int32_t* targets_out, cudaStream_t st);
int launch_polar_to_cart(radar_ctx* ctx, const cf* image, int S, int n_dirs, int n_elev, const double* range_bins,
                         const double* az_bins, const int32_t* az_order, int n_az, double x0, double dx, int nx, double y0,
                         double dy, int ny, int diagonal, cf* out, uint8_t* inside_out, cudaStream_t st);
int launch_sar_image(radar_ctx* ctx, const cf* aperture, int num_samples, int n_elems, const double* positions,
                     const double* dirs, int n_dirs, double lambda_m, int g_begin, int g_end, cf* out,
                     cudaStream_t st);

}  // namespace rb
