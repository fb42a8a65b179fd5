/*
 * radar_b200.h - C ABI of libradar_b200.so: the B200 (sm_100a) ADC-cube -> point-cloud path.
 *
 * This is the drop-in boundary.  The reference is pure Python/numpy and has
 * no FFI today; each entry point below names the reference method whose arithmetic it replaces
 * (paths relative to mmwave_radar_processing/ in the reference tree).  INTEGRATION.md shows the
 * ctypes stub a maintainer of the reference would add.
 *
 * Conventions
 *  - every function returns 0 on success, a negative RADAR_E_* code otherwise; nothing throws or
 *    aborts across the boundary; radar_last_error() returns the message of the last failure;
 *  - all data pointers are DEVICE pointers unless the parameter name starts with "h_";
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream); calls are asynchronous
 *    with respect to the host and ordered on that stream;
 *  - one ctx per (radar shape, device); calls on one ctx must be serialised by the caller;
 *  - complex values are interleaved float pairs (re, im) = numpy complex64;
 *  - ADC ingest layouts (`in_kind`):
 *        RADAR_IN_IQ16    int16  [frame][chirp][rx][sample][2]   (I,Q)  - raw capture order
 *        RADAR_IN_CUBE64  float2 [frame][virtual_antenna][sample][loop] - the reference's
 *                         complex cube after VirtualArrayReformatter (cast to complex64)
 *    with chirp = loop*chirps_per_loop + tx_slot and virtual_antenna = tx_slot*num_rx + rx
 *    (processors/virtual_array_reformater.py:44-65).
 */
#ifndef RADAR_B200_H
#define RADAR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RADAR_B200_ABI_VERSION 1

#define RADAR_OK              0
#define RADAR_E_INVALID      -1   /* bad argument */
#define RADAR_E_UNSUPPORTED  -2   /* shape outside what the kernels handle (e.g. tile exceeds shared memory) */
#define RADAR_E_CUDA         -3   /* CUDA runtime error, see radar_last_error */
#define RADAR_E_NOMEM        -4

#define RADAR_IN_IQ16    0
#define RADAR_IN_CUBE64  1

#define RADAR_WIN_RANGE    1      /* Hann over ADC samples */
#define RADAR_WIN_DOPPLER  2      /* Hann over chirp loops */
#define RADAR_WIN_ANTENNA  4      /* Hann over antennas (range-angle / DBS) */

#define RADAR_RD_RANGE_ONLY 8      /* radar_range_doppler: stop after the range transform (chirp axis stays in loop order,
                                      no Doppler FFT, no fftshift) - the input of radar_chirp_zoom */

#define RADAR_REDUCE_MAX  0
#define RADAR_REDUCE_MEAN 1

#define RADAR_CFAR_CA 0
#define RADAR_CFAR_OS 1
#define RADAR_CFAR_GO 2
#define RADAR_CFAR_SO 3

typedef struct radar_ctx radar_ctx;

typedef struct radar_shape {
    int32_t num_rx;            /* R   enabled RX antennas                      (cfgManager.py:266-278) */
    int32_t num_samples;       /* S   ADC samples per chirp = range FFT length                       */
    int32_t chirps_per_loop;   /* cpl TDM TX slots = frameCfg end-start+1      (cfgManager.py:160-169) */
    int32_t num_loops;         /* L   frameCfg loops = Doppler FFT length                            */
} radar_shape;

/* CFAR configuration (detectors/base.py, ca_cfar.py, os_cfar.py, go_so_cfar.py).
 * 2-D detectors use (train_r, guard_r) along range rows and (train_d, guard_d) along Doppler columns;
 * 1-D detectors use train_d / guard_d along the processed axis.
 * alpha: OS uses it as given.  CA/GO/SO: if alpha <= 0 it is derived from pfa exactly like the reference
 * (alpha = N*(pfa^(-1/N)-1), N = all training cells for CA, ONE side for GO/SO).
 * os_rank: 1-based k of the k-th smallest training cell; if <= 0 it is derived as clip(int(rho*N),1,N). */
typedef struct radar_cfar_params {
    int32_t kind;
    int32_t train_r, train_d;
    int32_t guard_r, guard_d;
    int32_t os_rank;
    double  pfa;
    double  rho;
    double  alpha;
} radar_cfar_params;

/* ---- context ------------------------------------------------------------------------------- */
int radar_abi_version(void);
int radar_ctx_create(const radar_shape* shape, int device, radar_ctx** out_ctx);
int radar_ctx_destroy(radar_ctx* ctx);
/* message of the last failed call on ctx (ctx == NULL: last failure of radar_ctx_create in this thread) */
const char* radar_last_error(const radar_ctx* ctx);
/* name of the kernel the last radar_range_doppler call on ctx launched ("rd_reg_kernel", "rd_stream_kernel", "rd_spec_kernel",
 * "rd_fused_kernel"; "" before the first call) - for benchmark lines and tests that must know which path ran */
const char* radar_last_rd_kernel(const radar_ctx* ctx);

/* Bin axes looked up by radar_point_cloud (host pointers, copied).  They must be produced with the
 * reference's numpy expressions (processors/range_doppler_resp.py:33-47, point_cloud_generator.py:89-105)
 * by the host layer so float-arange quirks match bit for bit. */
int radar_ctx_set_axes(radar_ctx* ctx, const double* h_range_bins, int n_range,
                       const double* h_vel_bins, int n_vel);
int radar_ctx_set_angle_bins(radar_ctx* ctx, const double* h_angle_bins, int n_bins);

/* ---- K1+K2: range-Doppler ------------------------------------------------------------------ *
 * Replaces VirtualArrayReformatter.process + RangeDopplerProcessor.process
 * (processors/virtual_array_reformater.py:44-65, processors/range_doppler_resp.py:49-110) and the
 * magnitude of RangeDopplerDetector._compute_range_doppler_response
 * (processors/range_doppler_detection/range_doppler_detector.py:62-80).
 *   rd_out  : float2 [frame][n_ant_out][S][L], Doppler axis fftshift-ed, or NULL
 *   mag_out : float; mag_antenna >= 0 -> [frame][S][L] = |RD[mag_antenna]|,
 *             mag_antenna == -1 -> [frame][n_ant_out][S][L]; or NULL
 *   ant_list: optional device int32[n_ant] subset of virtual antennas (NULL -> all V, n_ant ignored)
 *   window_flags: RADAR_WIN_RANGE | RADAR_WIN_DOPPLER is the reference behaviour; 0 = un-windowed
 *             (StripMapSARProcessor / MicroDopplerProcessor use that). */
int radar_range_doppler(radar_ctx* ctx, const void* adc, int in_kind, int n_frames,
                        const int32_t* ant_list, int n_ant, int window_flags,
                        void* rd_out, float* mag_out, int mag_antenna, void* stream);

/* ---- a10: coarse range profile ------------------------------------------------------------- *
 * RangeProcessor.coarse_fft (processors/range_resp.py:32-57): chirp `loop_idx` of the virtual cube,
 * Hann(S), |FFT|, mean over all V antennas.   profile_out: float [frame][S]. */
int radar_range_profile(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx,
                        float* profile_out, void* stream);

/* ---- a10b: zoom range profile ------------------------------------------------ *
 * RangeProcessor.zoom_fft (processors/range_resp.py:59-103): chirp `loop_idx`, Hann(S), then
 * scipy.signal.ZoomFFT(S, [f0, f1], fs) with m = S points and endpoint=False, i.e.
 *   X[v][k] = sum_s hann[s] x[v][s] exp(-2 pi i s r_k / range_max_m),  r_k = start + k (stop - start)/S,
 * |.|, mean over all V antennas.  start_m / stop_m: DEVICE double [n_frames] (one window per frame: the altimeter
 * centres it on the frame's coarse ground peak).   zoom_out: float [frame][S]; its range axis is
 * linspace(start, stop, S) as in the reference. */
int radar_range_zoom(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx,
                     const double* start_m, const double* stop_m, double range_max_m,
                     float* zoom_out, void* stream);

/* ---- prominence peak picking ------------------------------------------------ *
 * RangeProcessor.find_peaks (processors/range_resp.py:105-150): scipy.signal.find_peaks(resp_db,
 * prominence=prominence_db) (local maxima with plateau midpoints, prominence over the whole signal), keep the peaks
 * within threshold_db of the strongest, strongest first, at most max_peaks.
 *   profiles   : float [n_profiles][n]; in_db == 0: linear magnitudes, 20*log10 is taken in fp64 on the device
 *                (Altimeter._get_coarse_peaks, altimeter.py:69-76); in_db != 0: already in dB
 *   peak_idx_out: int32 [n_profiles][max_peaks] sample indices, -1 padded;  peak_db_out: double, same shape, or NULL
 *   count_out  : int32 [n_profiles] or NULL */
int radar_find_peaks(radar_ctx* ctx, const float* profiles, int n_profiles, int n, int in_db,
                     double prominence_db, double threshold_db, int max_peaks,
                     int32_t* peak_idx_out, double* peak_db_out, int32_t* count_out, void* stream);

/* ---- K4: CFAR + ordered compaction ---------------------------------------------------------- *
 * 2-D: BaseCFAR2D.detect with CaCFAR2D / OsCFAR2D (detectors/base.py:208-230, ca_cfar.py:85-155,
 * os_cfar.py:94-195) on n_maps maps float [n_maps][rows][cols].
 *   noise_out : double [n_maps][rows][cols] noise estimate (0 outside the valid interior) or NULL;
 *               thresholds = alpha*noise inside, +inf outside: the host layer forms them with the same fp64
 *               product the kernel compares against
 *   mask_out  : uint8 [n_maps][rows][cols] (x > alpha*noise, compared in fp64) or NULL
 *   dets_out  : int32 [n_maps][max_dets][2] (row, col) in row-major order (np.where order) or NULL
 *   count_out : int32 [n_maps] number of detections (may exceed max_dets: list truncated) or NULL */
int radar_cfar_2d(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols,
                  const radar_cfar_params* params, double* noise_out, uint8_t* mask_out,
                  int32_t* dets_out, int32_t* count_out, int max_dets, void* stream);

/* 1-D CFAR along the columns of every selected row: BaseCFAR1D.detect with Ca/Os/Go/So
 * (detectors/base.py:38-65, ca_cfar.py:11-77, os_cfar.py:11-86, go_so_cfar.py:11-123); used on the
 * range profile (rows = 1) and per range bin along Doppler by the sequential and ground detectors
 * (processors/range_doppler_detection/range_doppler_detector_sequential.py:72-108,
 *  range_doppler_ground_detector.py:84-119).
 *   row_select: optional uint8 [n_maps][rows]; rows with 0 are skipped (no detections, noise 0). */
int radar_cfar_1d_rows(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols,
                       const radar_cfar_params* params, const uint8_t* row_select,
                       double* noise_out, uint8_t* mask_out,
                       int32_t* dets_out, int32_t* count_out, int max_dets, void* stream);

/* alpha = N*(pfa^(-1/N)-1) (detectors/base.py:155-169) - host helper so callers print the same number */
double radar_cfar_ca_alpha(int n_train_cells, double pfa);

/* ---- K3b: per-detection angle FFT + point cloud --------------------------------------------- *
 * PointCloudGenerator._compute_angle_estimation + _convert_to_cartesian
 * (processors/point_cloud_generator.py:143-248).
 *   rd       : float2 [frame][V][S][L] from radar_range_doppler (all V antennas)
 *   dets     : int32 [frame][max_dets][2], counts int32 [frame]
 *   az_idx/el_idx : device int32 antenna lists (n == 0 -> angle 0, as the reference)
 *   offsets  : NULL -> padded output: row of detection i of frame f is f*max_dets + i;
 *              else int32 [frame+1] from radar_det_offsets -> packed output: row offsets[f] + i
 *              (frame after frame, np.where order inside a frame - what the NCCL allgather and the
 *              host download move)
 *   points_out : double [rows][4] (x, y, z, velocity), FLU frame (fp64 like the reference)
 *   az_bin_out / el_bin_out : int32 [rows] argmax bins (for bit-exact checks) or NULL
 * Needs radar_ctx_set_axes and radar_ctx_set_angle_bins (n_bins >= num_angle_bins). */
int radar_point_cloud(radar_ctx* ctx, const void* rd, int n_frames,
                      const int32_t* dets, const int32_t* counts, int max_dets,
                      const int32_t* az_idx, int n_az, const int32_t* el_idx, int n_el,
                      int shift_az, int shift_el, int num_angle_bins, const int32_t* offsets,
                      double* points_out, int32_t* az_bin_out, int32_t* el_bin_out, void* stream);

/* Multi-GPU exchange fused into the point-cloud kernel (the variable-length point clouds are the one
 * exchange of the streaming path).  After this call every PACKED row (offsets != NULL) radar_point_cloud writes is also
 * stored at mirrors[i] + 4*row for row < cap_rows: mirrors are device pointers the kernel can store to - peer-mapped
 * buffers of the other ranks (NVLink / NVSwitch peer memory, e.g. torch symmetric memory) and this rank's own slot of
 * the gathered array - so the all-gather of the points happens inside the kernel, row by row, with no staging copy and
 * no collective kernel.  h_mirrors: host array of n_mirrors (<= 16) device pointers; n_mirrors = 0 switches it off.
 * Completion across ranks is the caller's barrier. */
int radar_ctx_set_point_mirrors(radar_ctx* ctx, const uint64_t* h_mirrors, int n_mirrors, int cap_rows);

/* Exclusive scan of min(counts[f], max_dets): offsets_out int32 [n_frames+1], offsets_out[n_frames] = total
 * number of points of the batch (the per-frame list lengths of PointCloudGenerator.process,
 * processors/point_cloud_generator.py:108-140, laid end to end). */
int radar_det_offsets(radar_ctx* ctx, const int32_t* counts, int n_frames, int max_dets,
                      int32_t* offsets_out, void* stream);

/* ---- K3a: range-angle heat map --------------------------------------------------------------- *
 * RangeAngleProcessor.process (processors/range_angle_resp.py:68-122).
 *   ant_list: device int32[n_ant] (NULL -> all V antennas in order); window_flags: RADAR_WIN_RANGE |
 *   RADAR_WIN_ANTENNA = perform_windowing True (the antenna Hann spans ALL V antennas and is indexed by
 *   the ORIGINAL antenna number, applied before the subset).   out: float [frame][S][num_angle_bins]. */
int radar_range_angle(radar_ctx* ctx, const void* adc, int in_kind, int n_frames, int loop_idx,
                      const int32_t* ant_list, int n_ant, int num_angle_bins, int window_flags,
                      float* out, void* stream);

/* ---- K5b: Doppler beam sharpening ------------------------------------------------------------ *
 * RangeAngleProcessorDBSEnhanced.process_dbs_enhanced (processors/range_angle_resp_dbs_enhanced.py:
 * 137-304).  The host layer selects, per output angle i, the Doppler bin vel_bin[i] and the coarse angle
 * bin ang_bin[i] (fp64 argmins, :239-258) and passes them as device int32[n_out].
 *   out: float [frame][S][n_out]. */
int radar_dbs(radar_ctx* ctx, const void* adc, int in_kind, int n_frames,
              const int32_t* ant_list, int n_ant, int num_angle_bins,
              const int32_t* vel_bin, const int32_t* ang_bin, int n_out,
              float* out, void* stream);

/* ---- K5a: multi-frame synthetic-array beamformer ---------------------------------------------- *
 * SyntheticArrayBeamformerProcessor.compute_synthetic_response
 * (processors/simple_synthetic_array_beamformer_processor_multiFrame.py:499-585).
 *   aperture : float2 [S][N]  history reshaped (frame-major, valid chirps), NOT yet Hamming-weighted
 *   positions: double [3][N]  element positions (x,y,z rows)
 *   dirs     : double [3][G]  unit pointing vectors, g = az*E + el
 *   [g_begin, g_end) : voxel slice computed by this call (multi-GPU sharding)
 *   out      : float2 [S][g_end-g_begin] (range bin major), complex response */
int radar_sar_image(radar_ctx* ctx, const void* aperture, int num_samples, int n_elems,
                    const double* positions, const double* dirs, int n_dirs, double lambda_m,
                    int g_begin, int g_end, void* out, void* stream);

/* Multi-GPU exchange fused into the beamformer (the voxel-sharded image).  After this call every cell
 * out[k][j] that radar_sar_image writes (k = range bin, j = direction - g_begin) is also stored at
 * mirrors[i] + k * row_stride + col_offset + j (complex64 cells): mirrors are device pointers the kernel can store to - the
 * full [S][G] images of the other ranks (NVLink / NVSwitch peer memory, e.g. torch symmetric memory) and this rank's own -
 * with row_stride = G and col_offset = this rank's g_begin, so the all-gather of the image tiles happens inside the
 * range-FFT kernel, cell by cell, already in the final layout: no collective kernel, no staging copy, no transpose.
 * h_mirrors: host array of n_mirrors (<= 16) device pointers (8-byte aligned); n_mirrors = 0 switches it off.  Completion
 * across ranks is the caller's barrier.  No counterpart in the reference (single process). */
int radar_ctx_set_image_mirrors(radar_ctx* ctx, const uint64_t* h_mirrors, int n_mirrors, int64_t row_stride, int64_t col_offset);

/* ---- SAR self-calibration of the synthetic-array geometry ---------------------- *
 * SyntheticArrayBeamformerProcessor.perform_array_calibration
 * (processors/simple_synthetic_array_beamformer_processor_multiFrame.py:673-816): range FFT of the windowed
 * aperture, three calibration targets (scipy find_peaks(height=0) on the mean range profile in dB, top three; per
 * target the strongest azimuth peak of |10 log10 |image[rng, :, 0]||), steered phase steps between neighbouring
 * elements (diff of np.unwrap), per-step least squares on (2 pi / lambda) dirs[0:2] and the running sum
 * subtracted from positions[0:2, 1:].  fp64 throughout.
 *   aperture     : float2 [S][N]   as for radar_sar_image
 *   window       : double [N]      element weights (np.hamming(N) in the reference), device
 *   positions    : double [3][N]
 *   image        : float2 [S][n_dirs]  the beamformed response of radar_sar_image over ALL directions
 *   dirs         : double [3][n_dirs], g = az*n_elev + el; only el = 0 is used
 *   positions_out: double [3][N]   calibrated geometry (a copy of positions when status != 3)
 *   targets_out  : int32 [7]       (range bin, azimuth bin) x 3, then status: 3 = calibrated, 0..2 = too few targets
 *                  with an azimuth peak (the reference then keeps the geometry), -1 = fewer than three range peaks
 *                  (the reference raises IndexError there) */
int radar_sar_calibrate(radar_ctx* ctx, const void* aperture, int num_samples, int n_elems, const double* window,
                        const double* positions, const void* image, int n_dirs, int n_elev, const double* dirs,
                        double lambda_m, double* positions_out, int32_t* targets_out, void* stream);

/* ---- polar -> cartesian resampling of the SAR image -------------------------------- *
 * SyntheticArrayBeamformerProcessor.get_interpolated_response_cart
 * (processors/simple_synthetic_array_beamformer_processor_multiFrame.py:587-613): scipy griddata(method="linear",
 * fill_value=0) of image[:, :, el 0] given at (r cos az, r sin az), evaluated on the grid x0 + ix*dx, y0 + iy*dy.
 * Every polar cell is a cyclic trapezoid, so both of its diagonals are Delaunay and Qhull's pick is arbitrary per cell;
 * `diagonal` (0: (i,j)-(i+1,j+1), 1: (i,j+1)-(i+1,j)) fixes the split.  Pixels outside the convex hull are 0.
 *   image     : float2 [S][n_dirs], n_dirs = n_az*n_elev, g = az*n_elev + el (el 0 is read)
 *   range_bins: double [S] ascending; az_bins: double [n_az] ascending (radians, inside (-pi, pi))
 *   az_order  : int32 [n_az] image column of each entry of az_bins (NULL = identity), for descending user axes
 *   out       : float2 [nx][ny]; inside_out: uint8 [nx][ny] or NULL (1 = inside the hull) */
int radar_polar_to_cart(radar_ctx* ctx, const void* image, int num_samples, int n_dirs, int n_elev,
                        const double* range_bins, const double* az_bins, const int32_t* az_order, int n_az,
                        double x0, double dx, int nx, double y0, double dy, int ny, int diagonal,
                        void* out, uint8_t* inside_out, void* stream);

/* ---- Doppler-azimuth response -------------------------------------------------- *
 * DopplerAzimuthProcessor.coarse_doppler_azimuth_fft + the tail of .process
 * (processors/doppler_azimuth_resp.py:296-334, 483-491):
 *   out[f][d][a] = mean over rows k in [k_begin, k_end) of | sum_v w[v] cube[f][v][k][d] exp(-2 pi i v a'/A) |,
 *   a' = (a - A/2) mod A if shift_angle else a  (zero-padded angle FFT, fftshift on the angle axis).
 *   cube      : float2 [frame][n_ant][rows][cols] - the windowed range-Doppler cube of radar_range_doppler
 *               (rows = S, cols = L, Doppler axis already shifted) or the zoom cube of radar_chirp_zoom
 *   ant_weight: device float[n_ant] (np.hanning(n_ant) for standard arrays with TDM-MIMO, :95-98) or NULL
 *   out       : float [frame][cols][num_angle_bins]; the host layer keeps the valid-angle columns. */
int radar_doppler_azimuth(radar_ctx* ctx, const void* cube, int n_frames, int n_ant, int rows, int cols,
                          int k_begin, int k_end, const float* ant_weight, int num_angle_bins, int shift_angle,
                          float* out, void* stream);

/* Zoom Doppler transform of the "precise" Doppler-azimuth response: scipy.signal.ZoomFFT over the chirp axis
 * (processors/doppler_azimuth_resp.py:130-162), X[j] = sum_l x[l] exp(-2 pi i l turns[j]).
 *   range_cube: float2 [frame][n_ant][S][L] from radar_range_doppler(... | RADAR_RD_RANGE_ONLY)
 *   turns     : device double[n_out], turns per chirp of every output bin (formed in fp64 by the host layer with
 *               ZoomFFT's expressions: (f0 + j (f1 - f0)/m)/fs)
 *   zoom_out  : float2 [frame][n_ant][k_end-k_begin][n_out] */
int radar_chirp_zoom(radar_ctx* ctx, const void* range_cube, int n_frames, int n_ant, int k_begin, int k_end,
                     const double* turns, int n_out, void* zoom_out, void* stream);

/* radar_chirp_zoom without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same kernels and same results as radar_chirp_zoom, but the per-call twiddle table [num_loops][n_out]
 * lives in a file-scope device array of the library (512 KiB, held by every GPU that loads the library until exit)
 * instead of the context's cudaMalloc'ed scratch; the call makes no cudaMalloc / cudaFree and calls no allocating
 * library routine.
 *   limit  : num_loops * n_out <= radar_chirp_zoom_static_capacity() (= 65536 twiddles); larger requests return
 *            RADAR_E_UNSUPPORTED with a message naming the limit - use radar_chirp_zoom for those
 *   set-up : the first call on a device looks the array's device address up (cudaGetSymbolAddress, once per device);
 *            a call with n_frames == 0 does only that, e.g. before a graph capture starts
 *   sharing: all calls on a device (any ctx) share the array; every call rewrites the entries it reads before reading
 *            them (nothing of the previous call survives into the result), so calls ordered on one stream are safe;
 *            calls that overlap on the GPU (different streams without a dependency) RACE on it. */
int radar_chirp_zoom_static(radar_ctx* ctx, const void* range_cube, int n_frames, int n_ant, int k_begin, int k_end,
                            const double* turns, int n_out, void* zoom_out, void* stream);
int radar_chirp_zoom_static_capacity(void);

/* Column-wise reduction over the selected rows of every map: MicroDopplerProcessor.process takes np.max of |RD|
 * over the range rows inside target_ranges (processors/micro_doppler_resp.py:104-108).
 *   maps float [n_maps][rows][cols]; row_select device uint8[rows] or NULL (all rows); op RADAR_REDUCE_*;
 *   out float [n_maps][cols]. */
int radar_row_reduce(radar_ctx* ctx, const float* maps, int n_maps, int rows, int cols, const uint8_t* row_select,
                     int op, float* out, void* stream);

/* ---- point cloud -> ego velocity ------------------------------------------------ *
 * VelocityEstimator.lsq_fit_ego_vel_ransac_points_2D / _3D (point_cloud_processing/vel_estimator.py:76-170):
 * y = -v, H = xyz[:dims]/|xyz[:dims]|, sklearn RANSACRegressor(LinearRegression(fit_intercept=False),
 * residual_threshold, random_state=42, max_trials, min_samples) - one warp per frame, fp64.
 *   points  : double [rows][4] packed point clouds (radar_point_cloud with offsets), offsets int32 [n_frames+1]
 *   dims    : 2 / 3 = fit (vx, vy) / (vx, vy, vz) to point rows (x, y, z, v);  -1 / -2 / -3 = the rows already hold the
 *             design matrix and target (h0, h1, h2, y) with that many columns in use (unused columns zero): the 1-D
 *             regressions of the ADC-domain estimator (processors/velocity_estimator.py:409-506)
 *   subsets : device int32 [n_blocks][max_trials][min_samples]; a block holds the point indices sklearn draws for a
 *             cloud of n points (random_state=42 re-seeds every fit, so they depend only on n and the trial);
 *             tabulated by the host layer with numpy's RandomState
 *   subset_index : device int32 [table_max_n + 1], block number for n points, -1 = not tabulated
 *   vel_out : double [n_frames][3] (vx, vy, vz; vz = 0 for dims == 2; zeros when there is no estimate)
 *   r2_out  : double [n_frames] R^2 of the final model on its inliers (0 with <= 3 inliers) or NULL
 *   inlier_frac_out : double [n_frames] inliers / points or NULL
 *   status_out : int32 [n_frames] inlier count of the final model; 0 = no estimate (fewer than min_samples points,
 *             non-finite input, no consensus set: the ValueErrors the reference swallows); -1 = no subset block for
 *             this frame's point count; or NULL */
int radar_ego_velocity(radar_ctx* ctx, const double* points, const int32_t* offsets, int n_frames, int dims,
                       const int32_t* subsets, const int32_t* subset_index, int table_max_n, int max_trials,
                       int min_samples, double residual_threshold, double stop_probability,
                       double* vel_out, double* r2_out, double* inlier_frac_out, int32_t* status_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RADAR_B200_H */
