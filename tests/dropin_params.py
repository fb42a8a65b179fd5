"""Parameter sets of the drop-in tests, shared by ``tests/golden/make_golden_dropin.py`` (which runs the REFERENCE
classes with them) and ``tests/test_gpu_dropin.py`` (which runs this repo's mirrors the same way).

The sets restate, as Python dicts, what the reference's consumers pass:
  * GUI / analyzers: ``gui_configs/processor_params.yaml`` - every processor is built as ``cls(config_manager, **params)``
    and called as ``process(adc_cube=cube, [velocity_ned=...,] **params)`` with the SAME dict
    (``visualization/backends/view_controller.py:39-129``), then ``getattr`` for the ``view_keys`` of
    ``visualization/backends/processor_registry.py:92-214``;
  * ``tests/verify_processors.py:121-191``: PointCloudGenerator with OS-CFAR-2D, RangeDopplerDetector2D with CA-CFAR-2D.
"""
import numpy as np

ALTIMETER = dict(min_altitude_m=25.0e-2, zoom_search_region_m=20.0e-2, altitude_search_limit_m=40e-2, range_bias=0.0,
                 precise_est_enabled=False)
GROUND_DET = dict(vel_cfar_type="os_cfar_1d", vel_cfar_params=dict(num_train=12, num_guard=4, rho=0.5, alpha=15),
                  altimeter_params=dict(ALTIMETER))

P = "processors."
D = P + "range_doppler_detection."
DET_KEYS = ["range_bins", "vel_bins", "dets", "rng_dop_resp"]


def specs(v: int):
    """Processor specs for a cube of ``v`` virtual antennas: (key, module, class, params, view_keys, velocity_ned)."""
    if v >= 12:        # the GUI defaults (6843 ODS)
        ra_ants, az, el = [0, 3, 4, 7], [0, 3, 4, 7], [9, 8, 5, 4]
    elif v >= 8:
        ra_ants, az, el = [0, 3, 4, 7], [0, 1, 2, 3, 4, 5, 6, 7], [1, 5]
    else:
        ra_ants, az, el = [0, 1, 3], [0, 1, 2, 3], [3, 2]
    out = [
        ("range_doppler_resp", P + "range_doppler_resp", "RangeDopplerProcessor", dict(rx_idx=0),
         ["range_bins", "vel_bins"], None),
        ("range_resp", P + "range_resp", "RangeProcessor", dict(chirp_idx=0), [], None),
        ("range_angle_resp", P + "range_angle_resp", "RangeAngleProcessor",
         dict(num_angle_bins=64, rx_antennas=ra_ants, chirp_idx=0), ["range_bins", "angle_bins"], None),
        ("range_angle_resp_256", P + "range_angle_resp", "RangeAngleProcessor",
         dict(num_angle_bins=256, rx_antennas=[], chirp_idx=1), ["range_bins", "angle_bins"], None),
        ("range_angle_resp_1024", P + "range_angle_resp", "RangeAngleProcessor",
         dict(num_angle_bins=1024, rx_antennas=ra_ants, chirp_idx=0, perform_windowing=False),
         ["range_bins", "angle_bins"], None),
        ("dbs_fast", P + "range_angle_resp_dbs_enhanced", "RangeAngleProcessorDBSEnhanced",
         dict(num_angle_bins_range_angle_response=64, num_angle_bins_dbs_enhanced_response=100, min_x_y_vel_dbs=0.25,
              rx_antennas=ra_ants), ["range_bins", "angle_bins"], [0.5, 0.1, 0.0]),
        ("dbs_gated", P + "range_angle_resp_dbs_enhanced", "RangeAngleProcessorDBSEnhanced",
         dict(num_angle_bins_range_angle_response=64, num_angle_bins_dbs_enhanced_response=100, min_x_y_vel_dbs=0.25,
              rx_antennas=ra_ants), ["range_bins", "angle_bins"], [0.1, 0.05, 0.3]),
        ("range_detector", P + "range_detector", "RangeDetector",
         dict(cfar_type="os_cfar_1d", cfar_params=dict(num_train=5, num_guard=3, rho=0.5, alpha=2)),
         ["range_bins", "dets", "thresholds", "range_resp"], None),
        ("range_detector_go", P + "range_detector", "RangeDetector",
         dict(cfar_type="go_cfar_1d", cfar_params=dict(num_train=6, num_guard=2, pfa=1e-2)),
         ["range_bins", "dets", "thresholds", "range_resp"], None),
        ("altimeter", P + "altimeter", "Altimeter", dict(ALTIMETER),
         ["range_bins", "coarse_fft_data", "current_altitude_corrected_m"], None),
        ("det2d_gui", D + "range_doppler_detector_2d", "RangeDopplerDetector2D",
         dict(cfar_type="os_cfar_2d", cfar_params=dict(num_train=[5, 5], num_guard=[3, 2], rho=0.7, alpha=2)),
         DET_KEYS, None),
        ("det2d_verify_ca", D + "range_doppler_detector_2d", "RangeDopplerDetector2D",
         dict(cfar_type="ca_cfar_2d", cfar_params=dict(num_train=(4, 4), num_guard=(2, 2), pfa=1e-5)), DET_KEYS, None),
        ("det2d_window_too_large", D + "range_doppler_detector_2d", "RangeDopplerDetector2D",
         dict(cfar_type="ca_cfar_2d", cfar_params=dict(num_train=(40, 40), num_guard=(2, 2), pfa=1e-5)), DET_KEYS, None),
        ("det_seq", D + "range_doppler_detector_sequential", "RangeDopplerDetectorSequential",
         dict(rng_cfar_type="os_cfar_1d", rng_cfar_params=dict(num_train=5, num_guard=3, rho=0.6, alpha=2),
              vel_cfar_type="os_cfar_1d", vel_cfar_params=dict(num_train=5, num_guard=2, rho=0.7, alpha=3)),
         DET_KEYS, None),
        ("det_seq_ca", D + "range_doppler_detector_sequential", "RangeDopplerDetectorSequential",
         dict(rng_cfar_type="ca_cfar_1d", rng_cfar_params=dict(num_train=6, num_guard=2, pfa=1e-2),
              vel_cfar_type="so_cfar_1d", vel_cfar_params=dict(num_train=4, num_guard=1, pfa=1e-3)),
         DET_KEYS, None),
        ("det_ground", D + "range_doppler_ground_detector", "RangeDopplerGroundDetector",
         dict(vel_cfar_type="os_cfar_1d", vel_cfar_params=dict(num_train=16, num_guard=4, rho=0.5, alpha=12),
              altimeter_params=dict(ALTIMETER)), DET_KEYS, None),
        ("pcg_gui", P + "point_cloud_generator", "PointCloudGenerator",
         dict(detector_type="range_doppler_ground_detector", detector_params=dict(GROUND_DET), az_antenna_idxs=az,
              el_antenna_idxs=el, shift_az_resp=True, shift_el_resp=False), ["angle_bins"], None),
        ("pcg_verify", P + "point_cloud_generator", "PointCloudGenerator",
         dict(detector_type="range_doppler_detector_2d",
              detector_params=dict(cfar_type="os_cfar_2d",
                                   cfar_params=dict(num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5)),
              num_angle_bins=64, az_antenna_idxs=[0, 1, 2, 3], el_antenna_idxs=[], shift_az_resp=True,
              shift_el_resp=False), ["angle_bins"], None),
        ("pcg_seq", P + "point_cloud_generator", "PointCloudGenerator",
         dict(detector_type="range_doppler_detector_sequential",
              detector_params=dict(rng_cfar_type="os_cfar_1d", rng_cfar_params=dict(num_train=5, num_guard=3, rho=0.6, alpha=2),
                                   vel_cfar_type="os_cfar_1d", vel_cfar_params=dict(num_train=5, num_guard=2, rho=0.7, alpha=3)),
              num_angle_bins=128, az_antenna_idxs=az, el_antenna_idxs=el, shift_az_resp=False, shift_el_resp=True),
         ["angle_bins"], None),
        ("pcg_one_antenna", P + "point_cloud_generator", "PointCloudGenerator",
         dict(detector_type="range_doppler_detector_2d",
              detector_params=dict(cfar_type="ca_cfar_2d", cfar_params=dict(num_train=(4, 4), num_guard=(2, 2), pfa=1e-5)),
              num_angle_bins=32, az_antenna_idxs=np.array([v - 1]), el_antenna_idxs=[0]), ["angle_bins"], None),
    ]
    return out


# RangeDopplerProcessor.process argument grid (rx_idx -1 / 0 / last x magnitude / complex)
def rd_variants(v: int):
    return [(rx, mag) for rx in (-1, 0, v - 1) for mag in (True, False)]


# stand-alone detector classes: registry key -> ctor params; run on the reference's fp64 RD map / RD rows / range profile
CFAR_2D = {
    "os_cfar_2d": dict(num_train=[5, 5], num_guard=[3, 2], rho=0.7, alpha=2),
    "ca_cfar_2d": dict(num_train=(4, 4), num_guard=(2, 2), pfa=1e-5),
}
CFAR_2D_EXTRA = {      # (registry key, params, tag)
    "os_rank_edge": ("os_cfar_2d", dict(num_train=(2, 3), num_guard=(1, 0), rho=0.999, alpha=1.5)),
    "ca_too_large": ("ca_cfar_2d", dict(num_train=(40, 40), num_guard=(1, 1), pfa=1e-3)),
}
CFAR_1D = {
    "ca_cfar_1d": dict(num_train=10, num_guard=2, pfa=1e-3),
    "go_cfar_1d": dict(num_train=10, num_guard=2, pfa=1e-3),
    "so_cfar_1d": dict(num_train=10, num_guard=2, pfa=1e-3),
    "os_cfar_1d": dict(num_train=5, num_guard=3, rho=0.5, alpha=2),
}
CFAR_1D_TOO_LARGE = ("ca_cfar_1d", dict(num_train=60, num_guard=5, pfa=1e-3))

# frames of every case: seeded random scenes + an all-zero cube (no detections anywhere)
SEEDS = [10, 11, 12]
CASES = {"6843_RadVel_ods_20Hz": "ods", "1843_RadVel": "standard", "RadCloud": "standard"}

# 20-frame scripted sequence for the ground detector: a strong ground return whose range drifts, two clutter targets
SEQ_FRAMES = 20


def seq_params(v: int, loops: int):
    """PointCloudGenerator kwargs of the scripted sequence: the GUI defaults (``pcg_gui``); on Doppler axes shorter than
    the GUI's 1-D window (2*(12+4)+1 = 33 cells, e.g. 1843_RadVel with 30 loops - the reference then detects nothing at
    all, which ``pcg_gui`` covers) the Doppler CFAR shrinks so that the sequence still produces clouds."""
    params = dict([sp for sp in specs(v) if sp[0] == "pcg_gui"][0][3])
    if loops < 33:
        det = dict(params["detector_params"])
        det["vel_cfar_params"] = dict(num_train=8, num_guard=2, rho=0.5, alpha=15)
        params["detector_params"] = det
    return params


def ground_sequence_targets(frame: int, range_res_m: float, num_samples: int, loops: int):
    """Targets of frame ``frame``: ground return from 0.45 m drifting +1.7 cm per frame with a jump the altimeter must
    reject at frames 8-9 (a return 0.9 m further: outside altitude_search_limit_m), plus off-nadir ground clutter."""
    alt = 0.45 + 0.017 * frame
    ground_bin = alt / range_res_m
    if frame in (8, 9):
        ground_bin += 0.9 / range_res_m
    t = [(900.0, ground_bin, 0.0, 0.0)]
    for k in range(1, 6):        # ring of ground clutter at growing slant range and Doppler
        t.append((500.0 - 60 * k, (alt + 0.06 * k) / range_res_m * 1.0 + 0.8 * k, (-1) ** k * (2.0 + 1.5 * k) * loops / 70.0,
                  0.15 * k - 0.4))
    t.append((400.0, min(num_samples - 5.0, 40.0), 6.0 * loops / 70.0, 0.5))
    return t
