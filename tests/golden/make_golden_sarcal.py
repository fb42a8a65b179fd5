"""Golden fixtures for SAR self-calibration, produced by running the REFERENCE's
SyntheticArrayBeamformerProcessor with enable_calibration=True on the seeded frames of the SAR goldens.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_sarcal.py

Runs only in the build container (imports /root/reference read-only).  Inputs are regenerated in the tests from the seeds;
only outputs are stored (tests/golden/sarcal_RadSAR_1843_official.npz).
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, REPO)
sys.dont_write_bytecode = True

SARCAL_CFG = "RadSAR_1843_official"
SARCAL_VEL = np.array([0.01, 0.5, 0.0])
# tag -> (azimuth bins, frames, stride, calibration iterations, first frame seed)
SARCAL_CASES = {
    "a": (np.deg2rad(np.linspace(-45, 45, 90)), 5, 1, 1, 1000),
    "b": (np.deg2rad(np.linspace(-60, 60, 121)), 3, 2, 2, 1100),
}
SARCAL_KW = dict(receiver_idx=1, chirp_cfg_idx=0, el_angle_bins_rad=np.array([0.0]), min_vel=np.array([0.0, 0.3, 0.0]),
                 max_vel=np.array([0.05, 0.7, 0.05]), max_vel_stdev=np.array([0.02, 0.02, 0.02]))


def sarcal_frames(cm, tag):
    """-> list of (raw complex cube (rx, S, chirps), velocity) fed to process() for SARCAL_CASES[tag]."""
    from mmwave_radar_processing_b200.synthetic import iq_to_complex_cube, synth_frame_iq
    _, nf, _, _, seed0 = SARCAL_CASES[tag]
    cpl = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
    return [(iq_to_complex_cube(synth_frame_iq(cm.num_rx_antennas, cm.get_num_adc_samples(0), cpl, cm.frameCfg_loops,
                                               seed0 + f)), SARCAL_VEL + np.array([0.0, 0.002 * f, 0.0]))
            for f in range(nf)]


def main():
    sys.path.insert(0, REF)
    from mmwave_radar_processing.config_managers.cfgManager import ConfigManager  # reference
    from mmwave_radar_processing.processors.simple_synthetic_array_beamformer_processor_multiFrame import \
        SyntheticArrayBeamformerProcessor  # reference
    cm = ConfigManager()
    cm.load_cfg(os.path.join(REF, "configs", SARCAL_CFG + ".cfg"))
    out = {}
    for tag, (az, nf, stride, iters, _) in SARCAL_CASES.items():
        p = SyntheticArrayBeamformerProcessor(cm, num_frames=nf, stride=stride, az_angle_bins_rad=az, enable_calibration=True,
                                              num_calibration_iters=iters, **SARCAL_KW)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            for raw, vel in sarcal_frames(cm, tag):
                resp = p.process(raw, current_vel=vel)
        assert p.array_geometry_valid and resp.shape[0] > 0
        out[f"{tag}_resp"] = np.array(resp)
        out[f"{tag}_interp"] = np.array(p.interpolated_beamformed_resp)          # scipy griddata of the calibrated image
        out[f"{tag}_geometry"] = np.array(p.array_geometry)
        out[f"{tag}_calibrated"] = np.array(p.array_geometry_calibrated)
        # the targets of one more calibration pass on the final state (what an extra iteration would pick)
        _, _, freq = p._prepare_calibration_data()
        out[f"{tag}_next_targets"] = np.array(p._find_calibration_targets(freq, num_targets=3))
        # and a single pass from the uncalibrated image, with its targets
        p.compute_synthetic_response(p.array_geometry)
        out[f"{tag}_first_targets"] = np.array(p._find_calibration_targets(freq, num_targets=3))
        out[f"{tag}_first_calibrated"] = np.array(p.perform_array_calibration())
        out[f"{tag}_first_resp"] = np.array(p.compute_synthetic_response(p.array_geometry))
        out[f"{tag}_first_interp"] = np.array(p.get_interpolated_response_cart())
        out[f"{tag}_interp_x0y0"] = np.array([p.interp_x_s[0, 0], p.interp_y_s[0, 0], p.interpolated_grid_resolution_m])
        shift = out[f"{tag}_calibrated"] - out[f"{tag}_geometry"]
        print(tag, "targets", out[f"{tag}_first_targets"].tolist(), "max |shift| m", np.abs(shift).max(),
              "resp max", np.abs(resp).max())
    np.savez_compressed(os.path.join(HERE, f"sarcal_{SARCAL_CFG}.npz"), **out)


if __name__ == "__main__":
    main()
