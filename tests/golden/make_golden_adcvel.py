"""Golden fixtures for the ADC-domain ego-velocity estimator, produced by running the REFERENCE's
processors.velocity_estimator.VelocityEstimator (sklearn RANSACRegressor + scipy find_peaks / ZoomFFT inside) on seeded
ego-motion scenes.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_adcvel.py

Runs only in the build container (imports /root/reference read-only).  Inputs are regenerated in the tests from the seeds
(adcvel_frame below); only outputs are stored (tests/golden/adcvel_*.npz).
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

ADCVEL_CFGS = [("1843_RadVel", "standard"), ("6843_RadVel_ods_20Hz", "ods")]
# (ego (boresight, azimuth-lateral, elevation-lateral) m/s, altitude m, precise, x_only, scatterers, mover fraction)
ADCVEL_CASES = [
    ((0.60, 0.30, 0.00), 1.0, False, False, 60, 0.15),
    ((0.60, 0.30, 0.00), 1.0, True, False, 60, 0.15),
    ((0.02, 0.45, 0.10), 1.0, False, False, 60, 0.10),      # small vx: the angle-on-Doppler model
    ((-0.50, -0.25, 0.20), 1.2, False, False, 80, 0.20),    # negative vx
    ((0.90, 0.00, -0.30), 0.8, True, False, 40, 0.40),      # many movers
    ((0.40, 0.20, 0.10), 1.0, False, True, 60, 0.15),       # x measurement only
    ((0.30, 0.10, 0.05), 1.0, False, False, 6, 0.0),        # too few peaks for min_samples
]
ADCVEL_BOUNDS = (0.5, 0.5)                                   # lower / upper range bound around the altitude
ADCVEL_SEED0 = 9100


def adcvel_frame(cm, case_idx):
    """-> int16 frame [chirp][rx][sample][2] of ADCVEL_CASES[case_idx] for a loaded config manager."""
    from mmwave_radar_processing_b200.synthetic import synth_ego_motion_iq
    ego, alt, _, _, k, movers = ADCVEL_CASES[case_idx]
    lo = max(2.0, (alt - 0.4) / cm.range_res_m)
    hi = (alt + 0.4) / cm.range_res_m
    cpl = cm.frameCfg_end_index - cm.frameCfg_start_index + 1          # both config managers have these
    return synth_ego_motion_iq(cm.num_rx_antennas, cm.get_num_adc_samples(0), cpl, cm.frameCfg_loops,
                               ADCVEL_SEED0 + case_idx, ego, cm.vel_res_m_s, geometry=cm.array_geometry,
                               range_bins=(lo, hi), num_scatterers=k, mover_frac=movers)


def main():
    sys.path.insert(0, REF)
    import sklearn
    from mmwave_radar_processing.config_managers.cfgManager import ConfigManager  # reference
    from mmwave_radar_processing.processors.velocity_estimator import VelocityEstimator  # reference
    from mmwave_radar_processing.processors.virtual_array_reformater import VirtualArrayReformatter  # reference
    from mmwave_radar_processing_b200.synthetic import iq_to_complex_cube
    for name, geom in ADCVEL_CFGS:
        cm = ConfigManager()
        cm.load_cfg(os.path.join(REF, "configs", name + ".cfg"), array_geometry=geom)
        reform = VirtualArrayReformatter(cm)
        out = {"sklearn_version": np.array(sklearn.__version__)}
        for i, (ego, alt, precise, x_only, _, _) in enumerate(ADCVEL_CASES):
            est = VelocityEstimator(cm, lower_range_bound=ADCVEL_BOUNDS[0], upper_range_bound=ADCVEL_BOUNDS[1],
                                    x_measurement_only=x_only)
            cube = reform.process(iq_to_complex_cube(adcvel_frame(cm, i)))
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                cur = est.process(adc_cube=cube, altitude=alt, enable_precise_responses=precise)
            out[f"current_{i}"] = np.asarray(cur, dtype=np.float64).copy()
            out[f"proposed_{i}"] = np.asarray(est.proposed_velocity_estimate, dtype=np.float64)
            out[f"vx_{i}"] = np.float64(est.ego_vx_estimate)
            out[f"az_{i}"] = np.array([est.azimuth_ego_vy_estimate, est.azimuth_estimate_R2, est.azimuth_inlier_percent])
            out[f"el_{i}"] = np.array([est.elevation_ego_vy_estimate, est.elevation_estimate_R2,
                                       est.elevation_inlier_percent])
            out[f"az_peaks_{i}"] = np.asarray(est.azimuth_peaks, dtype=np.float64).reshape(-1, 2)
            out[f"el_peaks_{i}"] = np.asarray(est.elevation_peaks, dtype=np.float64).reshape(-1, 2)
            out[f"az_zero_{i}"] = np.asarray(est.azimuth_peak_zero_az, dtype=np.float64)
            print(name, i, "vx", float(out[f"vx_{i}"]), "az", out[f"az_{i}"], "el", out[f"el_{i}"], "peaks",
                  out[f"az_peaks_{i}"].shape[0], out[f"el_peaks_{i}"].shape[0], "cur", out[f"current_{i}"])
        np.savez_compressed(os.path.join(HERE, f"adcvel_{name}.npz"), **out)


if __name__ == "__main__":
    main()
