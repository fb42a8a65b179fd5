"""Golden fixtures for SURVEY 8f next-3 / next-4 (Doppler-azimuth coarse + precise, micro-Doppler, strip-map SAR),
produced by running the REFERENCE's DopplerAzimuthProcessor / MicroDopplerProcessor / StripMapSARProcessor on seeded
synthetic frames.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_da.py

Runs only in the build container (imports /root/reference read-only).  Inputs are regenerated in the tests from the
seeds; only outputs are stored (tests/golden/da_<cfg>.npz).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, REPO)
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

from mmwave_radar_processing.config_managers.cfgManager import ConfigManager  # noqa: E402  (reference)
from mmwave_radar_processing.processors.virtual_array_reformater import VirtualArrayReformatter  # noqa: E402
from mmwave_radar_processing.processors.doppler_azimuth_resp import DopplerAzimuthProcessor  # noqa: E402
from mmwave_radar_processing.processors.micro_doppler_resp import MicroDopplerProcessor  # noqa: E402
from mmwave_radar_processing.processors.strip_map_SAR_processor import StripMapSARProcessor  # noqa: E402

from mmwave_radar_processing_b200.synthetic import synth_frame_iq, iq_to_complex_cube  # noqa: E402

CASES = {"1843_RadVel": "standard", "6843_RadVel_ods_20Hz": "ods", "RadCloud": "standard"}
SEED0 = 2000
N_MICRO = 5

# (tag, kwargs of DopplerAzimuthProcessor.process); rx_antennas filled per config
DA_CALLS = [
    ("coarse_default", dict()),
    ("coarse_window", dict(range_window=[0.5, 3.0])),
    ("coarse_noshift", dict(shift_angle=False)),
    ("precise_default", dict(use_precise_fft=True, precise_vel_range=[-0.25, 0.25])),
    ("precise_pos", dict(use_precise_fft=True, precise_vel_range=[0.05, 0.6], range_window=[0.5, 3.0])),
    ("precise_narrow", dict(use_precise_fft=True, precise_vel_range=[-0.02, 0.03])),
]


def main():
    for name, geom in CASES.items():
        cm = ConfigManager()
        cm.load_cfg(os.path.join(REF, "configs", name + ".cfg"), array_geometry=geom)
        cpl = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
        r, s, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.frameCfg_loops
        reform = VirtualArrayReformatter(cm)

        def cube_v_of(seed):
            cube = iq_to_complex_cube(synth_frame_iq(r, s, cpl, l, seed))
            return reform.process(cube) if cm.virtual_antennas_enabled else cube, cube

        out = {}
        cube_v, cube_raw = cube_v_of(SEED0)
        n_v = cube_v.shape[0]
        subsets = {"all": [], "sub": [0, 3, 4, 7] if n_v >= 8 else [0, 1, 2]}
        da = DopplerAzimuthProcessor(cm, num_angle_bins=64)
        out["valid_angle_mask"] = da.valid_angle_mask
        out["angle_bins"] = da.angle_bins
        for tag, kw in DA_CALLS:
            for sub_tag, ants in subsets.items():
                resp = da.process(adc_cube=cube_v.copy(), rx_antennas=list(ants), **{k: (list(v) if isinstance(v, list) else v) for k, v in kw.items()})
                out[f"da_{tag}_{sub_tag}"] = resp
                if kw.get("use_precise_fft"):
                    out[f"da_{tag}_{sub_tag}_zoomed_vel_bins"] = np.array(da.zoomed_vel_bins)
        # peak helpers on the default coarse response
        resp = da.process(adc_cube=cube_v.copy())
        out["peaks_rows"] = da.detect_peaks_rows(resp.copy(), da.vel_bins, min_threshold_dB=30.0) \
            if True else None
        out["peak_zero_az"] = np.asarray(da.detect_peak_zero_az(resp.copy(), da.vel_bins, min_threshold_dB=30.0))
        da32 = DopplerAzimuthProcessor(cm, num_angle_bins=32, valid_angle_range=np.array([np.deg2rad(-45), np.deg2rad(30)]))
        out["da32_coarse"] = da32.process(adc_cube=cube_v.copy())
        out["da32_valid_angle_mask"] = da32.valid_angle_mask

        # micro-Doppler: history over N_MICRO frames
        md = MicroDopplerProcessor(cm, target_ranges=[0.4, 2.5], num_frames_history=4)
        for f in range(N_MICRO):
            cv, _ = cube_v_of(SEED0 + 10 + f)
            res = md.process(cv, rx_idx=1)
            out[f"micro_{f}"] = res.copy()
        out["micro_time_bins"] = md.time_bins

        # strip-map SAR (takes the RAW cube and reformats itself)
        sm = StripMapSARProcessor(cm)
        resp = sm.process(adc_cube=cube_raw.copy(), vel_m_per_s=3.0, sensor_height_m=0.24, rx_index=1, max_SAR_distance=1.5)
        out["strip_resp"] = resp
        out["strip_x_s"] = sm.x_s
        out["strip_y_s"] = sm.y_s
        out["strip_slices"] = np.array([sm.valid_ranges_slice.start, sm.valid_ranges_slice.stop,
                                        sm.valid_angles_slice.start, sm.valid_angles_slice.stop])
        np.savez_compressed(os.path.join(HERE, f"da_{name}.npz"), **out)
        print(name, {k: v.shape for k, v in out.items() if k.startswith("da_") and "vel_bins" not in k})
        print(name, "peaks_rows", out["peaks_rows"].shape, "zero_az", out["peak_zero_az"], "strip", out["strip_resp"].shape)


if __name__ == "__main__":
    main()
