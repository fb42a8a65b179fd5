"""Golden fixtures for SURVEY 8f next-1 (zoom range profile, prominence peak picking, altimeter recurrence), produced by
running the REFERENCE's RangeProcessor / Altimeter (scipy ZoomFFT + find_peaks) on seeded synthetic frames.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_altimeter.py

Runs only in the build container (imports /root/reference read-only).  Inputs are regenerated in the tests from the
seeds; only outputs are stored (tests/golden/altimeter_<cfg>.npz).
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
from mmwave_radar_processing.processors.range_resp import RangeProcessor  # noqa: E402
from mmwave_radar_processing.processors.altimeter import Altimeter  # noqa: E402

from mmwave_radar_processing_b200.synthetic import synth_frame_iq, iq_to_complex_cube  # noqa: E402

ALT_PARAMS = dict(min_altitude_m=0.25, zoom_search_region_m=0.2, altitude_search_limit_m=0.4, range_bias=0.05)
N_FRAMES = 12
CASES = {"1843_RadVel": "standard", "6843_RadVel_ods_20Hz": "ods", "RadCloud": "standard"}


def main():
    for name, geom in CASES.items():
        cm = ConfigManager()
        cm.load_cfg(os.path.join(REF, "configs", name + ".cfg"), array_geometry=geom)
        cpl = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
        r, s, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.frameCfg_loops
        reform = VirtualArrayReformatter(cm)
        rp = RangeProcessor(cm)
        alt_fast = Altimeter(cm, **ALT_PARAMS)
        alt_prec = Altimeter(cm, **ALT_PARAMS)
        out = {}
        rmax = float(np.max(rp.range_bins))
        windows = []
        for f in range(N_FRAMES):
            cube = iq_to_complex_cube(synth_frame_iq(r, s, cpl, l, 1000 + f))
            cube_v = reform.process(cube) if cm.virtual_antennas_enabled else cube
            prof = rp.coarse_fft(cube_v, chirp_idx=0)
            with np.errstate(divide="ignore"):
                db = 20 * np.log10(prof)
            pk_m, pk_db = rp.find_peaks(rng_resp_db=db, rng_bins=rp.range_bins, max_peaks=3)
            out[f"coarse_peaks_m_{f}"] = np.asarray(pk_m, dtype=np.float64)
            out[f"coarse_peaks_db_{f}"] = np.asarray(pk_db, dtype=np.float64)
            # zoom windows: around the strongest coarse peak, a wide one and one clipped at the axis ends
            centre = float(pk_m[0]) if len(pk_m) else 0.3 * rmax
            for w, (lo, hi) in enumerate([(max(1e-6, centre - 0.2), min(rmax - 1e-6, centre + 0.2)),
                                          (1e-6, rmax - 1e-6), (0.1 * rmax, 0.35 * rmax)]):
                z, zb = rp.zoom_fft(cube_v, lo, hi, chirp_idx=0)
                zp_m, zp_db = rp.find_peaks(rng_resp_db=20 * np.log10(z), rng_bins=zb, max_peaks=2)
                out[f"zoom_{f}_{w}"] = z
                out[f"zoom_peaks_m_{f}_{w}"] = np.asarray(zp_m, dtype=np.float64)
                if f == 0:
                    windows.append((lo, hi))
                out[f"zoom_win_{f}_{w}"] = np.array([lo, hi])
            out[f"alt_fast_{f}"] = np.float64(alt_fast.process(cube_v, precise_est_enabled=False))
            out[f"alt_prec_{f}"] = np.float64(alt_prec.process(cube_v, precise_est_enabled=True))
        np.savez_compressed(os.path.join(HERE, f"altimeter_{name}.npz"), **out)
        print(name, "alt_fast", [round(float(out[f'alt_fast_{f}']), 4) for f in range(N_FRAMES)])
        print(name, "alt_prec", [round(float(out[f'alt_prec_{f}']), 4) for f in range(N_FRAMES)])


if __name__ == "__main__":
    main()
