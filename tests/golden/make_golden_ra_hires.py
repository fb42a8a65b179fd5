"""High-resolution range-azimuth goldens (BASELINE.json configs[2]: RadarHD / RadCloud, num_angle_bins = 1024) from the
REFERENCE's RangeAngleProcessor (processors/range_angle_resp.py:27-122).  Runs only in the build container.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_ra_hires.py

To keep the fixture small only every 8th range row is stored (float32), together with fp64 row sums and the per-row argmax
of EVERY row, which pin the rows that are not stored.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, "/root/reference")
sys.dont_write_bytecode = True

from mmwave_radar_processing.config_managers.cfgManager import ConfigManager  # noqa: E402  (reference)
from mmwave_radar_processing.processors.virtual_array_reformater import VirtualArrayReformatter  # noqa: E402
from mmwave_radar_processing.processors.range_angle_resp import RangeAngleProcessor  # noqa: E402

from mmwave_radar_processing_b200.synthetic import synth_frame_iq, iq_to_complex_cube  # noqa: E402

CASES = {"RadarHD": (0, 1), "RadCloud": (0,)}
A = 1024
ROW_STEP = 8


def main():
    for name, seeds in CASES.items():
        cm = ConfigManager()
        cm.load_cfg(os.path.join("/root/reference", "configs", name + ".cfg"))
        r, s = cm.num_rx_antennas, cm.get_num_adc_samples(0)
        cpl, l = cm.frameCfg_end_index - cm.frameCfg_start_index + 1, cm.frameCfg_loops
        out = {}
        for seed in seeds:
            cube = iq_to_complex_cube(synth_frame_iq(r, s, cpl, l, seed))
            if cm.virtual_antennas_enabled:
                cube = VirtualArrayReformatter(cm).process(cube)
            rap = RangeAngleProcessor(cm, num_angle_bins=A)
            for tag, kw in (("all", dict(chirp_idx=0, rx_antennas=[])), ("sub_c1", dict(chirp_idx=1, rx_antennas=[0, 1, 3]))):
                resp = rap.process(cube, **kw)
                key = f"s{seed}_{tag}"
                out[key + "_rows"] = resp[::ROW_STEP].astype(np.float32)
                out[key + "_rowsum"] = resp.sum(axis=1)
                out[key + "_argmax"] = resp.argmax(axis=1).astype(np.int32)
                out[key + "_shape"] = np.array(resp.shape, dtype=np.int32)
        path = os.path.join(HERE, f"ra1024_{name}.npz")
        np.savez_compressed(path, **out)
        print(path, os.path.getsize(path) // 1024, "KiB", {k: v.shape for k, v in out.items() if k.endswith("_rows")})


if __name__ == "__main__":
    main()
