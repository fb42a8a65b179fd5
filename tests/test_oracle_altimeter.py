"""Oracle (oracle/radar_oracle.py: zoom_range_profile, find_range_peaks, altimeter_step) against the goldens produced by
the reference's RangeProcessor / Altimeter (tests/golden/make_golden_altimeter.py; scipy ZoomFFT + find_peaks there,
direct sums and a restated prominence walk here)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg, rel_l2
from oracle import radar_oracle as orc
from mmwave_radar_processing_b200.synthetic import synth_frame_iq

ALT_PARAMS = dict(min_altitude_m=0.25, zoom_search_region_m=0.2, altitude_search_limit_m=0.4, range_bias=0.05)
ALT_CASES = [("1843_RadVel", "standard"), ("6843_RadVel_ods_20Hz", "ods"), ("RadCloud", "standard")]
N_FRAMES = 12


def alt_frames(cm):
    r, s, cpl, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    iq = np.stack([synth_frame_iq(r, s, cpl, l, 1000 + f) for f in range(N_FRAMES)])
    cubes = [orc.virtual_array(orc.iq_to_cube(iq[f]), r, cpl, l) for f in range(N_FRAMES)]
    return iq, cubes


def coarse_axis(cm):
    return np.arange(start=0, step=cm.range_res_m, stop=cm.range_max_m - cm.range_res_m / 2)


@pytest.mark.parametrize("name,geom", ALT_CASES)
def test_zoom_peaks_altimeter_match_reference(name, geom):
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"altimeter_{name}.npz"))
    _, cubes = alt_frames(cm)
    rb = coarse_axis(cm)
    st_fast = dict(ALT_PARAMS, measured=ALT_PARAMS["min_altitude_m"], corrected=ALT_PARAMS["min_altitude_m"])
    st_prec = dict(st_fast)
    for f in range(N_FRAMES):
        prof = orc.coarse_range_profile(cubes[f], 0)
        with np.errstate(divide="ignore"):
            db = 20 * np.log10(prof)
        pm, pdb, _ = orc.find_range_peaks(db, rb, max_peaks=3)
        np.testing.assert_allclose(pm, g[f"coarse_peaks_m_{f}"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(pdb, g[f"coarse_peaks_db_{f}"], rtol=1e-9)
        for w in range(3):
            lo, hi = g[f"zoom_win_{f}_{w}"]
            z, zb = orc.zoom_range_profile(cubes[f], lo, hi, cm.range_max_m, 0)
            assert rel_l2(z, g[f"zoom_{f}_{w}"]) < 1e-10
            zp, _, _ = orc.find_range_peaks(20 * np.log10(z), zb, max_peaks=2)
            np.testing.assert_allclose(zp, g[f"zoom_peaks_m_{f}_{w}"], rtol=0, atol=1e-12)
        a_fast = orc.altimeter_step(st_fast, prof, rb, precise=False)
        a_prec = orc.altimeter_step(st_prec, prof, rb, cube_v=cubes[f], range_max_m=cm.range_max_m, precise=True)
        assert a_fast == pytest.approx(float(g[f"alt_fast_{f}"]), abs=1e-12)
        assert a_prec == pytest.approx(float(g[f"alt_prec_{f}"]), abs=1e-12)


def test_peak_finding_plateaus_and_edges():
    """scipy semantics the restatement must keep: plateau -> midpoint, edge samples never peaks, a higher neighbour ends
    the prominence walk, equal heights do not."""
    x = np.array([0, 5, 5, 5, 0, 8, 0, 8, 8, 1, 9], dtype=float)
    np.testing.assert_array_equal(orc.local_maxima_1d(x), [2, 5, 7])
    np.testing.assert_allclose(orc.peak_prominences(x, np.array([2, 5, 7])), [5.0, 8.0, 7.0])
    assert orc.local_maxima_1d(np.array([3.0, 2.0, 1.0])).size == 0
    assert orc.local_maxima_1d(np.array([1.0, 2.0, 2.0])).size == 0          # plateau running into the edge
    m, d, idx = orc.find_range_peaks(np.array([0, 30, 0, 9, 0, 29, 0], float), np.arange(7.0), max_peaks=2)
    np.testing.assert_array_equal(idx, [1, 5])                               # 9 dB peak is > 20 dB below the strongest
