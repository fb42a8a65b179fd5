"""Oracle (doppler_azimuth_coarse / _precise, micro_doppler_column, strip_map_response) against the goldens produced by the
reference's DopplerAzimuthProcessor / MicroDopplerProcessor / StripMapSARProcessor (tests/golden/make_golden_da.py)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg, rel_l2
from oracle import radar_oracle as orc
from mmwave_radar_processing_b200.synthetic import synth_frame_iq

DA_CASES = [("1843_RadVel", "standard"), ("6843_RadVel_ods_20Hz", "ods"), ("RadCloud", "standard")]
SEED0 = 2000
DA_CALLS = [
    ("coarse_default", dict()),
    ("coarse_window", dict(range_window=[0.5, 3.0])),
    ("coarse_noshift", dict(shift_angle=False)),
    ("precise_default", dict(use_precise_fft=True, precise_vel_range=[-0.25, 0.25])),
    ("precise_pos", dict(use_precise_fft=True, precise_vel_range=[0.05, 0.6], range_window=[0.5, 3.0])),
    ("precise_narrow", dict(use_precise_fft=True, precise_vel_range=[-0.02, 0.03])),
]


def da_cube(cm, seed):
    r, s, cpl, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    iq = synth_frame_iq(r, s, cpl, l, seed)
    return iq, orc.virtual_array(orc.iq_to_cube(iq), r, cpl, l)


def da_axes(cm):
    range_bins, vel_bins = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
    return vel_bins, range_bins


def da_subsets(n_v):
    return {"all": [], "sub": [0, 3, 4, 7] if n_v >= 8 else [0, 1, 2]}


@pytest.mark.parametrize("name,geom", DA_CASES)
def test_doppler_azimuth_matches_reference(name, geom):
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"da_{name}.npz"))
    _, cube_v = da_cube(cm, SEED0)
    vel_bins, range_bins = da_axes(cm)
    ant_win = geom == "standard" and cm.virtual_antennas_enabled
    mask = g["valid_angle_mask"]
    for tag, kw in DA_CALLS:
        for sub_tag, ants in da_subsets(cube_v.shape[0]).items():
            common = dict(num_angle_bins=64, rx_antennas=ants, range_window=kw.get("range_window", ()),
                          shift_angle=kw.get("shift_angle", True), antenna_window=ant_win, range_max_m=cm.range_max_m)
            if kw.get("use_precise_fft"):
                resp, zb = orc.doppler_azimuth_precise(cube_v, range_bins, vel_bins, cm.vel_res_m_s, cm.vel_max_m_s,
                                                       vel_range=kw["precise_vel_range"], **common)
                np.testing.assert_allclose(zb, g[f"da_{tag}_{sub_tag}_zoomed_vel_bins"], rtol=0, atol=1e-12)
            else:
                resp = orc.doppler_azimuth_coarse(cube_v, range_bins, **common)
            want = g[f"da_{tag}_{sub_tag}"]
            got = resp[:, mask]
            assert got.shape == want.shape, (tag, sub_tag)
            assert rel_l2(got, want) < 1e-9, (tag, sub_tag, rel_l2(got, want))
    resp32 = orc.doppler_azimuth_coarse(cube_v, range_bins, num_angle_bins=32, antenna_window=ant_win,
                                        range_max_m=cm.range_max_m)
    assert rel_l2(resp32[:, g["da32_valid_angle_mask"]], g["da32_coarse"]) < 1e-10


@pytest.mark.parametrize("name,geom", DA_CASES)
def test_micro_doppler_and_strip_map_match_reference(name, geom):
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"da_{name}.npz"))
    vel_bins, range_bins = da_axes(cm)
    hist = np.zeros((vel_bins.shape[0], 4))
    for f in range(5):
        _, cube_v = da_cube(cm, SEED0 + 10 + f)
        col = orc.micro_doppler_column(cube_v, range_bins, (0.4, 2.5), rx_idx=1)
        hist[:, 1:] = hist[:, 0:-1]
        hist[:, 0] = col
        assert rel_l2(hist, g[f"micro_{f}"]) < 1e-12
    _, cube_v = da_cube(cm, SEED0)
    r0, r1, a0, a1 = (int(v) for v in g["strip_slices"])
    full = orc.strip_map_response(cube_v, rx_index=1)
    assert rel_l2(full[r0:r1, a0:a1], g["strip_resp"]) < 1e-12
