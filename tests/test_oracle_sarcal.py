"""Oracle restatement of SAR self-calibration (reference ...multiFrame.py:673-816) against the reference goldens
(tests/golden/sarcal_*.npz, made by tests/golden/make_golden_sarcal.py)."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg, rel_l2

sys.path.insert(0, GOLDEN_DIR)
from make_golden_sarcal import SARCAL_CASES, SARCAL_CFG, SARCAL_KW, sarcal_frames  # noqa: E402

from oracle import radar_oracle as orc  # noqa: E402


def sarcal_inputs(cm, g, tag):
    """-> (history (F,S,Lv), geometry (F,3,Lv), d (3,A,1), lambda)."""
    az, nf, stride, _, _ = SARCAL_CASES[tag]
    prof = cm.profile_cfgs[0]
    mask, _ = orc.sar_chirp_times_us(cm.chirp_cfgs_per_loop, cm.frameCfg_loops, prof["idleTime_us"] + prof["rampEndTime_us"],
                                     SARCAL_KW["chirp_cfg_idx"], stride)
    hist = np.stack([raw[SARCAL_KW["receiver_idx"]][:, mask] for raw, _ in sarcal_frames(cm, tag)])
    d = orc.sar_steering(az, SARCAL_KW["el_angle_bins_rad"])
    return hist, g[f"{tag}_geometry"], d, orc.C_M_S / (prof["startFreq_GHz"] * 1e9)


@pytest.mark.parametrize("tag", sorted(SARCAL_CASES))
def test_sar_calibration_oracle_matches_reference(tag):
    cm = load_cfg(SARCAL_CFG)
    g = np.load(os.path.join(GOLDEN_DIR, f"sarcal_{SARCAL_CFG}.npz"))
    hist, geo, d, lam = sarcal_inputs(cm, g, tag)
    f, s, lv = hist.shape
    aperture = hist.transpose(1, 0, 2).reshape(s, -1)
    flat = geo.transpose(1, 0, 2).reshape(3, -1)
    image = orc.sar_response(hist, geo, d, lam)
    assert rel_l2(image, g[f"{tag}_first_resp"]) < 1e-11
    new, targets = orc.sar_calibrate(aperture, flat, image[:, :, 0], d[:, :, 0], lam)
    np.testing.assert_array_equal(np.array(targets), g[f"{tag}_first_targets"])
    np.testing.assert_allclose(new, g[f"{tag}_first_calibrated"], rtol=0, atol=1e-10)
    # the processor's own loop: iteration i calibrates the ORIGINAL geometry against the image of iteration i-1
    cal = new
    for _ in range(SARCAL_CASES[tag][3] - 1):
        image = orc.sar_response(hist, cal.reshape(3, f, lv).transpose(1, 0, 2), d, lam)
        cal, _ = orc.sar_calibrate(aperture, flat, image[:, :, 0], d[:, :, 0], lam)
    np.testing.assert_allclose(cal.reshape(3, f, lv).transpose(1, 0, 2), g[f"{tag}_calibrated"], rtol=0, atol=1e-10)
    image = orc.sar_response(hist, g[f"{tag}_calibrated"], d, lam)
    assert rel_l2(image, g[f"{tag}_resp"]) < 1e-9


def test_sar_calibration_target_rules():
    """Fewer than three range peaks -> IndexError (as the reference); a target row without azimuth peak is skipped."""
    rng = np.random.default_rng(0)
    freq = np.ones((16, 8), dtype=complex)
    freq[5] *= 100.0
    with pytest.raises(IndexError):
        orc.sar_calibration_targets(freq, np.ones((16, 9)))
    freq[9] *= 50.0
    freq[12] *= 20.0
    image = np.ones((16, 9))
    image[5, 3] = 40.0
    image[9, 6] = 30.0                      # row 12 stays flat: no azimuth peak
    assert orc.sar_calibration_targets(freq, image) == [(5, 3), (9, 6)]
    new, targets = orc.sar_calibrate(freq * rng.standard_normal((16, 8)), np.zeros((3, 8)), image, np.zeros((3, 9)), 0.004)
    assert new is None and len(targets) < 3


@pytest.mark.parametrize("tag", sorted(SARCAL_CASES))
def test_polar_to_cart_brackets_reference_griddata(tag):
    """Every pixel of the reference's scipy.griddata output equals one of the two Delaunay splits of its polar cell, and the
    hull (non-zero support) is the same."""
    cm = load_cfg(SARCAL_CFG)
    g = np.load(os.path.join(GOLDEN_DIR, f"sarcal_{SARCAL_CFG}.npz"))
    az = SARCAL_CASES[tag][0]
    image = g[f"{tag}_first_resp"][:, :, 0]
    want = g[f"{tag}_first_interp"]
    x0, y0, res = g[f"{tag}_interp_x0y0"]
    xs = x0 + res * np.arange(want.shape[0])
    ys = y0 + res * np.arange(want.shape[1])
    xq, yq = np.meshgrid(xs, ys, indexing="ij")
    r = np.linspace(0, cm.range_max_m, image.shape[0])
    a, ia = orc.polar_to_cart(image, r, az, xq, yq, 0)
    b, ib = orc.polar_to_cart(image, r, az, xq, yq, 1)
    scale = np.abs(image).max()
    ma = np.abs(want.ravel() - a) < 1e-9 * scale
    mb = np.abs(want.ravel() - b) < 1e-9 * scale
    assert (ma | mb).all(), (~(ma | mb)).sum()
    assert ma.mean() < 1.0 and mb.mean() < 1.0          # the reference really mixes both splits
    np.testing.assert_array_equal(ia, ib)
    np.testing.assert_array_equal(ia, (want.ravel() != 0))
