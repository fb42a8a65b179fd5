"""Oracle restatement of the ADC-domain ego-velocity estimator (reference processors/velocity_estimator.py) against the
reference goldens (tests/golden/adcvel_*.npz, made by tests/golden/make_golden_adcvel.py)."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg
from test_oracle_da import da_axes

sys.path.insert(0, GOLDEN_DIR)
from make_golden_adcvel import ADCVEL_BOUNDS, ADCVEL_CASES, ADCVEL_CFGS, adcvel_frame  # noqa: E402

from oracle import radar_oracle as orc  # noqa: E402

VALID = np.array([np.deg2rad(-70), np.deg2rad(70)])
ATOL = 1e-9          # fp64 restatement against fp64 reference (different lstsq / FFT routines)


def oracle_case(cm, geom, i):
    ego, alt, precise, x_only, _, _ = ADCVEL_CASES[i]
    iq = adcvel_frame(cm, i)
    cube_v = orc.virtual_array(orc.iq_to_cube(iq), cm.num_rx_antennas, cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    vel_bins, range_bins = da_axes(cm)
    window = [max(0, alt - ADCVEL_BOUNDS[0]), min(cm.range_max_m, alt + ADCVEL_BOUNDS[1])]
    return orc.adc_velocity_estimate(cube_v, geom, cm.virtual_antennas_enabled, range_bins, vel_bins, cm.vel_res_m_s,
                                     cm.vel_max_m_s, cm.range_max_m, window, VALID, np.zeros(3), precise=precise,
                                     x_only=x_only)


def check_against_golden(g, i, res, atol):
    x_only = ADCVEL_CASES[i][3]
    np.testing.assert_allclose(res["vx"], g[f"vx_{i}"], rtol=0, atol=atol)
    np.testing.assert_allclose(res["proposed"], g[f"proposed_{i}"], rtol=atol, atol=atol)
    np.testing.assert_allclose(res["current"], g[f"current_{i}"], rtol=atol, atol=atol)
    if not x_only:
        for key in ("az", "el"):
            want = g[f"{key}_peaks_{i}"]
            got = res[f"{key}_peaks"]
            got = np.empty((0, 2)) if got is None else got
            assert got.shape == want.shape, (key, i)
            np.testing.assert_allclose(got, want, rtol=0, atol=1e-12)
            np.testing.assert_allclose(np.array(res[key]), g[f"{key}_{i}"], rtol=atol, atol=atol)


@pytest.mark.parametrize("name,geom", ADCVEL_CFGS)
def test_adc_velocity_oracle_matches_reference(name, geom):
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"adcvel_{name}.npz"))
    for i in range(len(ADCVEL_CASES)):
        check_against_golden(g, i, oracle_case(cm, geom, i), ATOL)


def test_ransac_fit_one_column_models():
    """The generic fit on a 1-column design: recovers the slope, rejects outliers, fails below min_samples."""
    rng = np.random.default_rng(5)
    h = rng.uniform(-1, 1, 40)
    y = 0.7 * h + 0.01 * rng.standard_normal(40)
    y[::7] += 3.0
    coef, r2, frac, mask = orc.ransac_fit(h, y, 0.15)
    assert abs(coef[0] - 0.7) < 0.01 and r2 > 0.99 and not mask[::7].any() and frac == mask.mean()
    assert orc.ransac_fit(h[:9], y[:9], 0.15) is None
    assert orc.ransac_fit(np.full(12, np.nan), y[:12], 0.15) is None
