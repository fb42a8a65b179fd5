"""Oracle (ego_velocity_ransac, ransac_subsets) against the goldens produced by the reference's
point_cloud_processing VelocityEstimator with sklearn's RANSACRegressor (tests/golden/make_golden_vel.py)."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle import radar_oracle as orc

sys.path.insert(0, GOLDEN_DIR)
from make_golden_vel import VEL_CASES, synth_point_cloud  # noqa: E402  (seeded input generator only)

VEL_CFGS = [("1843_RadVel", 2), ("6843_RadVel_ods_20Hz", 3)]


@pytest.mark.parametrize("name,dims", VEL_CFGS)
def test_ego_velocity_matches_reference(name, dims):
    g = np.load(os.path.join(GOLDEN_DIR, f"vel_{name}.npz"))
    current = np.array([0.0, 0.0, 0.0])
    for i, (n, frac, sigma) in enumerate(VEL_CASES):
        pts = synth_point_cloud(n, frac, sigma, 7000 + i)
        if n > 0:
            coef, r2, ratio, _ = orc.ego_velocity_ransac(pts, dims=dims)
            proposed = np.array([coef[0], coef[1], 0.0]) if dims == 2 else coef
            np.testing.assert_allclose(proposed, g[f"proposed_{i}"], rtol=1e-9, atol=1e-12, err_msg=f"case {i}")
            assert r2 == pytest.approx(float(g[f"r2_{i}"]), rel=1e-9, abs=1e-12)
            assert ratio == pytest.approx(float(g[f"inlier_{i}"]), abs=1e-15)      # inlier COUNT is bit-exact
            if r2 >= 0.6 and ratio >= 0.75:
                current = proposed
        np.testing.assert_allclose(current, g[f"current_{i}"], rtol=1e-9, atol=1e-12)


def test_ransac_subsets_follow_sklearn_sampling():
    """Known answer from sklearn's own docstring (utils/_random.pyx): sample_without_replacement(10, 5,
    random_state=42) == [8, 1, 5, 0, 7]; and the three sampling regimes are distinct, in range and duplicate free."""
    np.testing.assert_array_equal(orc.ransac_subsets(10, 5, 1, 42)[0], [8, 1, 5, 0, 7])
    for n in (10, 11, 49, 51, 400, 1001, 1500):
        s = orc.ransac_subsets(n, 10, 20, 42)
        assert s.shape == (20, 10) and s.min() >= 0 and s.max() < n
        assert all(len(set(row)) == 10 for row in s)
    np.testing.assert_array_equal(orc.ransac_subsets(10, 10, 3, 42), np.tile(np.arange(10), (3, 1)))   # ratio 1: reservoir, no draws
