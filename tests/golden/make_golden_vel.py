"""Golden fixtures for SURVEY 8f next-2 (point cloud -> ego velocity), produced by running the REFERENCE's
point_cloud_processing.vel_estimator.VelocityEstimator (sklearn RANSACRegressor inside) on seeded synthetic point clouds.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_vel.py

Runs only in the build container (imports /root/reference read-only; sklearn version recorded in the fixture).  Inputs
are regenerated in the tests from the seeds (synth_point_cloud below); only outputs are stored (tests/golden/vel_*.npz).
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, REPO)
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

# (n_points, outlier fraction, noise sigma m/s); covers n < min_samples, n == min_samples, the permutation band and
# n > 1000 (min_samples/n < 0.01: tracking selection), clean / noisy / outlier-dominated clouds
VEL_CASES = [(0, 0.0, 0.02), (5, 0.0, 0.02), (10, 0.0, 0.02), (11, 0.2, 0.02), (24, 0.3, 0.03), (40, 0.1, 0.05),
             (75, 0.5, 0.05), (150, 0.25, 0.04), (150, 0.7, 0.04), (400, 0.35, 0.06), (1200, 0.3, 0.05), (60, 0.0, 0.4)]
EGO = np.array([0.8, -0.35, 0.15])


def synth_point_cloud(n, outlier_frac, sigma, seed):
    """(x, y, z, v) rows: static scatterers seen from a platform moving at EGO (v = -dir.EGO + noise), a fraction
    replaced by movers with unrelated radial velocities.  One point of every 50 is duplicated (ties)."""
    rng = np.random.default_rng(seed)
    az = rng.uniform(-1.0, 1.0, n)
    el = rng.uniform(-0.4, 0.4, n)
    r = rng.uniform(0.5, 8.0, n)
    xyz = np.stack([r * np.cos(el) * np.cos(az), r * np.cos(el) * np.sin(az), r * np.sin(el)], axis=1)
    d = xyz / np.maximum(np.linalg.norm(xyz, axis=1, keepdims=True), 1e-12)
    v = -(d @ EGO) + sigma * rng.standard_normal(n)
    out = rng.random(n) < outlier_frac
    v[out] = rng.uniform(-2.0, 2.0, int(out.sum()))
    pts = np.concatenate([xyz, v[:, None]], axis=1)
    for i in range(50, n, 50):
        pts[i] = pts[i - 1]
    return pts


def main():
    import sklearn
    from mmwave_radar_processing.config_managers.cfgManager import ConfigManager  # reference
    from mmwave_radar_processing.point_cloud_processing.vel_estimator import VelocityEstimator  # reference
    for name, geom in (("1843_RadVel", "standard"), ("6843_RadVel_ods_20Hz", "ods")):
        cm = ConfigManager()
        cm.load_cfg(os.path.join(REF, "configs", name + ".cfg"), array_geometry=geom)
        est = VelocityEstimator(cm, min_R2_threshold=0.6, min_inlier_percent=0.75)
        out = {"sklearn_version": np.array(sklearn.__version__)}
        for i, (n, frac, sigma) in enumerate(VEL_CASES):
            pts = synth_point_cloud(n, frac, sigma, 7000 + i)
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                cur = est.process(points=pts.copy())
            out[f"proposed_{i}"] = np.asarray(est.proposed_velocity_estimate, dtype=np.float64)
            out[f"current_{i}"] = np.asarray(cur, dtype=np.float64).copy()
            out[f"r2_{i}"] = np.float64(est.estimated_R2)
            out[f"inlier_{i}"] = np.float64(est.inlier_percent)
            print(name, i, n, out[f"proposed_{i}"], float(out[f"r2_{i}"]), float(out[f"inlier_{i}"]), out[f"current_{i}"])
        np.savez_compressed(os.path.join(HERE, f"vel_{name}.npz"), **out)


if __name__ == "__main__":
    main()
