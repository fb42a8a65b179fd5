"""GPU parity of SURVEY 8f next-2 through the C ABI: radar_ego_velocity and the VelocityEstimator mirror against the
reference goldens (sklearn RANSACRegressor inside the reference, tests/golden/vel_*.npz) and the oracle."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg

sys.path.insert(0, GOLDEN_DIR)
from make_golden_vel import VEL_CASES, synth_point_cloud  # noqa: E402  (seeded input generator only)

pytestmark = pytest.mark.gpu
VEL_CFGS = [("1843_RadVel", "standard", 2), ("6843_RadVel_ods_20Hz", "ods", 3)]


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


@pytest.mark.parametrize("name,geom,dims", VEL_CFGS)
def test_velocity_estimator_matches_reference(torch_cuda, name, geom, dims):
    """The reference-facing call, frame after frame (the accepted estimate is carried like the reference does)."""
    from mmwave_radar_processing_b200.point_cloud_processing.vel_estimator import VelocityEstimator
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"vel_{name}.npz"))
    est = VelocityEstimator(cm, min_R2_threshold=0.6, min_inlier_percent=0.75)
    for i, (n, frac, sigma) in enumerate(VEL_CASES):
        cur = est.process(points=synth_point_cloud(n, frac, sigma, 7000 + i))
        want = g[f"proposed_{i}"]
        assert np.asarray(est.proposed_velocity_estimate).shape == want.shape, f"case {i}"
        np.testing.assert_allclose(est.proposed_velocity_estimate, want, rtol=1e-9, atol=1e-12, err_msg=f"case {i}")
        assert est.estimated_R2 == pytest.approx(float(g[f"r2_{i}"]), rel=1e-9, abs=1e-12)
        assert est.inlier_percent == pytest.approx(float(g[f"inlier_{i}"]), abs=1e-15)     # inlier count bit-exact
        np.testing.assert_allclose(cur, g[f"current_{i}"], rtol=1e-9, atol=1e-12)
        est.update_history(estimated=np.asarray(cur))
    assert len(est.history_R2_statistics) == len(VEL_CASES)


@pytest.mark.parametrize("dims", [2, 3])
def test_ego_velocity_batch_matches_oracle(torch_cuda, dims):
    """300 packed clouds of mixed sizes in one launch against the oracle cloud by cloud; degenerate inputs."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.point_cloud_processing.vel_estimator import VelocityEstimator
    cm = load_cfg("1843_RadVel" if dims == 2 else "6843_RadVel_ods_20Hz", "standard" if dims == 2 else "ods")
    rng = np.random.default_rng(11)
    clouds = []
    for i in range(300):
        n = int(rng.choice([0, 3, 9, 10, 12, 17, 33, 64, 65, 130, 257]))
        clouds.append(synth_point_cloud(n, float(rng.choice([0.0, 0.2, 0.5, 0.8])), 0.04, 9000 + i))
    big = [c for c in clouds if c.shape[0] >= 33]
    big[0][:, 0:3] = big[0][0, 0:3] * np.linspace(1, 2, big[0].shape[0])[:, None]      # all directions parallel: rank 1
    rank1 = next(i for i, c in enumerate(clouds) if c is big[0])
    big[1][3, 0:3] = 0.0                                                              # a point at the origin: 0/0 -> nan
    pts = np.concatenate(clouds, axis=0)
    offsets = np.concatenate([[0], np.cumsum([c.shape[0] for c in clouds])]).astype(np.int32)
    est = VelocityEstimator(cm)
    vel, r2, frac, status = est.estimate_batch(torch.from_numpy(pts).cuda(), torch.from_numpy(offsets).cuda(), dims=dims)
    vel, r2, frac, status = vel.cpu().numpy(), r2.cpu().numpy(), frac.cpu().numpy(), status.cpu().numpy()
    assert (status >= 0).all()
    n_est = 0
    for i, c in enumerate(clouds):
        coef, want_r2, want_frac, mask = orc.ego_velocity_ransac(c, dims=dims) if c.shape[0] else (np.zeros(2), 0.0, 0.0, None)
        if i == rank1:        # H is rank 1 up to rounding: LAPACK's singular-value cut-off decides the reference's answer
            assert np.isfinite(vel[i]).all() and status[i] >= 0
            continue
        if mask is None:
            assert status[i] == 0 and not vel[i].any() and r2[i] == 0.0 and frac[i] == 0.0, i
            continue
        n_est += 1
        assert status[i] == int(mask.sum()), i                                           # inlier count bit-exact
        np.testing.assert_allclose(vel[i, :dims], coef, rtol=1e-8, atol=1e-10, err_msg=f"cloud {i}")
        assert r2[i] == pytest.approx(want_r2, rel=1e-8, abs=1e-10) and frac[i] == pytest.approx(want_frac, abs=1e-15)
    assert n_est > 150


def test_pipeline_point_clouds_feed_the_estimator(torch_cuda):
    """ADC cubes -> packed device point clouds (FramePipeline) -> ego velocity without leaving the device; equals the
    per-frame host call."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.point_cloud_processing.vel_estimator import VelocityEstimator
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    cm = load_cfg("1843_RadVel")
    iq = synth_batch_for_cfg(cm, range(6))
    pipe = FramePipeline(cm, detector=DetectorSpec("os", (8, 4), (2, 1), rho=0.75, alpha=3.5), az_antenna_idxs=(0, 1, 2, 3),
                         el_antenna_idxs=(), max_dets=512)
    out = pipe.run(torch.from_numpy(iq).cuda())
    est = VelocityEstimator(cm)
    vel, r2, frac, status = est.estimate_batch(out["points"], out["offsets"])
    clouds = pipe.gather_host(out)
    for f, c in enumerate(clouds):
        e = VelocityEstimator(cm)
        e.process(points=c)
        if c.shape[0] > 0:
            np.testing.assert_array_equal(vel[f].cpu().numpy(), np.resize(e.proposed_velocity_estimate, 3) if status[f] > 0 else np.zeros(3))
