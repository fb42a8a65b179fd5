"""GPU parity of the ADC-domain ego-velocity estimator: the VelocityEstimator mirror in
processors/velocity_estimator.py (radar_doppler_azimuth / radar_chirp_zoom / radar_find_peaks / radar_ego_velocity behind
it) against the reference goldens and the oracle, and radar_ego_velocity's raw design-row mode against the oracle."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg
from test_oracle_adcvel import ADCVEL_BOUNDS, ADCVEL_CASES, ADCVEL_CFGS, adcvel_frame, check_against_golden, oracle_case

pytestmark = pytest.mark.gpu
ATOL = 1e-7          # fp64 fits on identical peak sets; the peaks themselves (bin values) must agree to 1e-12


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def run_mirror(cm, i, cube):
    from mmwave_radar_processing_b200.processors.velocity_estimator import VelocityEstimator
    ego, alt, precise, x_only, _, _ = ADCVEL_CASES[i]
    est = VelocityEstimator(cm, lower_range_bound=ADCVEL_BOUNDS[0], upper_range_bound=ADCVEL_BOUNDS[1],
                            x_measurement_only=x_only)
    cur = est.process(adc_cube=cube, altitude=alt, enable_precise_responses=precise)
    res = dict(current=np.asarray(cur, dtype=np.float64), proposed=est.proposed_velocity_estimate, vx=est.ego_vx_estimate,
               az=(est.azimuth_ego_vy_estimate, est.azimuth_estimate_R2, est.azimuth_inlier_percent),
               el=(est.elevation_ego_vy_estimate, est.elevation_estimate_R2, est.elevation_inlier_percent),
               az_peaks=np.asarray(est.azimuth_peaks).reshape(-1, 2), el_peaks=np.asarray(est.elevation_peaks).reshape(-1, 2))
    return est, res


@pytest.mark.parametrize("name,geom", ADCVEL_CFGS)
def test_velocity_estimator_matches_reference(torch_cuda, name, geom):
    """The reference-facing call: host complex virtual-array cube in, [vx, vy, vz] out, every golden case."""
    from oracle import radar_oracle as orc
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"adcvel_{name}.npz"))
    for i in range(len(ADCVEL_CASES)):
        iq = adcvel_frame(cm, i)
        cube_v = orc.virtual_array(orc.iq_to_cube(iq), cm.num_rx_antennas, cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
        est, res = run_mirror(cm, i, cube_v)
        check_against_golden(g, i, res, ATOL)
        if i == 0:
            assert est.azimuth_response_mag.dtype == np.float64
            np.testing.assert_allclose(est.get_range_window(1.0), [0.5, 1.5])
            assert est.get_estimated_velocity_measurement_predictions("azimuth").shape == est.valid_angle_bins.shape
            est.update_history(estimated=res["current"])
            assert len(est.history_R2_statistics) == 1


@pytest.mark.parametrize("name,geom", ADCVEL_CFGS)
def test_velocity_estimator_device_frame_matches_oracle(torch_cuda, name, geom):
    """int16 device frame in (no host cube), against the oracle on the same frame."""
    torch = torch_cuda
    cm = load_cfg(name, geom)
    for i in (0, 1, 3):
        iq = torch.from_numpy(adcvel_frame(cm, i)).cuda()
        _, res = run_mirror(cm, i, iq)
        want = oracle_case(cm, geom, i)
        np.testing.assert_allclose(res["current"], want["current"], rtol=ATOL, atol=ATOL)
        np.testing.assert_allclose(res["proposed"], want["proposed"], rtol=ATOL, atol=ATOL)
        np.testing.assert_allclose(res["az_peaks"], want["az_peaks"], rtol=0, atol=1e-12)


def test_points_branch_and_raw_rows(torch_cuda):
    """process(points=...) (standard arrays) against the oracle's 2-D fit; raw design rows (dims -1 / -2 / -3) against
    ransac_fit on a batch of ragged systems."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from make_golden_vel import synth_point_cloud
    from mmwave_radar_processing_b200.engine import RadarEngine
    from mmwave_radar_processing_b200.point_cloud_processing.ransac_subsets import SubsetTable
    from mmwave_radar_processing_b200.processors.velocity_estimator import VelocityEstimator
    cm = load_cfg("1843_RadVel")
    est = VelocityEstimator(cm, lower_range_bound=0.5, upper_range_bound=0.5)
    pts = synth_point_cloud(150, 0.25, 0.04, 7007)
    cur = est.process(points=pts)
    coef, r2, frac, _ = orc.ego_velocity_ransac(pts, 2)
    np.testing.assert_allclose(est.proposed_velocity_estimate, [coef[0], coef[1], 0.0], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose([est.azimuth_estimate_R2, est.azimuth_inlier_percent], [r2, frac], rtol=1e-9)
    assert cur is est.proposed_velocity_estimate or np.allclose(cur, 0.0)
    with pytest.raises(NotImplementedError):
        VelocityEstimator(load_cfg("6843_RadVel_ods_20Hz", "ods"), 0.5, 0.5).process(points=pts)

    eng = RadarEngine.for_config(cm)
    rng = np.random.default_rng(11)
    for dims, d in ((-1, 1), (-2, 2), (-3, 3)):
        counts = [12, 40, 9, 133, 10, 64]
        rows, want = [], []
        for n in counts:
            h = rng.uniform(-1, 1, (n, d))
            y = h @ np.array([0.6, -0.3, 0.2])[:d] + 0.02 * rng.standard_normal(n)
            y[rng.random(n) < 0.2] += rng.uniform(-2, 2)
            block = np.zeros((n, 4))
            block[:, :d] = h
            block[:, 3] = y
            rows.append(block)
            want.append(orc.ransac_fit(h, y, 0.2))
        table = SubsetTable()
        table.ensure(counts)
        off = torch.tensor(np.concatenate([[0], np.cumsum(counts)]), dtype=torch.int32).cuda()
        vel, r2, frac, status = eng.ego_velocity(torch.from_numpy(np.concatenate(rows)).cuda(), off, dims, table,
                                                 residual_threshold=0.2)
        vel, r2, frac, status = vel.cpu().numpy(), r2.cpu().numpy(), frac.cpu().numpy(), status.cpu().numpy()
        for f, w in enumerate(want):
            if w is None:
                assert status[f] == 0 and not vel[f].any()
                continue
            assert status[f] == int(w[3].sum())
            np.testing.assert_allclose(vel[f, :d], w[0], rtol=1e-9, atol=1e-9)
            np.testing.assert_allclose([r2[f], frac[f]], [w[1], w[2]], rtol=1e-9, atol=1e-9)
            assert not vel[f, d:].any()
