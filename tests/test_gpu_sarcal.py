"""GPU parity of SAR self-calibration through the C ABI: radar_sar_calibrate against the oracle and the
reference goldens, and SyntheticArrayBeamformerProcessor(enable_calibration=True) end to end."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg, rel_l2
from test_oracle_sarcal import SARCAL_CASES, SARCAL_CFG, SARCAL_KW, sarcal_frames, sarcal_inputs

pytestmark = pytest.mark.gpu
POS_ATOL = 1e-8      # metres; fp64 on the device, direct DFT sums against numpy's FFT
IMG_TOL = 1e-4       # relative L2 of the fp32 image against the fp64 reference


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


@pytest.mark.parametrize("tag", sorted(SARCAL_CASES))
def test_sar_calibrate_matches_reference(torch_cuda, tag):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.processors._device import engine_for_cube_shape
    cm = load_cfg(SARCAL_CFG)
    g = np.load(os.path.join(GOLDEN_DIR, f"sarcal_{SARCAL_CFG}.npz"))
    hist, geo, d, lam = sarcal_inputs(cm, g, tag)
    f, s, lv = hist.shape
    eng = engine_for_cube_shape(1, s, 1)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    ap = up(hist.transpose(1, 0, 2).reshape(s, -1).astype(np.complex64))
    pos = up(geo.transpose(1, 0, 2).reshape(3, -1))
    dirs = up(d.reshape(3, -1))
    image = eng.sar_image(ap, pos, dirs, lam)
    new, targets = eng.sar_calibrate(ap, up(np.hamming(f * lv)), pos, image, dirs, 1, lam)
    targets = targets.cpu().numpy()
    assert targets[6] == 3
    np.testing.assert_array_equal(targets[:6].reshape(3, 2), g[f"{tag}_first_targets"])
    np.testing.assert_allclose(new.cpu().numpy(), g[f"{tag}_first_calibrated"], rtol=0, atol=POS_ATOL)
    # the same call on the fp64 oracle image (no dependence on the fp32 beamformer): identical targets and geometry
    want, _ = orc.sar_calibrate(hist.transpose(1, 0, 2).reshape(s, -1), geo.transpose(1, 0, 2).reshape(3, -1),
                                orc.sar_response(hist, geo, d, lam)[:, :, 0], d[:, :, 0], lam)
    np.testing.assert_allclose(new.cpu().numpy(), want, rtol=0, atol=POS_ATOL)


@pytest.mark.parametrize("tag", sorted(SARCAL_CASES))
def test_sar_processor_with_calibration(torch_cuda, tag):
    """The reference-facing call: process() frame by frame with enable_calibration=True."""
    from mmwave_radar_processing_b200.processors.simple_synthetic_array_beamformer_processor_multiFrame import \
        SyntheticArrayBeamformerProcessor
    cm = load_cfg(SARCAL_CFG)
    g = np.load(os.path.join(GOLDEN_DIR, f"sarcal_{SARCAL_CFG}.npz"))
    az, nf, stride, iters, _ = SARCAL_CASES[tag]
    p = SyntheticArrayBeamformerProcessor(cm, num_frames=nf, stride=stride, az_angle_bins_rad=az, enable_calibration=True,
                                          num_calibration_iters=iters, **SARCAL_KW)
    for raw, vel in sarcal_frames(cm, tag):
        resp = p.process(raw, current_vel=vel)
    assert p.array_geometry_valid and resp.shape == g[f"{tag}_resp"].shape
    np.testing.assert_allclose(p.array_geometry_calibrated, g[f"{tag}_calibrated"], rtol=0, atol=POS_ATOL)
    assert rel_l2(resp, g[f"{tag}_resp"]) < IMG_TOL
    assert p.interpolated_beamformed_resp.shape == p.interp_x_s.shape


def test_sar_calibrate_degenerate_inputs(torch_cuda):
    """Too few range peaks -> status -1 and IndexError in the mirror's terms; flat azimuth rows -> status < 3 and the
    geometry comes back unchanged; argument validation."""
    torch = torch_cuda
    from mmwave_radar_processing_b200 import _native as nat
    from mmwave_radar_processing_b200.processors._device import engine_for_cube_shape
    s, n, a = 16, 8, 9
    eng = engine_for_cube_shape(1, s, 1)
    x = np.ones((s, n), dtype=np.complex64)
    pos = np.arange(3 * n, dtype=np.float64).reshape(3, n) * 1e-3
    dirs = np.zeros((3, a))
    dirs[0] = 1.0
    up = lambda v: torch.from_numpy(np.ascontiguousarray(v)).cuda()
    flat = np.ones((s, a), dtype=np.complex64)
    new, targets = eng.sar_calibrate(up(x), up(np.ones(n)), up(pos), up(flat), up(dirs), 1, 0.004)
    assert targets.cpu().numpy()[6] == -1                    # a constant aperture has a single spectral line: no 3 peaks
    np.testing.assert_array_equal(new.cpu().numpy(), pos)
    t = np.arange(s)[:, None]
    x = (100 * np.exp(2j * np.pi * 3 * t / s) + 50 * np.exp(2j * np.pi * 7 * t / s) + 20 * np.exp(2j * np.pi * 11 * t / s)
         + 0 * np.arange(n)[None, :]).astype(np.complex64)
    new, targets = eng.sar_calibrate(up(x), up(np.ones(n)), up(pos), up(flat), up(dirs), 1, 0.004)
    targets = targets.cpu().numpy()
    assert targets[6] == 0 and (targets[:6] == -1).all()     # three range peaks, no azimuth peak on flat rows
    np.testing.assert_array_equal(new.cpu().numpy(), pos)
    with pytest.raises(nat.RadarError):
        eng.sar_calibrate(up(x), up(np.ones(n)), up(pos), up(flat), up(dirs), 2, 0.004)      # 9 dirs, 2 elevations


@pytest.mark.parametrize("tag", sorted(SARCAL_CASES))
def test_polar_to_cart_matches_reference_griddata(torch_cuda, tag):
    """radar_polar_to_cart with either diagonal against the oracle; every pixel of the reference's scipy.griddata image
    equals one of the two device images (Qhull's per-cell pick is arbitrary), and the hull is the same."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.processors._device import engine_for_cube_shape
    cm = load_cfg(SARCAL_CFG)
    g = np.load(os.path.join(GOLDEN_DIR, f"sarcal_{SARCAL_CFG}.npz"))
    az = SARCAL_CASES[tag][0]
    image = g[f"{tag}_first_resp"][:, :, 0]
    want = g[f"{tag}_first_interp"]
    x0, y0, res = g[f"{tag}_interp_x0y0"]
    nx, ny = want.shape
    r = np.linspace(0, cm.range_max_m, image.shape[0])
    eng = engine_for_cube_shape(1, image.shape[0], 1)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dev = up(image.astype(np.complex64))
    xq, yq = np.meshgrid(x0 + res * np.arange(nx), y0 + res * np.arange(ny), indexing="ij")
    scale = np.abs(image).max()
    got = []
    for diag in (0, 1):
        out, inside = eng.polar_to_cart(dev, 1, up(r), up(az), x0, res, nx, y0, res, ny, diagonal=diag, want_inside=True)
        ref, ref_in = orc.polar_to_cart(image, r, az, xq, yq, diag)
        np.testing.assert_array_equal(inside.cpu().numpy().ravel().astype(bool), ref_in)
        assert np.abs(out.cpu().numpy().ravel() - ref).max() < 1e-6 * scale
        got.append(out.cpu().numpy())
    close = (np.abs(want - got[0]) < 1e-6 * scale) | (np.abs(want - got[1]) < 1e-6 * scale)
    assert close.all(), int((~close).sum())
    # descending user axis through az_order: same image
    order = np.arange(len(az))[::-1].copy()
    flipped = eng.polar_to_cart(up(image[:, ::-1].astype(np.complex64)), 1, up(r), up(az), x0, res, nx, y0, res, ny,
                                az_order=up(order.astype(np.int32)))
    inner = np.hypot(xq, yq) > r[1]                      # the origin sample is column 0 of the image as given
    np.testing.assert_array_equal(flipped.cpu().numpy()[inner], got[0][inner])


def test_sar_processor_interpolated_image(torch_cuda):
    """process() leaves interpolated_beamformed_resp = the device resampling of the final image (calibrated case 'a')."""
    from mmwave_radar_processing_b200.processors.simple_synthetic_array_beamformer_processor_multiFrame import \
        SyntheticArrayBeamformerProcessor
    cm = load_cfg(SARCAL_CFG)
    g = np.load(os.path.join(GOLDEN_DIR, f"sarcal_{SARCAL_CFG}.npz"))
    az, nf, stride, iters, _ = SARCAL_CASES["a"]
    p = SyntheticArrayBeamformerProcessor(cm, num_frames=nf, stride=stride, az_angle_bins_rad=az, enable_calibration=True,
                                          num_calibration_iters=iters, **SARCAL_KW)
    for raw, vel in sarcal_frames(cm, "a"):
        p.process(raw, current_vel=vel)
    want = g["a_interp"]
    d0 = p.interpolated_beamformed_resp.copy()
    d1 = p.get_interpolated_response_cart(diagonal=1)
    assert d0.shape == want.shape and d0.dtype == np.complex128
    scale = np.abs(g["a_resp"]).max()
    close = (np.abs(want - d0) < 2e-4 * scale) | (np.abs(want - d1) < 2e-4 * scale)      # fp32 image under the resampling
    assert close.mean() > 0.999, close.mean()
    np.testing.assert_array_equal(d0 != 0, want != 0)
