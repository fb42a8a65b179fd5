"""GPU parity of SURVEY 8f next-3 / next-4 through the C ABI: radar_doppler_azimuth, radar_chirp_zoom, radar_row_reduce and
the DopplerAzimuthProcessor / MicroDopplerProcessor / StripMapSARProcessor mirrors, against the oracle on the same seeded
frames and against the reference goldens (tests/golden/da_*.npz)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg, rel_l2
from test_oracle_da import DA_CALLS, DA_CASES, SEED0, da_axes, da_cube, da_subsets

pytestmark = pytest.mark.gpu
TOL = 1e-4          # relative L2, fp32 device arithmetic against the fp64 reference (measured ~2e-7)


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


@pytest.mark.parametrize("name,geom", DA_CASES)
def test_doppler_azimuth_processor_matches_reference(torch_cuda, name, geom):
    """The reference-facing call: host complex cube in, fp64 [vel, valid angle] out, every golden parameter set."""
    from mmwave_radar_processing_b200.processors.doppler_azimuth_resp import DopplerAzimuthProcessor
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"da_{name}.npz"))
    _, cube_v = da_cube(cm, SEED0)
    da = DopplerAzimuthProcessor(cm, num_angle_bins=64)
    np.testing.assert_array_equal(da.valid_angle_mask, g["valid_angle_mask"])
    np.testing.assert_array_equal(da.angle_bins, g["angle_bins"])
    for tag, kw in DA_CALLS:
        for sub_tag, ants in da_subsets(cube_v.shape[0]).items():
            resp = da.process(adc_cube=cube_v.copy(), rx_antennas=list(ants),
                              **{k: (list(v) if isinstance(v, list) else v) for k, v in kw.items()})
            want = g[f"da_{tag}_{sub_tag}"]
            assert resp.dtype == np.float64 and resp.shape == want.shape, (tag, sub_tag)
            assert rel_l2(resp, want) < TOL, (tag, sub_tag, rel_l2(resp, want))
            if kw.get("use_precise_fft"):
                np.testing.assert_allclose(da.zoomed_vel_bins, g[f"da_{tag}_{sub_tag}_zoomed_vel_bins"], rtol=0, atol=1e-12)
    da32 = DopplerAzimuthProcessor(cm, num_angle_bins=32, valid_angle_range=np.array([np.deg2rad(-45), np.deg2rad(30)]))
    assert rel_l2(da32.process(adc_cube=cube_v.copy()), g["da32_coarse"]) < TOL
    # peak helpers on the default coarse response (indices must agree: same angles / velocities)
    resp = da.process(adc_cube=cube_v.copy())
    np.testing.assert_allclose(da.detect_peaks_rows(resp.copy(), da.vel_bins, 30.0), g["peaks_rows"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(da.detect_peak_zero_az(resp.copy(), da.vel_bins, 30.0), g["peak_zero_az"], rtol=0, atol=1e-12)


@pytest.mark.parametrize("name,geom", DA_CASES)
def test_doppler_azimuth_int16_batch_matches_oracle(torch_cuda, name, geom):
    """int16 ingest, a batch of frames, coarse and precise, against the oracle frame by frame; batch == per-frame."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.processors.doppler_azimuth_resp import DopplerAzimuthProcessor
    cm = load_cfg(name, geom)
    vel_bins, range_bins = da_axes(cm)
    ant_win = geom == "standard" and cm.virtual_antennas_enabled
    frames = [da_cube(cm, SEED0 + 50 + f) for f in range(5)]
    iq = torch.from_numpy(np.stack([f[0] for f in frames])).cuda()
    da = DopplerAzimuthProcessor(cm, num_angle_bins=64)
    mask = da.valid_angle_mask
    ants = da_subsets(frames[0][1].shape[0])["sub"]
    coarse = da.process(iq, rx_antennas=ants, range_window=[0.3, 2.0]).cpu().numpy()
    precise = da.process(iq, use_precise_fft=True, precise_vel_range=[-0.3, 0.4]).cpu().numpy()
    for f, (_, cube_v) in enumerate(frames):
        want = orc.doppler_azimuth_coarse(cube_v, range_bins, 64, ants, [0.3, 2.0], True, ant_win, cm.range_max_m)[:, mask]
        assert rel_l2(coarse[f], want) < TOL
        want, _ = orc.doppler_azimuth_precise(cube_v, range_bins, vel_bins, cm.vel_res_m_s, cm.vel_max_m_s, 64, (), (), True,
                                              ant_win, cm.range_max_m, vel_range=[-0.3, 0.4])
        assert rel_l2(precise[f], want[:, mask]) < TOL
    one = da.process(iq[2], rx_antennas=ants, range_window=[0.3, 2.0])
    np.testing.assert_array_equal(one, coarse[2].astype(np.float64))       # a single int16 frame comes back as host fp64


@pytest.mark.parametrize("name,geom", DA_CASES)
def test_micro_doppler_and_strip_map(torch_cuda, name, geom):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.processors.micro_doppler_resp import MicroDopplerProcessor
    from mmwave_radar_processing_b200.processors.strip_map_SAR_processor import StripMapSARProcessor
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"da_{name}.npz"))
    vel_bins, range_bins = da_axes(cm)
    md = MicroDopplerProcessor(cm, target_ranges=[0.4, 2.5], num_frames_history=4)
    np.testing.assert_allclose(md.time_bins, g["micro_time_bins"])
    iqs = []
    for f in range(5):
        iq, cube_v = da_cube(cm, SEED0 + 10 + f)
        iqs.append(iq)
        res = md.process(cube_v, rx_idx=1)
        assert res.shape == g[f"micro_{f}"].shape and rel_l2(res, g[f"micro_{f}"]) < TOL
    cols = md.process(torch.from_numpy(np.stack(iqs)).cuda(), rx_idx=1).cpu().numpy()       # int16 batch: newest-first columns
    assert rel_l2(cols[::-1][:4].T, g["micro_4"]) < TOL
    md.reset()
    assert not md.micro_doppler_resp.any()
    with pytest.raises(ValueError):
        MicroDopplerProcessor(cm, target_ranges=[1e6, 2e6]).process(cube_v)

    iq, cube_v = da_cube(cm, SEED0)
    raw = orc.iq_to_cube(iq)                                  # the reference hands strip-map the RAW (rx, S, chirps) cube
    sm = StripMapSARProcessor(cm)
    resp = sm.process(adc_cube=raw, vel_m_per_s=3.0, sensor_height_m=0.24, rx_index=1, max_SAR_distance=1.5)
    assert resp.dtype == np.complex128 and resp.shape == g["strip_resp"].shape
    assert rel_l2(resp, g["strip_resp"]) < TOL
    np.testing.assert_allclose(sm.x_s, g["strip_x_s"], rtol=1e-12)
    np.testing.assert_allclose(sm.y_s, g["strip_y_s"], rtol=1e-12)
    dev = sm.process(adc_cube=torch.from_numpy(iq[None]).cuda(), vel_m_per_s=3.0, rx_index=1)
    assert rel_l2(dev[0].cpu().numpy(), g["strip_resp"]) < TOL


def test_row_reduce_and_range_only(torch_cuda):
    """radar_row_reduce against numpy; RADAR_RD_RANGE_ONLY against the oracle's windowed range FFT."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200 import _native as nat
    from mmwave_radar_processing_b200.engine import RadarEngine
    cm = load_cfg("1843_RadVel")
    eng = RadarEngine.for_config(cm)
    rng = np.random.default_rng(3)
    maps = rng.random((7, 100, 30), dtype=np.float32)
    sel = rng.random(100) < 0.4
    t = torch.from_numpy(maps).cuda()
    np.testing.assert_array_equal(eng.row_reduce(t, sel, "max").cpu().numpy(), maps[:, sel].max(axis=1))
    np.testing.assert_allclose(eng.row_reduce(t, None, "mean").cpu().numpy(), maps.mean(axis=1), rtol=1e-6)
    iq, cube_v = da_cube(cm, SEED0)
    rc, _ = eng.range_doppler(torch.from_numpy(iq[None]).cuda(), window_flags=3 | nat.RADAR_RD_RANGE_ONLY, want_rd=True)
    want = np.fft.fft(orc.doppler_azimuth_windowed(cube_v), axis=1)
    assert rel_l2(rc[0].cpu().numpy(), want) < TOL
