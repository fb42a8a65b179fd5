"""GPU parity of SURVEY 8f next-1 through the C ABI: radar_range_zoom, radar_find_peaks and the Altimeter / RangeProcessor
mirrors, against the oracle on the same seeded frames and against the reference goldens
(tests/golden/altimeter_*.npz, produced by the reference's scipy path)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_cfg, rel_l2
from test_oracle_altimeter import ALT_CASES, ALT_PARAMS, N_FRAMES, alt_frames, coarse_axis

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


@pytest.mark.parametrize("name,geom", ALT_CASES)
def test_zoom_profile_both_layouts(torch_cuda, name, geom):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import RadarEngine
    from mmwave_radar_processing_b200.processors.range_resp import RangeProcessor
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"altimeter_{name}.npz"))
    iq, cubes = alt_frames(cm)
    eng = RadarEngine.for_config(cm)
    iq_dev = torch.from_numpy(iq).to(eng.device)
    rp = RangeProcessor(cm)
    for w in range(3):
        lo = np.array([g[f"zoom_win_{f}_{w}"][0] for f in range(N_FRAMES)])
        hi = np.array([g[f"zoom_win_{f}_{w}"][1] for f in range(N_FRAMES)])
        zoom = eng.range_zoom(iq_dev, lo, hi, cm.range_max_m, loop_idx=0).cpu().numpy()     # int16 ingest, per-frame windows
        for f in range(N_FRAMES):
            ref = g[f"zoom_{f}_{w}"]
            want, _ = orc.zoom_range_profile(cubes[f], lo[f], hi[f], cm.range_max_m, 0)
            assert rel_l2(zoom[f], want) < 1e-4 and rel_l2(zoom[f], ref) < 1e-4        # 1e-4 rel L2 in fp32 (measured ~2e-7)
        # the reference-facing call: host complex cube in, fp64 arrays out
        z, zb = rp.zoom_fft(cubes[0], float(lo[0]), float(hi[0]), chirp_idx=0)
        assert z.dtype == np.float64 and rel_l2(z, g[f"zoom_0_{w}"]) < 1e-4
        np.testing.assert_allclose(zb, np.linspace(lo[0], hi[0], len(z)))


@pytest.mark.parametrize("name,geom", ALT_CASES)
def test_find_peaks_matches_reference(torch_cuda, name, geom):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import RadarEngine
    from mmwave_radar_processing_b200.processors.range_resp import RangeProcessor
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"altimeter_{name}.npz"))
    iq, cubes = alt_frames(cm)
    rb = coarse_axis(cm)
    eng = RadarEngine.for_config(cm)
    prof = eng.range_profile(torch.from_numpy(iq).to(eng.device), loop_idx=0)
    idx, db, cnt = eng.find_peaks(prof, in_db=False, max_peaks=3)
    idx, db, cnt = idx.cpu().numpy(), db.cpu().numpy(), cnt.cpu().numpy()
    rp = RangeProcessor(cm)
    for f in range(N_FRAMES):
        want_m = g[f"coarse_peaks_m_{f}"]
        assert cnt[f] == len(want_m)
        np.testing.assert_array_equal(rb[idx[f, :cnt[f]]], want_m)                      # indices bit-exact
        np.testing.assert_allclose(db[f, :cnt[f]], g[f"coarse_peaks_db_{f}"], rtol=1e-5)
        assert (idx[f, cnt[f]:] == -1).all()
        # reference-facing call on a host dB array (zoom response of window 1)
        z = g[f"zoom_{f}_1"]
        lo, hi = g[f"zoom_win_{f}_1"]
        m, d = rp.find_peaks(rng_resp_db=20 * np.log10(z), rng_bins=np.linspace(lo, hi, len(z)), max_peaks=2)
        np.testing.assert_allclose(m, g[f"zoom_peaks_m_{f}_1"], rtol=0, atol=1e-12)


def test_find_peaks_edge_cases(torch_cuda):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import RadarEngine
    eng = RadarEngine.for_config(load_cfg("1843_RadVel"))
    rows = [
        [0, 5, 5, 5, 0, 8, 0, 8, 8, 1, 9, 0],          # plateaus, a higher neighbour ending the walk
        [3, 2, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0],          # monotone: no peak
        [0, 30, 0, 9, 0, 29, 0, 30, 0, 0, 0, 0],       # 20 dB gate, equal strongest peaks (later one first)
        [1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1],          # flat
        [0, 7, 0, 7, 0, 7, 0, 7, 0, 7, 0, 7],          # more prominent peaks than max_peaks; last sample never a peak
    ]
    x = torch.tensor(rows, dtype=torch.float32, device=eng.device)
    idx, db, cnt = eng.find_peaks(x, in_db=True, max_peaks=3)
    idx, cnt = idx.cpu().numpy(), cnt.cpu().numpy()
    for r, row in enumerate(rows):
        _, _, want = orc.find_range_peaks(np.array(row, float), np.arange(len(row)), max_peaks=3)
        assert cnt[r] == len(want), (r, idx[r], want)
        np.testing.assert_array_equal(idx[r, :cnt[r]], want)
    # random profiles against the oracle (which is pinned to scipy by the goldens)
    rng = np.random.default_rng(5)
    y = (20 * np.log10(rng.rayleigh(size=(64, 257)) + 1e-3)).astype(np.float32)
    idx, db, cnt = eng.find_peaks(torch.from_numpy(y).to(eng.device), in_db=True, max_peaks=5)
    idx, cnt = idx.cpu().numpy(), cnt.cpu().numpy()
    for r in range(y.shape[0]):
        _, _, want = orc.find_range_peaks(y[r].astype(np.float64), np.arange(y.shape[1]), max_peaks=5)
        np.testing.assert_array_equal(idx[r, :cnt[r]], want)


@pytest.mark.parametrize("name,geom", ALT_CASES)
def test_altimeter_recurrence(torch_cuda, name, geom):
    torch = torch_cuda
    from mmwave_radar_processing_b200.engine import RadarEngine
    from mmwave_radar_processing_b200.processors.altimeter import Altimeter
    cm = load_cfg(name, geom)
    g = np.load(os.path.join(GOLDEN_DIR, f"altimeter_{name}.npz"))
    iq, cubes = alt_frames(cm)
    want_fast = np.array([float(g[f"alt_fast_{f}"]) for f in range(N_FRAMES)])
    want_prec = np.array([float(g[f"alt_prec_{f}"]) for f in range(N_FRAMES)])
    # reference call pattern: one host cube per call
    fast, prec = Altimeter(cm, **ALT_PARAMS), Altimeter(cm, **ALT_PARAMS)
    got_fast = np.array([fast.process(cubes[f], precise_est_enabled=False) for f in range(N_FRAMES)])
    got_prec = np.array([prec.process(cubes[f], precise_est_enabled=True) for f in range(N_FRAMES)])
    np.testing.assert_allclose(got_fast, want_fast, rtol=0, atol=1e-12)      # coarse: a bin centre + bias, exact
    np.testing.assert_allclose(got_prec, want_prec, rtol=0, atol=1e-12)      # zoom: a linspace point + bias, exact
    # device batch (extension): int16 frames in one call, altitude after every frame
    eng = RadarEngine.for_config(cm)
    batch = Altimeter(cm, **ALT_PARAMS)
    got = batch.process(torch.from_numpy(iq).to(eng.device), precise_est_enabled=True)
    np.testing.assert_allclose(got, want_prec, rtol=0, atol=1e-12)
    batch.reset()
    assert batch.current_altitude_measured_m == ALT_PARAMS["min_altitude_m"]
