"""GPU parity of the DROP-IN layer: this repo's mirrors of the reference's ``processors/`` + ``detectors/`` classes, driven
exactly as the reference's consumers drive them, against outputs of the REFERENCE classes run with the same arguments.

How the reference calls its processors (what this file imitates):
  * GUI / analyzers (``visualization/backends/view_controller.py:39-129``): ``proc = spec.processor_cls(config_manager,
    **params)``, then ``proc.process(adc_cube=cube, [velocity_ned=...,] **params)`` with the SAME dict, then ``getattr`` for
    the view keys of ``visualization/backends/processor_registry.py:92-214``;
  * ``scripts/test_vel_estimation.py:145-154``: ``VirtualArrayReformatter.process`` -> ``PointCloudGenerator.process``;
  * ``tests/verify_processors.py:121-191``: PointCloudGenerator + OS-CFAR-2D, RangeDopplerDetector2D + CA-CFAR-2D.

Goldens: ``tests/golden/dropin_<cfg>.npz`` (made by ``tests/golden/make_golden_dropin.py`` from the reference classes;
parameter sets in ``tests/dropin_params.py``).  Every spec runs twice: host complex128 cubes one frame at a time (the
reference's calling convention) and ONE device int16 batch of all frames (this repo's extension).

Tolerances (BASELINE.json north_star): detection / bin indices bit-exact, dtypes and shapes identical (``(0,2)`` int,
``(0,4)`` float); heat maps <= 1e-4 relative L2; CFAR masks identical except cells within 1e-5 of the threshold, counted
and printed; thresholds / noise estimates 1e-6 per cell when the detector is handed the reference's own fp64 map, 1e-4 relative L2 when it
runs on this repo's fp32 RD map.
"""
import importlib
import time

import numpy as np
import pytest

import dropin_params as dp
from conftest import GOLDEN_DIR, load_cfg, rel_l2

pytestmark = pytest.mark.gpu

MAP_TOL = 1e-4
NEAR_THR = 1e-5
STAT_RTOL = 1e-6
PKG = "mmwave_radar_processing_b200."


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


def _gold(name):
    import os
    return np.load(os.path.join(GOLDEN_DIR, f"dropin_{name}.npz"))


def _cls(module, cls_name):
    return getattr(importlib.import_module(PKG + module), cls_name)


def _shape(cm):
    return cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops


def _frames(cm):
    """Same frames as make_golden_dropin.frames_of: seeded scenes + an all-zero frame."""
    from mmwave_radar_processing_b200.synthetic import synth_frame_iq
    r, s, cpl, l = _shape(cm)
    frames = [(f"s{sd}", synth_frame_iq(r, s, cpl, l, sd)) for sd in dp.SEEDS]
    frames.append(("zero", np.zeros_like(frames[0][1])))
    return frames


def _host_cubes(cm, frames):
    """iq -> complex128 (rx, samples, chirps) -> virtual-array cube through the mirror of VirtualArrayReformatter, as
    scripts/test_vel_estimation.py:145-147 does."""
    from mmwave_radar_processing_b200.processors.virtual_array_reformater import VirtualArrayReformatter
    from mmwave_radar_processing_b200.synthetic import iq_to_complex_cube
    var = VirtualArrayReformatter(cm)
    return [(tag, var.process(iq_to_complex_cube(iq))) for tag, iq in frames]


def _call(proc, cube, params, vel):
    if vel is not None:
        return proc.process(adc_cube=cube, velocity_ned=np.array(vel), **params)
    return proc.process(adc_cube=cube, **params)


def _same_array(got, want, what):
    got = np.asarray(got)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    assert got.dtype.kind == want.dtype.kind, f"{what}: dtype {got.dtype} vs {want.dtype}"
    assert np.array_equal(got, want), f"{what}: values differ"


def _map_close(got, want, what):
    got = np.asarray(got)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    if not np.any(want):
        assert not np.any(got), f"{what}: expected an all-zero map"
        return
    err = rel_l2(got, want)
    assert err < MAP_TOL, f"{what}: rel L2 {err:.3e}"


def _axes_equal(proc, g, key, tag, view_keys):
    for attr in view_keys:
        k = f"{key}|{tag}|{attr}"
        if k not in g or attr in ("dets", "rng_dop_resp", "thresholds", "range_resp", "coarse_fft_data",
                                  "current_altitude_corrected_m"):
            continue
        got = np.asarray(getattr(proc, attr))
        assert got.shape == g[k].shape and np.array_equal(got, g[k], equal_nan=True), f"{key}|{tag}|{attr} axis differs"


def _stats_close(thr, noise, mask, g, prefix, x, what, own_map=False):
    """thresholds / noise estimates / detection mask of a CFAR detector object against the reference's cached arrays.
    -> number of mask cells that differ (all of them must sit within NEAR_THR of the threshold).
    own_map: the detector ran on this repo's fp32 RD map (error ~1e-7 of the map's NORM, not of every cell), so the
    statistics are held to the map tolerance; otherwise (same fp64 input as the reference) to STAT_RTOL per cell."""
    gthr, gnoise, gmask = g[prefix + "thresholds"], g[prefix + "noise_estimates"], g[prefix + "detections"]
    assert thr.shape == gthr.shape and noise.shape == gnoise.shape and mask.shape == gmask.shape, what
    assert mask.dtype == np.bool_, what
    fin = np.isfinite(gthr)
    assert np.array_equal(np.isfinite(thr), fin), f"{what}: +inf border of thresholds differs"
    if own_map:
        assert not fin.any() or rel_l2(thr[fin], gthr[fin]) < MAP_TOL, f"{what}: thresholds"
        assert not gnoise.any() or rel_l2(noise, gnoise) < MAP_TOL, f"{what}: noise_estimates"
    else:
        assert np.allclose(thr[fin], gthr[fin], rtol=STAT_RTOL, atol=0), f"{what}: thresholds"
        assert np.allclose(noise, gnoise, rtol=STAT_RTOL, atol=0), f"{what}: noise_estimates"
    assert not np.any(noise[~fin]), f"{what}: noise estimates outside the valid interior must be 0"
    diff = np.nonzero(mask != gmask)
    if diff[0].size:
        rel = np.abs(np.asarray(x, dtype=np.float64)[diff] - gthr[diff]) / np.abs(gthr[diff])
        assert np.all(rel < NEAR_THR), f"{what}: mask differs away from the threshold (rel {rel})"
    return int(diff[0].size)


def _dets_equal_or_near(got, want, what, thr=None, x=None):
    """Detection lists must be identical; when thresholds are known, cells within NEAR_THR may flip (counted)."""
    got = np.asarray(got)
    assert got.dtype.kind == "i" and got.ndim == 2 and got.shape[1] == 2, f"{what}: {got.dtype} {got.shape}"
    if got.shape == want.shape and np.array_equal(got, want):
        return 0
    assert thr is not None, f"{what}: detection lists differ ({got.shape[0]} vs {want.shape[0]})"
    a = {tuple(r) for r in got.tolist()} ^ {tuple(r) for r in want.tolist()}
    for r, c in a:
        assert abs(x[r, c] - thr[r, c]) / abs(thr[r, c]) < NEAR_THR, f"{what}: cell {(r, c)} flipped away from threshold"
    return len(a)


# ------------------------------------------------------------------------------------------------ every processor spec
@pytest.mark.parametrize("name", list(dp.CASES))
def test_processors_as_the_gui_drives_them(torch_cuda, name):
    """``cls(cm, **params)`` / ``process(adc_cube=cube, **params)`` / ``getattr(view_key)`` for every spec, host cubes frame
    by frame AND one device int16 batch, against the reference classes' outputs."""
    torch = torch_cuda
    cm = load_cfg(name, dp.CASES[name])
    g = _gold(name)
    frames = _frames(cm)
    cubes = _host_cubes(cm, frames)
    v = cubes[0][1].shape[0]
    batch = torch.from_numpy(np.stack([iq for _, iq in frames])).cuda()
    near_total = 0
    for key, module, cls_name, params, view_keys, vel in dp.specs(v):
        cls = _cls(module, cls_name)
        assert f"{key}|ctor|error" not in g
        # ---------------- host cubes, one frame at a time (stateful processors see the frames in golden order)
        proc = cls(cm, **params)
        for tag, cube in cubes:
            what = f"{name} {key} {tag} host"
            res = _call(proc, cube, params, vel)
            want = g[f"{key}|{tag}|data"]
            _axes_equal(proc, g, key, tag, view_keys)
            if cls_name in ("RangeDopplerProcessor", "RangeProcessor", "RangeAngleProcessor",
                            "RangeAngleProcessorDBSEnhanced"):
                assert isinstance(res, np.ndarray) and res.dtype == np.float64, what
                _map_close(res, want, what)
            elif cls_name == "RangeDetector":
                res = np.asarray(res)
                assert res.shape == want.shape and res.dtype == want.dtype, f"{what}: {res.dtype}{res.shape}"
                _map_close(proc.range_resp, g[f"{key}|{tag}|range_resp"], what + " range_resp")
                gthr = g[f"{key}|{tag}|thresholds"]
                fin = np.isfinite(gthr)
                assert np.array_equal(np.isfinite(proc.thresholds), fin), what
                assert np.allclose(proc.thresholds[fin], gthr[fin], rtol=1e-5, atol=0), what
                if not np.array_equal(res, want):
                    flips = set(res.tolist()) ^ set(want.tolist())
                    prof = g[f"{key}|{tag}|range_resp"]
                    assert all(abs(prof[i] - gthr[i]) / gthr[i] < NEAR_THR for i in flips), what
                    near_total += len(flips)
                assert np.array_equal(np.asarray(proc.dets), res)
            elif cls_name == "Altimeter":
                assert isinstance(res, float), what
                assert res == float(want), f"{what}: altitude {res} != {float(want)}"
                assert proc.current_altitude_corrected_m == float(g[f"{key}|{tag}|current_altitude_corrected_m"])
                _map_close(proc.coarse_fft_data, g[f"{key}|{tag}|coarse_fft_data"], what + " coarse_fft_data")
            elif cls_name.startswith("RangeDopplerDetector") or cls_name == "RangeDopplerGroundDetector":
                _map_close(proc.rng_dop_resp, g[f"{key}|{tag}|rng_dop_resp"], what + " rng_dop_resp")
                assert proc.rng_dop_resp.dtype == np.float64 and proc.rng_dop_resp_raw.dtype == np.complex128
                assert proc.rng_dop_resp_raw.shape == (v,) + proc.rng_dop_resp.shape
                if f"{key}|{tag}|rng_dop_resp_raw" in g:
                    _map_close(proc.rng_dop_resp_raw, g[f"{key}|{tag}|rng_dop_resp_raw"], what + " raw")
                    assert np.array_equal(np.abs(proc.rng_dop_resp_raw[0]).astype(np.float32),
                                          proc.rng_dop_resp.astype(np.float32)) or \
                        rel_l2(np.abs(proc.rng_dop_resp_raw[0]), proc.rng_dop_resp) < 1e-6
                thr = x = None
                if cls_name == "RangeDopplerDetector2D":
                    d = proc.detector
                    near_total += _stats_close(d.thresholds, d.noise_estimates, d.detections, g, f"{key}|{tag}|",
                                               g[f"{key}|{tag}|rng_dop_resp"], what, own_map=True)
                    thr, x = g[f"{key}|{tag}|thresholds"], g[f"{key}|{tag}|rng_dop_resp"].astype(np.float64)
                if cls_name == "RangeDopplerGroundDetector":
                    assert proc.altimeter.current_altitude_corrected_m == float(g[f"{key}|{tag}|altitude"]), what
                near_total += _dets_equal_or_near(res, want, what, thr, x)
                assert res is proc.dets and res.shape[1:] == (2,) and res.dtype.kind == "i"
                rv, vv, ri, vi = proc._map_detections_to_bins(res)
                if res.shape[0]:
                    assert np.array_equal(rv, proc.range_bins[res[:, 0]]) and np.array_equal(vv, proc.vel_bins[res[:, 1]])
            elif cls_name == "PointCloudGenerator":
                assert isinstance(res, np.ndarray) and res.dtype == np.float64 and res.ndim == 2 and res.shape[1] == 4, what
                gd = g[f"{key}|{tag}|dets"]
                if _dets_equal_or_near(proc.detector.dets, gd, what + " dets",
                                       *(_thr_of(proc) if hasattr(proc.detector, "detector") else (None, None))):
                    near_total += 1
                    continue                      # a threshold flip changes the cloud's rows; counted, not compared
                assert res.shape == want.shape, f"{what}: {res.shape} != {want.shape}"
                if want.shape[0]:
                    assert np.allclose(res, want, rtol=1e-9, atol=1e-12), f"{what}: points differ"
                    # per-detection angles through the reference's helper (ref :143-214): bit-exact bins
                    _, _, ri, di = proc.detector._map_detections_to_bins(proc.detector.dets)
                    az, el = proc._compute_angle_estimation(proc.detector.rng_dop_resp_raw, ri, di)
                    assert np.array_equal(az, g[f"{key}|{tag}|az"], equal_nan=True), what + " az angles"
                    assert np.array_equal(el, g[f"{key}|{tag}|el"], equal_nan=True), what + " el angles"
            else:
                raise AssertionError(f"unhandled spec {cls_name}")

        # ---------------- ONE device int16 batch of all frames (fresh instance)
        if cls_name == "RangeDetector":
            continue                              # host-only class (its CFAR object takes numpy), as in the reference
        proc = cls(cm, **params)
        res = _call(proc, batch, params, vel)
        for f, (tag, _) in enumerate(cubes):
            what = f"{name} {key} {tag} device batch"
            want = g[f"{key}|{tag}|data"]
            if cls_name in ("RangeDopplerProcessor", "RangeProcessor", "RangeAngleProcessor",
                            "RangeAngleProcessorDBSEnhanced"):
                assert res.is_cuda and res.dtype == torch.float32 and res.shape[0] == len(cubes)
                _map_close(res[f].cpu().numpy(), want, what)
            elif cls_name == "Altimeter":
                assert float(res[f]) == float(want), what
            elif cls_name == "PointCloudGenerator":
                pts, counts = res
                n = int(counts[f])
                gd = g[f"{key}|{tag}|dets"]
                got_d = proc.detector._dets_dev[f, :n].cpu().numpy()
                if n != gd.shape[0] or not np.array_equal(got_d, gd):
                    near_total += 1
                    assert abs(n - gd.shape[0]) <= 2, f"{what}: {n} detections vs {gd.shape[0]}"
                    continue
                assert pts.dtype == torch.float64
                if n:
                    assert np.allclose(pts[f, :n].cpu().numpy(), want, rtol=1e-9, atol=1e-12), what
            else:
                n = int(proc._counts_dev[f])
                got_d = res[f, :n].cpu().numpy().astype(np.int64)
                thr = x = None
                if cls_name == "RangeDopplerDetector2D":
                    thr, x = g[f"{key}|{tag}|thresholds"], g[f"{key}|{tag}|rng_dop_resp"].astype(np.float64)
                near_total += _dets_equal_or_near(got_d.reshape(-1, 2), want, what, thr, x)
                _map_close(proc.rng_dop_resp[f].cpu().numpy(), g[f"{key}|{tag}|rng_dop_resp"], what + " rng_dop_resp")
    print(f"{name}: {near_total} CFAR cells within {NEAR_THR} of their threshold flipped across all specs")


def _thr_of(pcg):
    """(thresholds, map) of a PointCloudGenerator whose detector is the 2-D one (host mode caches them)."""
    d = pcg.detector.detector
    return d.thresholds, pcg.detector.rng_dop_resp


# ------------------------------------------------------------------------------------------------ RangeDopplerProcessor grid
@pytest.mark.parametrize("name", list(dp.CASES))
def test_range_doppler_processor_argument_grid(torch_cuda, name):
    """rx_idx -1 / 0 / last x magnitude / complex (reference: processors/range_doppler_resp.py:49-110)."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.processors.range_doppler_resp import RangeDopplerProcessor
    cm = load_cfg(name, dp.CASES[name])
    g = _gold(name)
    frames = _frames(cm)
    tag0, cube0 = _host_cubes(cm, frames[:1])[0]
    v = cube0.shape[0]
    dev = torch.from_numpy(frames[0][1][None]).cuda()
    for rx, mag in dp.rd_variants(v):
        want = g[f"rd_variant_rx{rx}_mag{int(mag)}|{tag0}|data"]
        res = RangeDopplerProcessor(cm).process(adc_cube=cube0, rx_idx=rx, return_magnitude=mag)
        assert res.dtype == (np.float64 if mag else np.complex128), (rx, mag, res.dtype)
        _map_close(res, want, f"{name} rx_idx={rx} magnitude={mag} host")
        res_d = RangeDopplerProcessor(cm).process(adc_cube=dev, rx_idx=rx, return_magnitude=mag)
        assert res_d.is_cuda and res_d.dtype == (torch.float32 if mag else torch.complex64)
        _map_close(res_d[0].cpu().numpy(), want, f"{name} rx_idx={rx} magnitude={mag} device")


# ------------------------------------------------------------------------------------------------ stand-alone detectors
@pytest.mark.parametrize("name", list(dp.CASES))
def test_cfar_detector_classes(torch_cuda, name):
    """``registry[key](**params).detect(x)`` + cached ``thresholds`` / ``noise_estimates`` / ``detections`` on the
    reference's fp64 RD map, three of its rows and the range profile (reference: detectors/base.py:38-65,208-230)."""
    from mmwave_radar_processing_b200.detectors.detector_registry import get_detector_registry
    g = _gold(name)
    reg = get_detector_registry()
    tag0 = f"s{dp.SEEDS[0]}"
    x2d = g[f"range_doppler_resp|{tag0}|data"]
    prof = g[f"range_resp|{tag0}|data"]
    assert x2d.dtype == np.float64 and prof.dtype == np.float64
    near = 0
    two_d = [(k, k, p) for k, p in dp.CFAR_2D.items()] + [(t, k, p) for t, (k, p) in dp.CFAR_2D_EXTRA.items()]
    for tag, kind, params in two_d:
        d = reg[kind](**params)
        lst = d.detect(x2d)
        assert isinstance(lst, list) and all(isinstance(t, tuple) and len(t) == 2 for t in lst)
        near += _stats_close(d.thresholds, d.noise_estimates, d.detections, g, f"cfar|{tag}|", x2d, f"{name} {tag}")
        got = np.array(lst, dtype=np.int64).reshape(-1, 2)
        assert np.array_equal(got, np.argwhere(d.detections)), f"{name} {tag}: list is not np.where order of the mask"
        near += _dets_equal_or_near(got, g[f"cfar|{tag}|list"], f"{name} {tag}", g[f"cfar|{tag}|thresholds"], x2d)
    rows = g["cfar|rows"]
    one_d = [(k, k, p) for k, p in dp.CFAR_1D.items()] + [("ca_too_large", *dp.CFAR_1D_TOO_LARGE)]
    for tag, kind, params in one_d:
        d = reg[kind](**params)
        for sig_tag, sig in [("prof", prof)] + [(f"row{j}", x2d[rr]) for j, rr in enumerate(rows)]:
            lst = d.detect(sig)
            assert isinstance(lst, list) and all(isinstance(i, int) for i in lst)
            pre = f"cfar|{tag}|{sig_tag}|"
            near += _stats_close(d.thresholds, d.noise_estimates, d.detections, g, pre, sig, f"{name} {tag} {sig_tag}")
            assert lst == np.nonzero(d.detections)[0].tolist()
            want = g[pre + "list"].tolist()
            if lst != want:
                gthr = g[pre + "thresholds"]
                assert all(abs(sig[i] - gthr[i]) / gthr[i] < NEAR_THR for i in set(lst) ^ set(want))
    with pytest.raises(ValueError):
        reg["ca_cfar_2d"](**dp.CFAR_2D["ca_cfar_2d"]).detect(prof)
    with pytest.raises(ValueError):
        reg["os_cfar_1d"](**dp.CFAR_1D["os_cfar_1d"]).detect(x2d)
    print(f"{name}: {near} stand-alone CFAR cells within {NEAR_THR} of their threshold flipped")


# ------------------------------------------------------------------------------------------------ ground-detector sequence
@pytest.mark.parametrize("name", list(dp.CASES))
def test_ground_detector_sequence(torch_cuda, name):
    """20 scripted frames through ONE PointCloudGenerator with the GUI's default (ground) detector: the altimeter
    recurrence (processors/altimeter.py:83-139), the per-frame range band (range_doppler_ground_detector.py:44-103) and the
    clouds, frame by frame (host cubes) and as one device batch."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.processors.altimeter import Altimeter
    from mmwave_radar_processing_b200.processors.point_cloud_generator import PointCloudGenerator
    from mmwave_radar_processing_b200.processors.virtual_array_reformater import VirtualArrayReformatter
    from mmwave_radar_processing_b200.synthetic import iq_to_complex_cube, synth_scene_iq
    cm = load_cfg(name, dp.CASES[name])
    g = _gold(name)
    r, s, cpl, l = _shape(cm)
    v = r * cpl
    params = dp.seq_params(v, l)
    iqs = [synth_scene_iq(r, s, cpl, l, 500 + f, dp.ground_sequence_targets(f, cm.range_res_m, s, l))
           for f in range(dp.SEQ_FRAMES)]
    var = VirtualArrayReformatter(cm)
    pcg = PointCloudGenerator(cm, **params)
    lat = []
    n_pts = 0
    for f, iq in enumerate(iqs):
        cube = var.process(iq_to_complex_cube(iq))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pts = pcg.process(adc_cube=cube, **params)
        lat.append(time.perf_counter() - t0)
        assert pcg.detector.altimeter.current_altitude_corrected_m == float(g[f"seq|f{f}|altitude"]), f"frame {f} altitude"
        _same_array(pcg.detector.dets, g[f"seq|f{f}|dets"], f"{name} seq frame {f} dets")
        want = g[f"seq|f{f}|data"]
        assert pts.shape == want.shape and pts.dtype == np.float64
        assert np.allclose(pts, want, rtol=1e-9, atol=1e-12), f"{name} seq frame {f} points"
        n_pts += pts.shape[0]
    assert n_pts > 0
    print(f"{name}: PointCloudGenerator.process(host cube) median {1e3 * float(np.median(lat[2:])):.3f} ms/frame "
          f"(reference on this path: ~11 ms/frame on one host core, BASELINE.md section 2)")
    # the same 20 frames as ONE device batch: fresh generator, fresh altimeter
    batch = torch.from_numpy(np.stack(iqs)).cuda()
    alts = Altimeter(cm, **dp.ALTIMETER).process(adc_cube=batch, **dp.ALTIMETER)
    assert np.array_equal(alts, np.array([float(g[f"seq|f{f}|altitude"]) for f in range(dp.SEQ_FRAMES)]))
    pcg = PointCloudGenerator(cm, **params)
    pts, counts = pcg.process(adc_cube=batch, **params)
    for f in range(dp.SEQ_FRAMES):
        n = int(counts[f])
        gd = g[f"seq|f{f}|dets"]
        assert n == gd.shape[0] and np.array_equal(pcg.detector._dets_dev[f, :n].cpu().numpy(), gd), f"batch frame {f}"
        assert np.allclose(pts[f, :n].cpu().numpy(), g[f"seq|f{f}|data"], rtol=1e-9, atol=1e-12)
    # reset() restores the initial altitude (reference: altimeter.py reset / ground detector reset)
    pcg.reset()
    assert pcg.detector.altimeter.current_altitude_measured_m == dp.ALTIMETER["min_altitude_m"]


# ------------------------------------------------------------------------------------------------ small contracts
def test_empty_cloud_and_error_contracts(torch_cuda):
    """(0, 4) float cloud / (0, 2) int detections on an empty frame; constructor / argument errors as the reference."""
    from mmwave_radar_processing_b200.processors.point_cloud_generator import PointCloudGenerator
    from mmwave_radar_processing_b200.processors.range_doppler_detection.range_doppler_detector_2d import \
        RangeDopplerDetector2D
    from mmwave_radar_processing_b200.processors.range_detector import RangeDetector
    cm = load_cfg("1843_RadVel")
    r, s, cpl, l = _shape(cm)
    zero = np.zeros((r * cpl, s, l), dtype=complex)
    pcg = PointCloudGenerator(cm, az_antenna_idxs=[0, 1, 2, 3], el_antenna_idxs=[],
                              detector_params=dict(cfar_type="ca_cfar_2d",
                                                   cfar_params=dict(num_train=(4, 4), num_guard=(2, 2), pfa=1e-5)))
    pts = pcg.process(adc_cube=zero)
    assert pts.shape == (0, 4) and pts.dtype == np.float64
    assert pcg.detector.dets.shape == (0, 2) and pcg.detector.dets.dtype.kind == "i"
    with pytest.raises(ValueError):
        PointCloudGenerator(cm, az_antenna_idxs=[0], el_antenna_idxs=[], detector_type="no_such_detector")
    with pytest.raises(ValueError):
        PointCloudGenerator(cm, az_antenna_idxs="0,1", el_antenna_idxs=[])
    with pytest.raises(ValueError):
        RangeDopplerDetector2D(cm, cfar_type="no_such_cfar")
    with pytest.raises(ValueError):
        RangeDetector(cm, cfar_type="no_such_cfar")
    with pytest.raises(ValueError):
        RangeDopplerDetector2D(cm, cfar_type="ca_cfar_2d",
                               cfar_params=dict(num_train=(4, 4), num_guard=(2, 2), pfa=1e-5)).process(zero[0])


def test_virtual_array_reformatter_is_the_oracles_regroup(torch_cuda):
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.processors.virtual_array_reformater import VirtualArrayReformatter
    from mmwave_radar_processing_b200.synthetic import iq_to_complex_cube
    for name, geom in dp.CASES.items():
        cm = load_cfg(name, geom)
        r, s, cpl, l = _shape(cm)
        iq = _frames(cm)[0][1]
        got = VirtualArrayReformatter(cm).process(iq_to_complex_cube(iq))
        want = orc.virtual_array(orc.iq_to_cube(iq), r, cpl, l)
        assert got.shape == (r * cpl, s, l) and got.dtype == np.complex128 and np.array_equal(got, want)
