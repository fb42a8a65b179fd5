"""Pin the numpy oracle against outputs of the reference classes (tests/golden/*.npz)."""
import json
import os

import numpy as np
import pytest

from conftest import (GOLDEN_CASES, GOLDEN_DIR, CFAR2D_SETS, CFAR1D_SETS, SEQ_PARAMS, GROUND_VEL,
                      ant_sets, frame_iq, golden, load_cfg, rel_l2)
from oracle import radar_oracle as orc


def _virtual(cm, seed):
    cube = orc.iq_to_cube(frame_iq(cm, seed))
    return orc.virtual_array(cube, cm.num_rx_antennas, cm.chirp_cfgs_per_loop, cm.frameCfg_loops)


@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_rd_and_profiles(name, seed):
    cm = load_cfg(name)
    g = golden(name, seed)
    cube = _virtual(cm, seed)
    rd = orc.range_doppler(cube)
    if "rd" in g:
        assert np.array_equal(cube.astype(np.complex64), g["virtual_cube"])
        assert rel_l2(rd, g["rd"]) < 1e-13
    assert rel_l2(np.abs(rd[0]), g["rd_mag_rx0"]) < 1e-13
    assert rel_l2(np.abs(rd[-1]), g["rd_mag_rx_last"]) < 1e-13
    assert rel_l2(orc.coarse_range_profile(cube, 0), g["range_profile"]) < 1e-13
    assert rel_l2(orc.coarse_range_profile(cube, 3), g["range_profile_c3"]) < 1e-13


@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_cfar2d_and_point_cloud(name, seed):
    cm = load_cfg(name)
    g = golden(name, seed)
    cube = _virtual(cm, seed)
    rd = orc.range_doppler(cube)
    mag0 = orc.rd_magnitude0(rd)
    rb, vb = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
    sets = ant_sets(cube.shape[0])
    for tag, (kind, params) in CFAR2D_SETS.items():
        thr, noise, mask = orc.cfar_2d(mag0, kind, **params)
        gthr = g[f"thr2d_{tag}"]
        assert np.array_equal(np.isinf(thr), np.isinf(gthr))
        fin = np.isfinite(gthr)
        assert np.allclose(thr[fin], gthr[fin], rtol=1e-12, atol=0)
        assert np.allclose(noise, g[f"noise2d_{tag}"], rtol=1e-12, atol=0)
        dets = orc.mask_to_dets(mask)
        assert dets.shape == g[f"dets2d_{tag}"].shape and np.array_equal(dets, g[f"dets2d_{tag}"])
        for ab in ("a", "b"):
            pts, az_i, el_i = orc.point_cloud(rd, dets, rb, vb, sets["az_" + ab], sets["el_" + ab])
            gp = g[f"pc_{tag}_{ab}"]
            assert pts.shape == gp.shape
            if gp.shape[0]:
                assert np.allclose(pts, gp, rtol=1e-12, atol=1e-12)
                _, bins = orc.angle_axis(64)
                if sets["az_" + ab]:
                    assert np.array_equal(bins[az_i], g[f"az_{tag}_{ab}"])
                if sets["el_" + ab]:
                    assert np.array_equal(bins[el_i], g[f"el_{tag}_{ab}"])


@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_cfar1d_family_and_detectors(name, seed):
    cm = load_cfg(name)
    g = golden(name, seed)
    cube = _virtual(cm, seed)
    rd = orc.range_doppler(cube)
    mag0 = orc.rd_magnitude0(rd)
    prof = orc.coarse_range_profile(cube, 0)
    rows = g["cfar1d_rows"]
    for tag, (kind, params) in CFAR1D_SETS.items():
        for label, x in [("prof", prof)] + [(f"row{j}", mag0[r]) for j, r in enumerate(rows)]:
            thr, noise, mask = orc.cfar_1d(x, kind, **params)
            gthr = g[f"c1_{label}_{tag}_thr"]
            assert np.array_equal(np.isinf(thr), np.isinf(gthr))
            fin = np.isfinite(gthr)
            assert np.allclose(thr[fin], gthr[fin], rtol=1e-12, atol=0)
            assert np.allclose(noise, g[f"c1_{label}_{tag}_noise"], rtol=1e-12, atol=0)
            assert np.array_equal(np.nonzero(mask)[0], g[f"c1_{label}_{tag}_idx"])
    # range detector (os 1-D on the profile)
    thr, _, mask = orc.cfar_1d(prof, "os", num_train=5, num_guard=3, rho=0.5, alpha=2)
    assert np.array_equal(np.nonzero(mask)[0].astype(float), g["range_detector_dets"])
    # sequential
    dets = orc.detect_sequential(cube, mag0, **SEQ_PARAMS)
    assert dets.shape == g["dets_seq"].shape and np.array_equal(dets, g["dets_seq"])
    # ground (CFAR part, altitude taken from the reference run)
    if "dets_ground" in g:
        rb, _ = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
        d2 = orc.detect_range_band(mag0, rb, float(g["ground_altitude"]), GROUND_VEL[0], GROUND_VEL[1])
        assert d2.shape == g["dets_ground"].shape and np.array_equal(d2, g["dets_ground"])


@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_range_angle(name, seed):
    cm = load_cfg(name)
    g = golden(name, seed)
    cube = _virtual(cm, seed)
    sets = ant_sets(cube.shape[0])
    for a in (64, 256):
        assert rel_l2(orc.range_angle(cube, a, 0, sets["ra"]), g[f"ra_{a}_sub"]) < 1e-13
        assert rel_l2(orc.range_angle(cube, a, 2, []), g[f"ra_{a}_all_c2"]) < 1e-13
    assert rel_l2(orc.range_angle(cube, 64, 1, sets["ra"], perform_windowing=False), g["ra_64_nowin"]) < 1e-13


RA_HIRES = {"RadarHD": (0, 1), "RadCloud": (0,)}


@pytest.mark.parametrize("name", sorted(RA_HIRES))
def test_range_angle_1024_bins(name):
    """BASELINE.json configs[2]: zero-padded range-azimuth maps with 1024 angle bins against the reference's
    RangeAngleProcessor (tests/golden/make_golden_ra_hires.py): stored rows, fp64 row sums and argmax of every row."""
    cm = load_cfg(name)
    g = np.load(os.path.join(GOLDEN_DIR, f"ra1024_{name}.npz"))
    for seed in RA_HIRES[name]:
        cube = orc.virtual_array(orc.iq_to_cube(frame_iq(cm, seed)), cm.num_rx_antennas, cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
        for tag, chirp, ants in (("all", 0, ()), ("sub_c1", 1, (0, 1, 3))):
            ra = orc.range_angle(cube, 1024, chirp, ants)
            key = f"s{seed}_{tag}"
            assert tuple(g[key + "_shape"]) == ra.shape
            assert rel_l2(ra[::8], g[key + "_rows"]) < 1e-6            # rows are stored as float32
            assert np.allclose(ra.sum(axis=1), g[key + "_rowsum"], rtol=1e-12)
            top2 = np.partition(ra, -2, axis=1)[:, -2:]
            clear = (top2[:, 1] - top2[:, 0]) > 1e-9 * top2[:, 1]        # a single live antenna gives a flat row: no argmax to pin
            assert np.array_equal(ra.argmax(axis=1)[clear], g[key + "_argmax"][clear])


@pytest.mark.parametrize("name", ["RadSAR_1843_official", "6843_RadVel_ods_20Hz"])
def test_dbs(name):
    cm = load_cfg(name)
    g = golden(name, 0)
    cube = _virtual(cm, 0)
    _, vb = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
    vel = g["dbs_velocity"]
    assert rel_l2(orc.dbs_enhanced(cube, vel, vb, 64, 100, [0, 3, 4, 7]), g["dbs_sub"]) < 1e-13
    assert rel_l2(orc.dbs_enhanced(cube, vel, vb, 64, 100, []), g["dbs_all"]) < 1e-13
    # below the speed gate the reference falls back to the plain range-angle map
    assert rel_l2(orc.range_angle(cube, 64, 1, [0, 3, 4, 7]), g["dbs_slow"]) < 1e-13


def test_sar():
    from mmwave_radar_processing_b200.synthetic import synth_frame_iq
    name = "RadSAR_1843_official"
    cm = load_cfg(name)
    g = golden(name, 0)
    r, s, cpl, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    prof = cm.profile_cfgs[0]
    for tag in ("a", "b"):
        rx, cfg_idx, nf, stride = [int(v) for v in g[f"sar_{tag}_params"]]
        mask, times = orc.sar_chirp_times_us(cpl, l, prof["idleTime_us"] + prof["rampEndTime_us"], cfg_idx, stride)
        assert np.allclose(times, g[f"sar_{tag}_chirp_times_us"], rtol=0, atol=0)
        hist = np.stack([orc.iq_to_cube(synth_frame_iq(r, s, cpl, l, 1000 + f))[rx][:, mask] for f in range(nf)])
        hv = g[f"sar_{tag}_hist_vel"]
        geo = orc.sar_geometry(hv, times, cm.frameCfg_periodicity_ms)
        assert np.allclose(geo, g[f"sar_{tag}_geometry"], rtol=1e-13, atol=1e-18)
        assert bool(g[f"sar_{tag}_valid"][-1]) == orc.sar_velocity_gate(
            hv, np.array([0.0, 0.3, 0.0]), np.array([0.05, 0.7, 0.05]), np.array([0.02, 0.02, 0.02]))
        d = orc.sar_steering(g[f"sar_{tag}_az"], g[f"sar_{tag}_el"])
        lam = orc.C_M_S / (prof["startFreq_GHz"] * 1e9)
        assert lam == float(g[f"sar_{tag}_lambda"])
        resp = orc.sar_response(hist, geo, d, lam)
        assert resp.shape == g[f"sar_{tag}_resp"].shape
        assert rel_l2(resp, g[f"sar_{tag}_resp"]) < 1e-11


def test_cfg_table_matches_reference():
    with open(os.path.join(GOLDEN_DIR, "cfg_table.json")) as fh:
        table = json.load(fh)
    for name, row in table["cfgs"].items():
        cm = load_cfg(name)
        assert cm.num_rx_antennas == row["num_rx"] and cm.num_tx_antennas == row["num_tx"]
        assert cm.chirp_cfgs_per_loop == row["cpl"] and cm.get_num_adc_samples(0) == row["S"]
        assert cm.frameCfg_loops == row["L"] and cm.virtual_antennas_enabled == row["virtual"]
        for k_ref, k_attr in (("range_res_m", "range_res_m"), ("range_max_m", "range_max_m"),
                              ("vel_res_m_s", "vel_res_m_s"), ("vel_max_m_s", "vel_max_m_s"),
                              ("range_bin_size_m", "range_bin_size_m")):
            assert float(getattr(cm, k_attr)) == row[k_ref], (name, k_ref)     # bit-for-bit
        rb, vb = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
        assert rb.shape[0] == row["n_range_bins"] == row["S"]
        assert vb.shape[0] == row["n_vel_bins"] == row["L"]
        assert rb[-1] == row["range_bins_last"] and vb[0] == row["vel_bins_first"] and vb[-1] == row["vel_bins_last"]
        assert cm.adcCfg_num_adc_bits == row["adc_bits"] and cm.adcbufCfg_adc_output_fmt == row["adcbuf_fmt"]
        assert cm.adcbufCfg_channel_interleave == row["interleave"]
        assert cm.frameCfg_periodicity_ms == row["frame_ms"] and cm.frameCfg_hardware_trigger_enabled == row["hw_trig"]
    for a, q in table["angle_axis"].items():
        with np.errstate(invalid="ignore"):
            _, bins = orc.angle_axis(int(a))
        assert bins.shape[0] == q["n"]
        for key, idx in (("first", 0), ("last", -1), ("mid", int(a) // 2)):
            ref = q[key]
            assert (np.isnan(ref) and np.isnan(bins[idx])) or bins[idx] == ref
