"""GPU parity tests: the CUDA path (through the C ABI) against the numpy oracle and the reference-generated
golden fixtures, on the same seeded inputs.

Tolerances (BASELINE.json north_star): bin / peak indices bit-exact; RD and RA magnitudes <= 1e-4 relative
L2 (fp32 device arithmetic vs fp64 reference); CFAR masks identical except cells within 1e-5 (relative) of
the threshold, which are counted and reported.
"""
import os

import numpy as np
import pytest

from conftest import (CFAR1D_SETS, CFAR2D_SETS, GOLDEN_CASES, GOLDEN_DIR, GROUND_VEL, SEQ_PARAMS, ant_sets, frame_iq, golden,
                      load_cfg, rel_l2, REPO)

pytestmark = pytest.mark.gpu

RD_TOL = 1e-4
NEAR_THR = 1e-5


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch


def _engine(cm):
    from mmwave_radar_processing_b200.engine import RadarEngine
    return RadarEngine.for_config(cm)


def _oracle_frame(cm, seed):
    from oracle import radar_oracle as orc
    iq = frame_iq(cm, seed)
    cube = orc.virtual_array(orc.iq_to_cube(iq), cm.num_rx_antennas, cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    return iq, cube, orc.range_doppler(cube)


def _mask_diff_report(x, thr, mask_dev, mask_ref):
    """Cells where the masks differ must all be within NEAR_THR of the threshold."""
    diff = np.nonzero(mask_dev != mask_ref)
    if diff[0].size == 0:
        return 0
    rel = np.abs(x[diff] - thr[diff]) / np.abs(thr[diff])
    assert np.all(rel < NEAR_THR), f"mask differs away from the threshold: rel distances {rel}"
    return int(diff[0].size)


# ------------------------------------------------------------------------------------------------ RD
@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_range_doppler_both_layouts(torch_cuda, name, seed):
    torch = torch_cuda
    cm = load_cfg(name)
    g = golden(name, seed)
    iq, cube, rd_ref = _oracle_frame(cm, seed)
    eng = _engine(cm)
    rd, mag = eng.range_doppler(torch.from_numpy(iq[None]).cuda(), want_rd=True, mag_antenna=0)
    rd = rd[0].cpu().numpy()
    assert rel_l2(rd, rd_ref) < RD_TOL
    assert rel_l2(mag[0].cpu().numpy(), np.abs(rd_ref[0])) < RD_TOL
    assert rel_l2(mag[0].cpu().numpy(), g["rd_mag_rx0"]) < RD_TOL
    if "rd" in g:
        assert rel_l2(rd, g["rd"]) < RD_TOL
    # complex64 cube layout (the reference's host cube)
    from mmwave_radar_processing_b200.processors._device import engine_for_cube_shape
    e2 = engine_for_cube_shape(*cube.shape)
    rd2, mag_all = e2.range_doppler(torch.from_numpy(cube.astype(np.complex64)[None]).cuda(), want_rd=True, mag_antenna=-1)
    assert rel_l2(rd2[0].cpu().numpy(), rd_ref) < RD_TOL
    assert rel_l2(mag_all[0].cpu().numpy(), np.abs(rd_ref)) < RD_TOL
    # the two ingest layouts see the same values (the int16 path may run a shape-specialised kernel with another
    # radix order, so agreement is to fp32 rounding, not bit for bit)
    assert rel_l2(rd2[0].cpu().numpy(), rd) < 1e-6
    # antenna subset + un-windowed variant
    sub = [cube.shape[0] - 1, 0]
    rd3, _ = eng.range_doppler(torch.from_numpy(iq[None]).cuda(), ant_list=sub, window_flags=0)
    ref3 = np.fft.fftshift(np.fft.fft2(cube[sub], axes=(-2, -1)), axes=-1)
    assert rel_l2(rd3[0].cpu().numpy(), ref3) < RD_TOL


@pytest.mark.parametrize("name", ["1843_RadVel_10Hz", "1843_RadVel_5Hz", "1843_vel_nav", "1843_RaGNNarok_UAV_5m",
                                  "1843_RaGNNarok_UAV_10m", "6843_vel_sr", "RadSAR_lr", "6843_hand_ods_30Hz",
                                  "1843_RadVel_10Hz_lr_SAR", "RadCloud_original", "1843_RaGNNarok"])
def test_range_doppler_every_fft_length(torch_cuda, name):
    """All (S, L) pairs of the 25 cfgs incl. odd tile sizes (63x115), primes (127, 23, 13) and 512-point FFTs."""
    torch = torch_cuda
    cm = load_cfg(name)
    iq, cube, rd_ref = _oracle_frame(cm, 7)
    eng = _engine(cm)
    rd, mag = eng.range_doppler(torch.from_numpy(iq[None]).cuda(), want_rd=True, mag_antenna=0)
    assert rel_l2(rd[0].cpu().numpy(), rd_ref) < RD_TOL
    assert rel_l2(mag[0].cpu().numpy(), np.abs(rd_ref[0])) < RD_TOL


def test_range_profile(torch_cuda):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    for name, seed in GOLDEN_CASES:
        cm = load_cfg(name)
        g = golden(name, seed)
        iq, cube, _ = _oracle_frame(cm, seed)
        eng = _engine(cm)
        dev = torch.from_numpy(iq[None]).cuda()
        assert rel_l2(eng.range_profile(dev, 0)[0].cpu().numpy(), g["range_profile"]) < RD_TOL
        assert rel_l2(eng.range_profile(dev, 3)[0].cpu().numpy(), orc.coarse_range_profile(cube, 3)) < RD_TOL


# ------------------------------------------------------------------------------------------------ CFAR
@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_cfar2d_and_point_cloud(torch_cuda, name, seed):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import cfar_params
    from mmwave_radar_processing_b200.pipeline import reference_axes
    cm = load_cfg(name)
    g = golden(name, seed)
    iq, cube, rd_ref = _oracle_frame(cm, seed)
    eng = _engine(cm)
    rb, vb, ab = reference_axes(cm, 64)
    eng.set_axes(rb, vb)
    eng.set_angle_bins(ab)
    rd, mag = eng.range_doppler(torch.from_numpy(iq[None]).cuda(), want_rd=True, mag_antenna=0)
    mag_h = mag[0].cpu().numpy().astype(np.float64)
    sets = ant_sets(cube.shape[0])
    near = 0
    for tag, (kind, params) in CFAR2D_SETS.items():
        p = cfar_params(kind, params["num_train"], params["num_guard"], pfa=params.get("pfa", 0.0),
                        rho=params.get("rho", 0.0), alpha=params.get("alpha", 0.0))
        res = eng.cfar(mag, p, want_noise=True, want_mask=True)
        noise = res["noise"][0].cpu().numpy()
        mask = res["mask"][0].cpu().numpy().astype(bool)
        # (1) exact against the oracle evaluated on the SAME fp32 map: isolates the CFAR kernel
        thr_o, noise_o, mask_o = orc.cfar_2d(mag_h, kind, **params)
        assert np.allclose(noise, noise_o, rtol=1e-12, atol=0)
        assert np.array_equal(mask, mask_o)
        n = int(res["counts"][0])
        dets = res["dets"][0, :n].cpu().numpy()
        assert np.array_equal(dets, orc.mask_to_dets(mask_o))
        # mask-only call (streaming path: OS-CFAR decides by counting, no order statistic): same mask, same list
        res2 = eng.cfar(mag, p, want_noise=False, want_mask=True)
        assert np.array_equal(res2["mask"][0].cpu().numpy().astype(bool), mask_o)
        assert int(res2["counts"][0]) == n and np.array_equal(res2["dets"][0, :n].cpu().numpy(), dets)
        # (2) against the fp64 reference run: identical except cells within 1e-5 of the threshold
        gthr = g[f"thr2d_{tag}"]
        ref_mask = np.zeros_like(mask)
        gd = g[f"dets2d_{tag}"]
        ref_mask[gd[:, 0], gd[:, 1]] = True
        fin = np.isfinite(gthr)
        assert rel_l2(noise[fin], g[f"noise2d_{tag}"][fin]) < RD_TOL
        flips = _mask_diff_report(g["rd_mag_rx0"], gthr, mask, ref_mask)
        near += flips
        if flips == 0:
            assert np.array_equal(dets, gd)
            for ab_tag in ("a", "b"):
                az, el = sets["az_" + ab_tag], sets["el_" + ab_tag]
                pts, az_bin, el_bin = eng.point_cloud(rd, res["dets"], res["counts"], az, el, True, False, 64)
                gp = g[f"pc_{tag}_{ab_tag}"]
                assert pts[0, :n].shape == gp.shape
                if n:
                    _, az_o, el_o = orc.point_cloud(rd_ref, gd, rb, vb, az, el)
                    if az:
                        assert np.array_equal(az_bin[0, :n].cpu().numpy(), az_o)        # bit-exact indices
                    if el:
                        assert np.array_equal(el_bin[0, :n].cpu().numpy(), el_o)
                    assert np.allclose(pts[0, :n].cpu().numpy(), gp, rtol=1e-12, atol=1e-12)
    print(f"{name} seed {seed}: {near} CFAR cells within {NEAR_THR} of threshold flipped")


def test_cfar_edge_cases(torch_cuda):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import cfar_params, utility_engine
    eng = utility_engine()
    rng = np.random.default_rng(3)
    # window larger than the map -> no detections, noise 0 (reference swallows the ValueError)
    x = rng.exponential(1.0, (9, 40)).astype(np.float32)
    res = eng.cfar(torch.from_numpy(x[None]).cuda(), cfar_params("os", (8, 4), (2, 1), rho=0.75, alpha=3.5),
                   want_noise=True, want_mask=True)
    assert int(res["counts"][0]) == 0 and not res["mask"].any() and not res["noise"].any()
    # heavy-detection map with truncation: count reports the true number, list holds the first max_dets in order
    x = rng.exponential(1.0, (40, 50)).astype(np.float32)
    p = cfar_params("ca", (3, 3), (1, 1), alpha=0.9)
    full = eng.cfar(torch.from_numpy(x[None]).cuda(), p, want_mask=True)
    n = int(full["counts"][0])
    _, _, mask_o = orc.cfar_2d(x.astype(np.float64), "ca", (3, 3), (1, 1), pfa=0.5)   # alpha overridden below
    cut = eng.cfar(torch.from_numpy(x[None]).cuda(), p, max_dets=17)
    assert int(cut["counts"][0]) == n > 17
    assert np.array_equal(cut["dets"][0].cpu().numpy(), full["dets"][0, :17].cpu().numpy())
    d = full["dets"][0, :n].cpu().numpy()
    lin = d[:, 0] * 50 + d[:, 1]
    assert np.all(np.diff(lin) > 0)                                     # row-major, strictly increasing
    assert np.array_equal(np.nonzero(full["mask"][0].cpu().numpy().ravel())[0], lin)
    # ties: constant map -> nothing exceeds alpha*noise for alpha >= 1, everything for alpha < 1 (strict >)
    c = np.full((1, 30, 30), 2.0, dtype=np.float32)
    for kind, kw in (("ca", dict(alpha=1.0)), ("os", dict(rho=0.5, alpha=1.0))):
        r = eng.cfar(torch.from_numpy(c).cuda(), cfar_params(kind, (2, 2), (1, 1), **kw))
        assert int(r["counts"][0]) == 0
    r = eng.cfar(torch.from_numpy(c).cuda(), cfar_params("os", (2, 2), (1, 1), rho=0.5, alpha=0.999))
    assert int(r["counts"][0]) == 24 * 24
    # OS rank extremes: rho -> k=1 (minimum) and k=N (maximum)
    x = rng.exponential(1.0, (25, 31)).astype(np.float32)
    for rho in (0.0, 1.0, 0.33):
        r = eng.cfar(torch.from_numpy(x[None]).cuda(), cfar_params("os", (3, 2), (1, 1), rho=rho, alpha=1.0), want_noise=True)
        _, noise_o, _ = orc.cfar_2d(x.astype(np.float64), "os", (3, 2), (1, 1), rho=rho, alpha=1.0)
        assert np.array_equal(r["noise"][0].cpu().numpy(), noise_o)
    # large OS window (> 256 and > 512 cells) exercises the wider per-lane register files
    x = rng.exponential(1.0, (64, 64)).astype(np.float32)
    for tr, gr in (((9, 9), (2, 2)), ((12, 12), (3, 3))):
        r = eng.cfar(torch.from_numpy(x[None]).cuda(), cfar_params("os", tr, gr, rho=0.7, alpha=2.0), want_noise=True)
        _, noise_o, _ = orc.cfar_2d(x.astype(np.float64), "os", tr, gr, rho=0.7, alpha=2.0)
        assert np.array_equal(r["noise"][0].cpu().numpy(), noise_o)


@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_cfar1d_family_and_detectors(torch_cuda, name, seed):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import cfar_params
    cm = load_cfg(name)
    g = golden(name, seed)
    iq, cube, rd_ref = _oracle_frame(cm, seed)
    eng = _engine(cm)
    dev = torch.from_numpy(iq[None]).cuda()
    _, mag = eng.range_doppler(dev, want_rd=False, mag_antenna=0)
    prof = eng.range_profile(dev, 0)
    mag_h = mag[0].cpu().numpy().astype(np.float64)
    prof_h = prof[0].cpu().numpy().astype(np.float64)
    for tag, (kind, params) in CFAR1D_SETS.items():
        p = cfar_params(kind, params["num_train"], params["num_guard"], pfa=params.get("pfa", 0.0),
                        rho=params.get("rho", 0.0), alpha=params.get("alpha", 0.0))
        # all rows of the RD map at once + the profile as a 1-row map
        res = eng.cfar(mag, p, one_d=True, want_noise=True, want_mask=True)
        noise = res["noise"][0].cpu().numpy()
        mask = res["mask"][0].cpu().numpy().astype(bool)
        for r in range(mag_h.shape[0]):
            _, n_o, m_o = orc.cfar_1d(mag_h[r], kind, **params)
            assert np.allclose(noise[r], n_o, rtol=1e-12, atol=0)
            assert np.array_equal(mask[r], m_o), (tag, r)
        resp = eng.cfar(prof[:, None, :].contiguous(), p, one_d=True, want_noise=True, want_mask=True)
        _, n_o, m_o = orc.cfar_1d(prof_h, kind, **params)
        assert np.allclose(resp["noise"][0, 0].cpu().numpy(), n_o, rtol=1e-12, atol=0)
        assert np.array_equal(resp["mask"][0, 0].cpu().numpy().astype(bool), m_o)
        # against the reference run (fp64 inputs): same indices unless a cell sits on the threshold
        gm = np.zeros_like(m_o)
        gm[g[f"c1_prof_{tag}_idx"]] = True
        _mask_diff_report(g["range_profile"], g[f"c1_prof_{tag}_thr"], m_o, gm)


# ------------------------------------------------------------------------------------------------ RA / DBS / SAR
@pytest.mark.parametrize("name,seed", GOLDEN_CASES)
def test_range_angle(torch_cuda, name, seed):
    torch = torch_cuda
    cm = load_cfg(name)
    g = golden(name, seed)
    iq, cube, _ = _oracle_frame(cm, seed)
    eng = _engine(cm)
    dev = torch.from_numpy(iq[None]).cuda()
    sets = ant_sets(cube.shape[0])
    for a in (64, 256):
        assert rel_l2(eng.range_angle(dev, 0, sets["ra"], a)[0].cpu().numpy(), g[f"ra_{a}_sub"]) < RD_TOL
        assert rel_l2(eng.range_angle(dev, 2, None, a)[0].cpu().numpy(), g[f"ra_{a}_all_c2"]) < RD_TOL
    assert rel_l2(eng.range_angle(dev, 1, sets["ra"], 64, windowing=False)[0].cpu().numpy(), g["ra_64_nowin"]) < RD_TOL


@pytest.mark.parametrize("name", ["RadarHD", "RadCloud"])
def test_range_angle_1024_bins(torch_cuda, name):
    """BASELINE.json configs[2] (RadarHD / RadCloud high-resolution heat maps, 2 MB fp32 per RadarHD frame): 1024 angle bins
    against the reference's RangeAngleProcessor goldens - stored rows and the row sums of all rows to 1e-4, argmax of every clear stored row."""
    torch = torch_cuda
    cm = load_cfg(name)
    g = np.load(os.path.join(GOLDEN_DIR, f"ra1024_{name}.npz"))
    eng = _engine(cm)
    seeds = (0, 1) if name == "RadarHD" else (0,)
    iq = torch.from_numpy(np.stack([frame_iq(cm, s) for s in seeds])).cuda()
    for tag, chirp, ants in (("all", 0, None), ("sub_c1", 1, [0, 1, 3])):
        ra = eng.range_angle(iq, chirp, ants, 1024).cpu().numpy().astype(np.float64)
        for i, seed in enumerate(seeds):
            key = f"s{seed}_{tag}"
            assert tuple(g[key + "_shape"]) == ra[i].shape
            assert rel_l2(ra[i][::8], g[key + "_rows"]) < RD_TOL
            assert rel_l2(ra[i].sum(axis=1), g[key + "_rowsum"]) < RD_TOL           # pins the rows that are not stored
            rows = g[key + "_rows"].astype(np.float64)
            top2 = np.partition(rows, -2, axis=1)[:, -2:]
            clear = (top2[:, 1] - top2[:, 0]) > 1e-4 * top2[:, 1]
            assert np.array_equal(ra[i][::8].argmax(axis=1)[clear], g[key + "_argmax"][::8][clear])


@pytest.mark.parametrize("name", ["RadSAR_1843_official", "6843_RadVel_ods_20Hz"])
def test_dbs(torch_cuda, name):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    cm = load_cfg(name)
    g = golden(name, 0)
    iq, cube, _ = _oracle_frame(cm, 0)
    eng = _engine(cm)
    dev = torch.from_numpy(iq[None]).cuda()
    _, vb = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
    _, coarse = orc.angle_axis(64)
    fine = np.linspace(coarse[0], coarse[-1], 100)
    vbin, abin = orc.dbs_bin_selection(coarse, fine, vb, g["dbs_velocity"])
    out = eng.dbs(dev, vbin.tolist(), abin.tolist(), ant_list=[0, 3, 4, 7], num_angle_bins=64)
    assert rel_l2(out[0].cpu().numpy(), g["dbs_sub"]) < RD_TOL
    out = eng.dbs(dev, vbin.tolist(), abin.tolist(), ant_list=None, num_angle_bins=64)
    assert rel_l2(out[0].cpu().numpy(), g["dbs_all"]) < RD_TOL


@pytest.mark.parametrize("name", ["RadSAR_1843_official", "6843_RadVel_ods_20Hz", "1843_RadVel"])
def test_dbs_fused_and_two_pass_equal_one_kernel(torch_cuda, name, monkeypatch):
    """Raw int16 batches of the shapes rd_reg_kernel serves: by default the sharpening runs inside the register-resident
    kernel (rd_reg_dbs_kernel: phase B accumulates the steered range transforms of a frame's antennas in shared memory);
    RADAR_B200_DBS=2 takes the two-pass path (range-Doppler cube of a block of frames in scratch + dbs_combine_kernel),
    RADAR_B200_DBS=1 the one-kernel dbs_kernel.  Blocks of 16 frames with a ragged last one, antenna subset and all
    antennas, more frames than SMs so CTAs walk several frames."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import RadarEngine
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    cm = load_cfg(name)
    iq = torch.from_numpy(np.tile(synth_batch_for_cfg(cm, range(40)), (5, 1, 1, 1, 1))[:173]).cuda()
    _, vb = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
    _, coarse = orc.angle_axis(64)
    fine = np.linspace(coarse[0], coarse[-1], 100)
    vbin, abin = orc.dbs_bin_selection(coarse, fine, vb, np.array([0.5, 0.1, 0.0]))
    shape = (cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    outs = {}
    for mode, chunk in (("1", "0"), ("2", "16"), ("2", "0"), ("0", "0")):
        monkeypatch.setenv("RADAR_B200_DBS", mode)
        monkeypatch.setenv("RADAR_B200_DBS_CHUNK", chunk)
        eng = RadarEngine(*shape)
        sub = [0, 1, 2, 3] if eng.V < 8 else [0, 3, 4, 7]
        outs[(mode, chunk)] = (eng.dbs(iq, vbin.tolist(), abin.tolist(), ant_list=sub, num_angle_bins=64).cpu().numpy(),
                               eng.dbs(iq, vbin.tolist(), abin.tolist(), ant_list=None, num_angle_bins=64).cpu().numpy(),
                               eng.dbs(iq, vbin.tolist(), abin.tolist(), ant_list=[sub[1]], num_angle_bins=64).cpu().numpy(),
                               # out of order and one antenna named twice: its two positions' steering coefficients add up
                               eng.dbs(iq[:7], vbin.tolist(), abin.tolist(), ant_list=[sub[2], sub[0], sub[2]], num_angle_bins=64).cpu().numpy())
    ref = outs[("1", "0")]
    for key in (("2", "16"), ("2", "0"), ("0", "0")):
        for got, want in zip(outs[key], ref):
            assert got.shape == want.shape
            assert np.isfinite(got).all()
            assert rel_l2(got, want) < 1e-5, (key, rel_l2(got, want))
    np.testing.assert_array_equal(outs[("2", "16")][0], outs[("2", "0")][0])       # the block size does not change a bit
    assert not np.array_equal(outs[("2", "0")][1], ref[1])                         # the two-pass path really ran
    assert not np.array_equal(outs[("0", "0")][1], outs[("2", "0")][1])            # and so did the fused one
    np.testing.assert_array_equal(outs[("0", "0")][1][:40], outs[("0", "0")][1][40:80])   # the tiled frames repeat: so must the maps


def test_sar(torch_cuda):
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.synthetic import synth_frame_iq
    from mmwave_radar_processing_b200.processors._device import engine_for_cube_shape
    name = "RadSAR_1843_official"
    cm = load_cfg(name)
    g = golden(name, 0)
    r, s, cpl, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    prof = cm.profile_cfgs[0]
    eng = engine_for_cube_shape(1, s, 1)
    for tag in ("a", "b"):
        rx, cfg_idx, nf, stride = [int(v) for v in g[f"sar_{tag}_params"]]
        mask, _ = orc.sar_chirp_times_us(cpl, l, prof["idleTime_us"] + prof["rampEndTime_us"], cfg_idx, stride)
        hist = np.stack([orc.iq_to_cube(synth_frame_iq(r, s, cpl, l, 1000 + f))[rx][:, mask] for f in range(nf)])
        geo = g[f"sar_{tag}_geometry"]
        d = orc.sar_steering(g[f"sar_{tag}_az"], g[f"sar_{tag}_el"])
        ap = torch.from_numpy(np.ascontiguousarray(hist.transpose(1, 0, 2).reshape(s, -1), dtype=np.complex64)).cuda()
        pos = torch.from_numpy(np.ascontiguousarray(geo.transpose(1, 0, 2).reshape(3, -1))).cuda()
        dirs = torch.from_numpy(np.ascontiguousarray(d.reshape(3, -1))).cuda()
        ref = g[f"sar_{tag}_resp"].reshape(s, -1)
        out = eng.sar_image(ap, pos, dirs, float(g[f"sar_{tag}_lambda"]))
        assert rel_l2(out.cpu().numpy(), ref) < RD_TOL
        # voxel sharding: any slice equals the same columns of the full image
        gtot = ref.shape[1]
        lo, hi = gtot // 3, gtot - 5
        part = eng.sar_image(ap, pos, dirs, float(g[f"sar_{tag}_lambda"]), lo, hi)
        assert np.array_equal(part.cpu().numpy(), out[:, lo:hi].cpu().numpy())


# ------------------------------------------------------------------------------------------------ batch properties
def test_batch_properties_full_size(torch_cuda):
    """Size-independent properties at bench batch size: every frame of a batch equals the same frame processed
    alone; linearity of the RD transform; permuting frames permutes results."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    cm = load_cfg("1843_RadVel")
    base = synth_batch_for_cfg(cm, range(16))
    reps = 64
    iq = np.concatenate([base] * reps)                       # 1024 frames
    perm = np.random.default_rng(0).permutation(iq.shape[0])
    pipe = FramePipeline(cm, detector=DetectorSpec("os", (8, 4), (2, 1), rho=0.75, alpha=3.5), max_dets=256)
    dev = torch.from_numpy(iq[perm]).cuda()
    out = pipe.run(dev)
    counts = out["counts"].cpu().numpy()
    offs = out["offsets"].cpu().numpy()
    assert offs[0] == 0 and np.array_equal(np.diff(offs), np.minimum(counts, 256))
    clouds = pipe.gather_host(out)
    single = FramePipeline(cm, detector=DetectorSpec("os", (8, 4), (2, 1), rho=0.75, alpha=3.5), max_dets=256)
    ref = single.run(torch.from_numpy(base).cuda())
    rc, rp = ref["counts"].cpu().numpy(), single.gather_host(ref)
    assert counts.min() > 0
    for i, src in enumerate(perm):
        j = src % 16
        assert counts[i] == rc[j]
        assert np.array_equal(clouds[i], rp[j])
    # end-to-end host path (chunks ping-ponging over two streams) returns the same clouds
    iq_h, cnt_h, pts_h = pipe.host_buffers(iq.shape[0], points_per_frame=int(counts.max()))
    iq_h.copy_(torch.from_numpy(iq[perm]))
    totals, regions, h2d, d2h = pipe.run_host(iq_h, cnt_h, pts_h, chunk_frames=96)   # ragged last chunk
    assert h2d == iq.nbytes and sum(totals) == offs[-1]
    host_clouds = FramePipeline.split_host(pts_h, cnt_h, totals, regions, 96, 256)
    assert len(host_clouds) == len(clouds)
    for a, b in zip(host_clouds, clouds):
        assert np.array_equal(a, b)
    # linearity: RD(a*x) == a*RD(x) up to fp32 rounding, RD(x+y) == RD(x)+RD(y)
    eng = pipe.engine
    x = torch.from_numpy(base[:2]).cuda()
    half = torch.from_numpy((base[:2] // 2 * 2)).cuda()
    rd_full, _ = eng.range_doppler(half)
    rd_half, _ = eng.range_doppler(torch.from_numpy(base[:2] // 2).cuda())
    assert rel_l2((2 * rd_half).cpu().numpy(), rd_full.cpu().numpy()) < 1e-6
    s = torch.from_numpy((base[0:1].astype(np.int32) // 2 + base[1:2].astype(np.int32) // 2).astype(np.int16)).cuda()
    a, _ = eng.range_doppler(torch.from_numpy(base[0:1] // 2).cuda())
    b, _ = eng.range_doppler(torch.from_numpy(base[1:2] // 2).cuda())
    c, _ = eng.range_doppler(s)
    assert rel_l2((a + b).cpu().numpy(), c.cpu().numpy()) < 1e-6
    del x


def test_os_cfar_counting_path_ties_and_edges(torch_cuda):
    """The counting OS-CFAR decision equals the order-statistic decision on maps full of ties, zeros, huge
    dynamic range and thresholds that land exactly on cell values (alpha a power of two), 1-D and 2-D."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import cfar_params, utility_engine
    eng = utility_engine()
    rng = np.random.default_rng(7)
    maps = []
    maps.append(rng.integers(0, 6, (40, 33)).astype(np.float32))                       # heavy ties
    maps.append((2.0 ** rng.integers(-20, 20, (40, 33))).astype(np.float32))            # powers of two
    maps.append(rng.exponential(1.0, (40, 33)).astype(np.float32))
    m = rng.exponential(1.0, (40, 33)).astype(np.float32); m[::3] = 0.0; maps.append(m)  # zero rows
    m = rng.exponential(1e30, (40, 33)).astype(np.float32); m[5, 5] = np.float32(3e38); maps.append(m)
    x = torch.from_numpy(np.stack(maps)).cuda()
    for alpha in (0.5, 1.0, 2.0, 3.5, 1e-3):
        for rho in (0.01, 0.5, 0.75, 1.0):
            for (t2, g2) in (((3, 2), (1, 1)), ((8, 4), (2, 1)), ((1, 1), (0, 0))):
                p = cfar_params("os", t2, g2, rho=rho, alpha=alpha)
                a = eng.cfar(x, p, want_noise=True, want_mask=True)
                b = eng.cfar(x, p, want_noise=False, want_mask=True)
                assert torch.equal(a["mask"], b["mask"]) and torch.equal(a["counts"], b["counts"])
                for i, mp in enumerate(maps):
                    _, _, mo = orc.cfar_2d(mp.astype(np.float64), "os", t2, g2, rho=rho, alpha=alpha)
                    assert np.array_equal(b["mask"][i].cpu().numpy().astype(bool), mo), (alpha, rho, t2, i)
            p1 = cfar_params("os", 6, 2, rho=rho, alpha=alpha)
            a = eng.cfar(x, p1, one_d=True, want_noise=True, want_mask=True)
            b = eng.cfar(x, p1, one_d=True, want_noise=False, want_mask=True)
            assert torch.equal(a["mask"], b["mask"])
            _, _, mo = orc.cfar_1d(maps[0][3].astype(np.float64), "os", 6, 2, rho=rho, alpha=alpha)
            assert np.array_equal(b["mask"][0, 3].cpu().numpy().astype(bool), mo)


def test_os_cfar_ladder_equals_counting_path(torch_cuda):
    """The level-ladder screen of the mask-only OS-CFAR (kernels_cfar.cu: os_ladder_cells) gives the mask of the plain
    counting path (RADAR_B200_CFAR_LADDER=0) and of the oracle: range-Doppler maps of both bench cfgs, pure noise, signed
    data (negative median), constant maps, a NaN / inf cell, detector settings of verify_processors.py and the GUI yaml."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import cfar_params
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    rng = np.random.default_rng(11)
    for name, geom in (("6843_RadVel_ods_20Hz", "ods"), ("1843_RadVel", "standard")):
        cm = load_cfg(name, geom)
        lad = _engine_with_env(cm, "RADAR_B200_CFAR_LADDER", 1)
        ref = _engine_with_env(cm, "RADAR_B200_CFAR_LADDER", 0)
        iq = torch.from_numpy(synth_batch_for_cfg(cm, range(12))).cuda()
        _, mag = lad.range_doppler(iq, want_rd=False, mag_antenna=0)
        s, l = mag.shape[-2:]
        extra = [np.abs(rng.normal(size=(s, l)) + 1j * rng.normal(size=(s, l))) * 1e3,        # Rayleigh noise only
                 rng.normal(size=(s, l)) - 0.5,                                               # signed, negative median
                 np.full((s, l), 3.0), np.zeros((s, l)),                                      # constant maps
                 rng.exponential(1.0, (s, l)) * np.exp(rng.normal(size=(s, l)) * 3.0)]        # heavy tailed
        m = rng.exponential(1.0, (s, l)); m[s // 2, l // 2] = np.nan; m[s // 3, l // 3] = np.inf; extra.append(m)
        maps = torch.cat([mag, torch.from_numpy(np.stack(extra).astype(np.float32)).cuda()])
        for (t2, g2, rho, alpha) in (((8, 4), (2, 1), 0.75, 3.5), ((5, 5), (3, 2), 0.7, 2.0), ((4, 4), (2, 2), 0.5, 1.0),
                                     ((2, 10), (1, 3), 0.9, 6.0), ((8, 4), (2, 1), 1.0, 0.25)):
            p = cfar_params("os", t2, g2, rho=rho, alpha=alpha)
            a = lad.cfar(maps, p, want_mask=True)
            b = ref.cfar(maps, p, want_mask=True)
            assert torch.equal(a["mask"], b["mask"]) and torch.equal(a["counts"], b["counts"]), (name, t2, g2, rho, alpha)
            for i in range(maps.shape[0]):
                n = int(a["counts"][i])
                assert torch.equal(a["dets"][i, :n], b["dets"][i, :n])
            for i in (0, 5, 12, 13, 16, 17):
                _, _, mo = orc.cfar_2d(maps[i].cpu().numpy().astype(np.float64), "os", t2, g2, rho=rho, alpha=alpha)
                assert np.array_equal(a["mask"][i].cpu().numpy().astype(bool), mo), (name, t2, i)


def test_cfar_tiled_large_maps(torch_cuda):
    """Maps beyond one CTA's shared memory run as tiles with halo + a separate ordered compaction: forced small tiles
    (RADAR_B200_CFAR_TILE) reproduce the single-CTA results bit for bit for every detector family, and genuinely large
    maps (300 x 200 2-D, a 50 000-cell 1-D signal) match the oracle."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.engine import cfar_params
    cm = load_cfg("1843_RadVel")
    plain = _engine_with_env(cm, "RADAR_B200_CFAR_TILE", 0)
    rng = np.random.default_rng(5)
    maps = torch.from_numpy(rng.exponential(1.0, (5, 61, 47)).astype(np.float32) *
                            np.exp(rng.normal(size=(5, 61, 47)) * 1.5).astype(np.float32)).cuda()
    sel = torch.from_numpy((rng.random((5, 61)) < 0.6).astype(np.uint8)).cuda()
    for tile in (7, 16, 23):
        tiled = _engine_with_env(cm, "RADAR_B200_CFAR_TILE", tile)
        for p, kw in ((cfar_params("os", (8, 4), (2, 1), rho=0.75, alpha=2.0), {}),
                      (cfar_params("os", (3, 2), (1, 1), rho=0.5, alpha=1.5), dict(want_noise=True)),
                      (cfar_params("ca", (4, 4), (2, 2), pfa=1e-2), dict(want_noise=True)),
                      (cfar_params("os", 6, 2, rho=0.75, alpha=2.0), dict(one_d=True, row_select=sel)),
                      (cfar_params("go", 5, 1, pfa=1e-2), dict(one_d=True, want_noise=True)),
                      (cfar_params("so", 5, 1, pfa=1e-1), dict(one_d=True, row_select=sel, want_noise=True))):
            a = tiled.cfar(maps, p, want_mask=True, **kw)
            b = plain.cfar(maps, p, want_mask=True, **kw)
            assert torch.equal(a["mask"], b["mask"]) and torch.equal(a["counts"], b["counts"]), (tile, kw)
            assert int(b["counts"].sum()) > 0
            if kw.get("want_noise"):
                assert torch.equal(a["noise"], b["noise"])
            for i in range(maps.shape[0]):
                n = int(a["counts"][i])
                assert torch.equal(a["dets"][i, :n], b["dets"][i, :n])
    big = rng.exponential(1.0, (2, 300, 200)).astype(np.float32)
    big[0, 150, 100] = 60.0; big[1, 20:24, 30] = 40.0
    x = torch.from_numpy(big).cuda()
    for kind, kw in (("os", dict(rho=0.75, alpha=3.5)), ("ca", dict(pfa=1e-4))):
        out = plain.cfar(x, cfar_params(kind, (8, 4), (2, 1), **kw), want_mask=True, want_noise=(kind == "ca"))
        for i in range(2):
            thr, noise, mo = orc.cfar_2d(big[i].astype(np.float64), kind, (8, 4), (2, 1), **kw)
            assert np.array_equal(out["mask"][i].cpu().numpy().astype(bool), mo), (kind, i)
            assert np.array_equal(out["dets"][i, :int(out["counts"][i])].cpu().numpy(), orc.mask_to_dets(mo))
            if kind == "ca":
                assert np.allclose(out["noise"][i].cpu().numpy(), noise, rtol=1e-12, atol=0)
    sig = rng.exponential(1.0, (1, 1, 50000)).astype(np.float32)
    sig[0, 0, 777] = 80.0
    out = plain.cfar(torch.from_numpy(sig).cuda(), cfar_params("os", 16, 4, rho=0.75, alpha=6.0), one_d=True, want_mask=True)
    _, _, mo = orc.cfar_1d(sig[0, 0].astype(np.float64), "os", 16, 4, rho=0.75, alpha=6.0)
    assert np.array_equal(out["mask"][0, 0].cpu().numpy().astype(bool), mo) and mo[777]


def test_error_paths(torch_cuda):
    torch = torch_cuda
    from mmwave_radar_processing_b200 import _native as nat
    from mmwave_radar_processing_b200.engine import RadarEngine, cfar_params, utility_engine
    eng = utility_engine()
    x = torch.zeros((1, 8, 8), dtype=torch.float32, device="cuda")
    with pytest.raises(nat.RadarError):
        eng.cfar(x, cfar_params("go", (1, 1), (1, 1), pfa=1e-3))          # GO/SO are 1-D only
    with pytest.raises(nat.RadarError):
        eng.cfar(x, cfar_params("os", (1, 1), (0, 0), rho=0.5, alpha=2.0, os_rank=99))
    with pytest.raises(nat.RadarError):
        RadarEngine(4, 257 * 3, 1, 8)                                     # prime factor 257 > 128
    e = RadarEngine(4, 64, 1, 8)
    with pytest.raises(ValueError):
        e.range_doppler(torch.zeros((1, 8, 4, 63, 2), dtype=torch.int16, device="cuda"))


# ------------------------------------------------------------------------------------------------ point-cloud kernels
def _engine_with_env(cm, var, value):
    """A private engine created with an environment switch set (switches are read when the context is created)."""
    from mmwave_radar_processing_b200.engine import RadarEngine
    old = os.environ.get(var)
    os.environ[var] = str(value)
    try:
        return RadarEngine(cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    finally:
        if old is None:
            del os.environ[var]
        else:
            os.environ[var] = old


@pytest.mark.parametrize("name", ["6843_RadVel_ods_20Hz", "1843_RadVel"])
def test_point_cloud_kernels_agree_and_match_oracle(torch_cuda, name):
    """Thread-per-detection kernel (default) and warp-per-detection kernel (RADAR_B200_PC=1): identical angle bins and
    points, both equal to the oracle's fp64 FFT argmax, over antenna sets of 1..8 (padded / exact powers of two), 12
    (warp kernel only), 8..256 bins, both shift flags, ragged per-frame counts."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200.pipeline import reference_axes
    cm = load_cfg(name)
    frames = [_oracle_frame(cm, 40 + f) for f in range(3)]
    iq = torch.from_numpy(np.stack([f[0] for f in frames])).cuda()
    v, s, l = frames[0][2].shape
    rng = np.random.default_rng(17)
    max_dets = 200
    counts = np.array([200, 0, 77], dtype=np.int32)
    dets = np.zeros((3, max_dets, 2), dtype=np.int32)
    dets[..., 0] = rng.integers(0, s, (3, max_dets))
    dets[..., 1] = rng.integers(0, l, (3, max_dets))
    engines = {"thread": _engine(cm), "warp": _engine_with_env(cm, "RADAR_B200_PC", 1)}
    cases = [([0], [], 64), ([1, 0], [3], 64), ([0, 1, 2], [3, 2, 1, 0], 128), ([0, 3, 1, 2], [2, 3], 8),
             (list(range(min(v, 8))), [1, 2, 3, 0, 2], 256), ([3, 2, 1, 0, 1, 2], list(range(min(v, 7))), 32)]
    if v >= 12:
        cases.append((list(range(12)), [9, 8, 5, 4], 64))            # 12 antennas: warp kernel on both engines
    for az, el, bins in cases:
        rb, vb, ab = reference_axes(cm, bins)
        got = {}
        for tag, eng in engines.items():
            eng.set_axes(rb, vb)
            eng.set_angle_bins(ab)
            rd, _ = eng.range_doppler(iq, want_rd=True, mag_antenna=0)
            rd_host = rd.cpu().numpy().astype(np.complex128)             # the oracle runs on the SAME fp32 cube
            for sh_az, sh_el in ((True, False), (False, True)):
                pts, az_bin, el_bin = eng.point_cloud(rd, torch.from_numpy(dets).cuda(), torch.from_numpy(counts).cuda(),
                                                      az, el, sh_az, sh_el, bins)
                got[tag, sh_az] = (pts.cpu().numpy(), az_bin.cpu().numpy(), el_bin.cpu().numpy())
        for sh_az, sh_el in ((True, False), (False, True)):
            a, b = got["thread", sh_az], got["warp", sh_az]
            for x, y in zip(a, b):
                np.testing.assert_array_equal(x, y)
            for f in range(3):
                n = counts[f]
                rd_ref = rd_host[f]
                np.testing.assert_array_equal(a[1][f, :n], orc.angle_indices(rd_ref, dets[f, :n, 0], dets[f, :n, 1], az, bins, sh_az))
                np.testing.assert_array_equal(a[2][f, :n], orc.angle_indices(rd_ref, dets[f, :n, 0], dets[f, :n, 1], el, bins, sh_el))


def test_stream_sized_batch_matches_small_batches(torch_cuda):
    """BASELINE.json configs[4] runs 100k-frame blocks through ONE launch per stage: element offsets pass 2^31 (60 000 frames
    x 12 antennas x 63 x 70 = 3.2e9 complex cells).  Frames at the far end of such a batch must equal the same frames run as a
    small batch, bit for bit (same kernels, same per-frame arithmetic)."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    if torch.cuda.mem_get_info()[0] < 60e9:
        pytest.skip("needs 60 GB of free device memory")
    cm = load_cfg("6843_RadVel_ods_20Hz", "ods")
    uniq = torch.from_numpy(synth_batch_for_cfg(cm, range(32))).cuda()
    f_big = 60000
    iq = uniq.repeat((f_big // 32, 1, 1, 1, 1))
    spec = DetectorSpec("os", (8, 4), (2, 1), rho=0.75, alpha=3.5)
    kw = dict(detector=spec, az_antenna_idxs=(0, 3, 4, 7), el_antenna_idxs=(9, 8, 5, 4), max_dets=256,
              with_range_angle=True, ra_antennas=(0, 3, 4, 7))
    small = FramePipeline(cm, **kw).run(uniq)
    ref = {k: small[k].clone() for k in ("counts", "offsets", "rd", "mag", "ra", "dets")}
    ref_pts = small["points"][:int(small["offsets"][32])].clone()
    big = FramePipeline(cm, **kw).run(iq)
    torch.cuda.synchronize()
    for lo in (0, f_big // 2, f_big - 32):          # f_big and f_big/2 are multiples of 32: the same 32 frames
        lo -= lo % 32
        assert torch.equal(big["counts"][lo:lo + 32], ref["counts"])
        assert torch.equal(big["rd"][lo:lo + 32], ref["rd"])
        assert torch.equal(big["mag"][lo:lo + 32], ref["mag"])
        assert torch.equal(big["ra"][lo:lo + 32], ref["ra"])
        assert torch.equal(big["dets"][lo:lo + 32], ref["dets"])
        o0, o1 = int(big["offsets"][lo]), int(big["offsets"][lo + 32])
        assert o1 - o0 == ref_pts.shape[0] and torch.equal(big["points"][o0:o1], ref_pts)
    assert int(big["offsets"][f_big]) == (f_big // 32) * ref_pts.shape[0]


def test_frame_file_to_point_clouds(torch_cuda, tmp_path):
    """int16 frame file -> memmap -> pinned chunks -> FramePipeline.run_host gives the clouds of the frames
    run directly from device memory (and of the oracle for the first frames)."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    from mmwave_radar_processing_b200 import ingest
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    cm = load_cfg("6843_RadVel_ods_20Hz", "ods")
    iq = synth_batch_for_cfg(cm, range(37))
    path = str(tmp_path / "stream.rb2")
    ingest.FrameFile.write(path, iq)
    pipe = FramePipeline(cm, detector=DetectorSpec(kind="os", num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5),
                         az_antenna_idxs=(0, 3, 4, 7), el_antenna_idxs=(9, 8, 5, 4), max_dets=256)
    clouds = ingest.process_file(pipe, path, batch_frames=16, chunk_frames=8, points_per_frame=256)
    direct = pipe.gather_host(pipe.run(torch.from_numpy(iq).cuda()))
    assert len(clouds) == len(direct) == 37
    for a, b in zip(clouds, direct):
        assert a.shape == b.shape and np.array_equal(a, b)
    for i in range(3):
        cube = orc.virtual_array(orc.iq_to_cube(iq[i]), cm.num_rx_antennas, cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
        rd = orc.range_doppler(cube)
        _, _, mask = orc.cfar_2d(np.abs(rd[0]), "os", (8, 4), (2, 1), rho=0.75, alpha=3.5)
        pts, _, _ = orc.point_cloud(rd, orc.mask_to_dets(mask), pipe.range_bins, pipe.vel_bins, [0, 3, 4, 7], [9, 8, 5, 4])
        assert clouds[i].shape == pts.shape and np.allclose(clouds[i], pts, rtol=1e-9, atol=1e-9)
    wrong = load_cfg("1843_RadVel")
    pipe2 = FramePipeline(wrong, detector=DetectorSpec(kind="os", num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5),
                          az_antenna_idxs=(0, 1, 2, 3), el_antenna_idxs=(), max_dets=64)
    with pytest.raises(ValueError):
        ingest.process_file(pipe2, path)


def test_nvtx_ranges_do_not_change_results(torch_cuda):
    """RADAR_B200_NVTX=1 (read at import) wraps the pipeline stages in NVTX ranges: a fresh interpreter runs a small batch
    with the ranges on and prints the point total, which must equal the in-process run."""
    import subprocess
    import sys
    torch = torch_cuda
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    cm = load_cfg("6843_RadVel_ods_20Hz", "ods")
    pipe = FramePipeline(cm, detector=DetectorSpec(kind="os", num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5),
                         az_antenna_idxs=(0, 3, 4, 7), el_antenna_idxs=(9, 8, 5, 4), max_dets=256)
    out = pipe.run(torch.from_numpy(synth_batch_for_cfg(cm, range(6))).cuda())
    want = int(out["offsets"][6])
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import torch\n"
        "from conftest import load_cfg\n"
        "from mmwave_radar_processing_b200 import pipeline\n"
        "from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg\n"
        "assert pipeline._NVTX\n"
        "cm = load_cfg('6843_RadVel_ods_20Hz', 'ods')\n"
        "p = pipeline.FramePipeline(cm, detector=pipeline.DetectorSpec(kind='os', num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5),\n"
        "                           az_antenna_idxs=(0, 3, 4, 7), el_antenna_idxs=(9, 8, 5, 4), max_dets=256, with_range_angle=True)\n"
        "marks = []\n"
        "o = p.run(torch.from_numpy(synth_batch_for_cfg(cm, range(6))).cuda(), marks.append)\n"
        "print('TOTAL', int(o['offsets'][6]), ','.join(marks))\n"
    ) % (REPO, os.path.join(REPO, "tests"))
    env = dict(os.environ, RADAR_B200_NVTX="1")
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=240)
    assert res.returncode == 0, res.stderr[-2000:]
    line = [l for l in res.stdout.splitlines() if l.startswith("TOTAL")][0].split()
    assert int(line[1]) == want and line[2] == "range_doppler,range_angle,cfar,point_cloud"


@pytest.mark.parametrize("name,geom", [("6843_RadVel_ods_20Hz", "ods"), ("1843_RadVel", "standard"), ("RadCloud", "standard"),
                                       ("RadSAR_1843_official", "standard")])
def test_register_resident_kernel_options_against_previous_kernels(torch_cuda, name, geom):
    """Every option of radar_range_doppler through rd_reg_kernel (all four instantiated shapes: whole / split Doppler and
    range columns, twiddle mode) against the previous kernels (RADAR_B200_RD_REG=0) and the oracle: window flags, magnitude
    of one / every antenna, magnitude only, batches of 1 / 3 / 301 frames (fewer and more items than CTAs), and a cube that
    is not 16-byte aligned (the bulk copies need the alignment: the call must fall back, not fail)."""
    torch = torch_cuda
    from oracle import radar_oracle as orc
    cm = load_cfg(name, geom)
    new = _engine_with_env(cm, "RADAR_B200_RD_REG", 1)
    old = _engine_with_env(cm, "RADAR_B200_RD_REG", 0)
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    iq_all = synth_batch_for_cfg(cm, range(7))
    big = torch.from_numpy(np.tile(iq_all, (43, 1, 1, 1, 1))).cuda()             # 301 frames
    for f in (1, 3, 301):
        x = big[:f]
        for flags in (3, 0, 1, 2):
            a, am = new.range_doppler(x, window_flags=flags, mag_antenna=0)
            assert new.last_rd_kernel() == "rd_reg_kernel"
            b, bm = old.range_doppler(x, window_flags=flags, mag_antenna=0)
            assert old.last_rd_kernel() != "rd_reg_kernel"
            assert rel_l2(a.cpu().numpy(), b.cpu().numpy()) < 1e-6 and rel_l2(am.cpu().numpy(), bm.cpu().numpy()) < 1e-6
            if f > 1 and flags != 3:
                continue
            _, all_mag = new.range_doppler(x, window_flags=flags, want_rd=False, mag_antenna=-1)
            assert rel_l2(all_mag.cpu().numpy(), a.abs().cpu().numpy()) < 1e-6
            last = new.V - 1
            _, one = new.range_doppler(x, window_flags=flags, want_rd=False, mag_antenna=last)
            assert torch.equal(one, all_mag[:, last])
    # frames repeat with period 7: identical inputs give identical outputs wherever they sit in the batch
    a, _ = new.range_doppler(big, mag_antenna=0)
    assert torch.equal(a[:7], a[294:301]) and torch.equal(a[7:14], a[140:147])
    cube = orc.virtual_array(orc.iq_to_cube(iq_all[2]), cm.num_rx_antennas, cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    assert rel_l2(a[2].cpu().numpy(), orc.range_doppler(cube)) < RD_TOL
    # misaligned cube (8 bytes off a 16-byte boundary)
    flat = torch.empty(big[:3].numel() + 8, dtype=torch.int16, device="cuda")
    off = flat[4:4 + big[:3].numel()].view(big[:3].shape)
    off.copy_(big[:3])
    assert off.data_ptr() % 16 == 8
    c, _ = new.range_doppler(off, mag_antenna=0)
    assert new.last_rd_kernel() != "rd_reg_kernel"
    assert rel_l2(c.cpu().numpy(), a[:3].cpu().numpy()) < 1e-6
