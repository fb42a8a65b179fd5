"""GPU parity of the tensor-core SAR beamformer (csrc/kernels_sar_tc.cu: tcgen05 TF32 GEMM with hi/lo operand splitting)
against a direct fp64 evaluation of the reference formula (compute_synthetic_response,
processors/simple_synthetic_array_beamformer_processor_multiFrame.py:543-585) and against the FP32-pipe kernel, over the
shapes that take the tensor-core path, the fallback shapes, both tile configurations and voxel slices."""
import os

import numpy as np
import pytest

from conftest import rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-4          # relative L2 (BASELINE.json); measured ~3e-7 for the split-TF32 path


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def sar_case(s, n, g, seed):
    rng = np.random.default_rng(seed)
    lam = 0.0039
    x = (rng.integers(-2000, 2000, (s, n)) + 1j * rng.integers(-2000, 2000, (s, n))).astype(np.complex64)
    t = np.arange(n) * 1e-4
    pos = np.stack([0.01 * t + 1e-5 * rng.standard_normal(n), 0.5 * t, 1e-5 * rng.standard_normal(n)])     # metres: 1e2..1e4 turns
    az = rng.uniform(-1.0, 1.0, g)
    el = rng.uniform(-0.3, 0.3, g)
    dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
    return x, pos, dirs, lam


def sar_direct(x, pos, dirs, lam, cols):
    """fp64: out[:, g] = FFT_S(hanning(S) * sum_n hamming(N)[n] x[s, n] exp(+2 pi i d_g.p_n / lambda)) for the columns asked."""
    s, n = x.shape
    w = np.hamming(n)[:, None] * np.exp(2j * np.pi * (pos.T @ dirs[:, cols]) / lam)
    return np.fft.fft((x.astype(np.complex128) @ w) * np.hanning(s)[:, None], axis=0)


def engine_with_mode(mode, s, dbg=None):
    from mmwave_radar_processing_b200.engine import RadarEngine
    env = {"RADAR_B200_SAR": str(mode)}
    if dbg is not None:
        env["RADAR_B200_SAR_DBG"] = str(dbg)
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        return RadarEngine(1, s, 1, 1)
    finally:
        for k, v in old.items():
            if v is None:
                del os.environ[k]
            else:
                os.environ[k] = v


@pytest.mark.parametrize("s,n,g", [(90, 500, 300), (63, 210, 129), (64, 37, 128), (100, 333, 200), (128, 100, 64), (120, 65, 257)])
def test_tensor_core_path_matches_fp64_and_fp32_kernel(torch_cuda, s, n, g):
    torch = torch_cuda
    x, pos, dirs, lam = sar_case(s, n, g, seed=s + n)
    want = sar_direct(x, pos, dirs, lam, np.arange(g))
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    outs = {}
    for mode in (1, 2, 3):              # FP32-pipe kernel, tcgen05 one tile per CTA, tcgen05 two tiles per CTA
        eng = engine_with_mode(mode, s)
        outs[mode] = eng.sar_image(xs, ps, ds, lam).cpu().numpy()
        err = rel_l2(outs[mode], want)
        assert err < TOL, (mode, err)
    assert rel_l2(outs[2], outs[1]) < 1e-5
    np.testing.assert_array_equal(outs[2], outs[3])                 # tiling must not change a single bit
    eng = engine_with_mode(2, s)
    lo, hi = g // 3, g - 5
    part = eng.sar_image(xs, ps, ds, lam, lo, hi).cpu().numpy()     # voxel slice (multi-GPU sharding unit)
    np.testing.assert_array_equal(part, outs[2][:, lo:hi])


@pytest.mark.parametrize("s,n,g", [(200, 64, 40), (130, 100, 150), (512, 300, 300)])
def test_more_than_128_range_samples_stay_on_the_tensor_cores(torch_cuda, s, n, g):
    """2 S > 256 does not fit one MMA: the samples are cut into column tiles of 96 (one pass of prep + GEMM each, one range
    FFT over all S).  Against fp64 and the FP32-pipe kernel; a voxel slice and a forced aperture split stay bit-identical."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.engine import RadarEngine
    x, pos, dirs, lam = sar_case(s, n, g, seed=s + n)
    want = sar_direct(x, pos, dirs, lam, np.arange(g))
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    tc = engine_with_mode(0, s).sar_image(xs, ps, ds, lam).cpu().numpy()
    fp32 = engine_with_mode(1, s).sar_image(xs, ps, ds, lam).cpu().numpy()
    assert rel_l2(tc, want) < TOL, rel_l2(tc, want)
    assert rel_l2(fp32, want) < TOL
    assert rel_l2(tc, fp32) < 1e-5
    assert not np.array_equal(tc, fp32)                      # the tensor-core path really ran
    lo, hi = g // 3, g - 5
    part = engine_with_mode(0, s).sar_image(xs, ps, ds, lam, lo, hi).cpu().numpy()
    np.testing.assert_array_equal(part, tc[:, lo:hi])
    os.environ["RADAR_B200_SAR_KSPLIT"] = "2"
    try:
        eng = RadarEngine(1, s, 1, 1)
    finally:
        del os.environ["RADAR_B200_SAR_KSPLIT"]
    np.testing.assert_array_equal(eng.sar_image(xs, ps, ds, lam).cpu().numpy(), tc)


def test_throughput_grid(torch_cuda):
    """The throughput grid of SURVEY.md section 8d (S = 90, N = 2000, G = 32768) is checked on a sample of directions."""
    torch = torch_cuda
    x, pos, dirs, lam = sar_case(90, 2000, 32768, seed=6)
    eng = engine_with_mode(0, 90)
    out = eng.sar_image(torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda(), lam)
    cols = np.random.default_rng(0).choice(32768, 96, replace=False)
    cols[:4] = [0, 127, 128, 32767]
    assert rel_l2(out[:, torch.from_numpy(cols).cuda()].cpu().numpy(), sar_direct(x, pos, dirs, lam, cols)) < TOL
    assert torch.isfinite(torch.view_as_real(out)).all()


@pytest.mark.parametrize("s,frames,per_frame,g", [(90, 5, 100, 300), (64, 7, 37, 200), (100, 3, 160, 129), (72, 4, 16, 140)])
def test_equispaced_chunks_step_the_phase(torch_cuda, s, frames, per_frame, g):
    """Constant velocity inside a frame (the reference's geometry): chunks of 16 equispaced elements take the stepping path
    (t0 + e delta), chunks straddling a frame boundary the general one.  Both against fp64 and against the general path
    alone (RADAR_B200_SAR_DBG=32), both tilings."""
    torch = torch_cuda
    rng = np.random.default_rng(s + per_frame)
    n = frames * per_frame
    lam = 0.0039
    x = (rng.integers(-2000, 2000, (s, n)) + 1j * rng.integers(-2000, 2000, (s, n))).astype(np.complex64)
    pos = np.zeros((3, n))
    start = np.zeros(3)
    for f in range(frames):
        vel = np.array([0.01, 0.5, 0.0]) + 0.05 * rng.standard_normal(3)
        t = np.arange(per_frame) * 1.1e-4
        pos[:, f * per_frame:(f + 1) * per_frame] = start[:, None] + vel[:, None] * t[None, :]
        start = pos[:, (f + 1) * per_frame - 1] + vel * 3e-3                     # inter-frame gap
    az = rng.uniform(-1.0, 1.0, g)
    el = rng.uniform(-0.3, 0.3, g)
    dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
    want = sar_direct(x, pos, dirs, lam, np.arange(g))
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    step2 = engine_with_mode(2, s).sar_image(xs, ps, ds, lam).cpu().numpy()
    step3 = engine_with_mode(3, s).sar_image(xs, ps, ds, lam).cpu().numpy()
    general = engine_with_mode(2, s, dbg=32).sar_image(xs, ps, ds, lam).cpu().numpy()
    assert rel_l2(step2, want) < 1e-5, rel_l2(step2, want)
    assert rel_l2(general, want) < 1e-5
    assert rel_l2(step2, general) < 5e-6, rel_l2(step2, general)
    np.testing.assert_array_equal(step2, step3)
    assert not np.array_equal(step2, general)                # the stepping path really ran


def test_split_k_does_not_change_a_bit(torch_cuda):
    """The aperture cut over CTA rows (split-K, chosen from the grid size so that voxel slices of a multi-GPU run fill the
    SMs): every accumulation segment is drained into its own array and summed in segment order, so any cut - and the
    in-place accumulation of ksplit = 1 - gives the same image bit for bit; a slice equals the full grid's columns."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.engine import RadarEngine
    s, n, g = 90, 2000, 700
    x, pos, dirs, lam = sar_case(s, n, g, seed=11)
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    outs = {}
    for ks in (1, 2, 3, 7, 0):
        os.environ["RADAR_B200_SAR_KSPLIT"] = str(ks)
        try:
            eng = RadarEngine(1, s, 1, 1)
        finally:
            del os.environ["RADAR_B200_SAR_KSPLIT"]
        outs[ks] = eng.sar_image(xs, ps, ds, lam).cpu().numpy()
        if ks == 0:
            part = eng.sar_image(xs, ps, ds, lam, 100, 333).cpu().numpy()
            np.testing.assert_array_equal(part, outs[0][:, 100:333])
    for ks in (2, 3, 7, 0):
        np.testing.assert_array_equal(outs[ks], outs[1])
    assert rel_l2(outs[1][:, :64], sar_direct(x, pos, dirs, lam, np.arange(64))) < TOL


def test_two_images_in_flight_on_one_gpu(torch_cuda):
    """distributed.LocalImagePipeline: two contexts / streams / CUDA graphs, consecutive images overlap; every image must be
    the plain call's image bit for bit, also after the inputs changed between replays."""
    torch = torch_cuda
    from mmwave_radar_processing_b200.distributed import LocalImagePipeline
    from mmwave_radar_processing_b200.engine import RadarEngine
    s, n, g = 90, 700, 1500
    x, pos, dirs, lam = sar_case(s, n, g, seed=21)
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    eng = RadarEngine(1, s, 1, 1)
    want = eng.sar_image(xs, ps, ds, lam).clone()
    pipe = LocalImagePipeline(eng, g, depth=2)
    for k in range(3):                                       # eager first (sizes the scratch of both contexts)
        img = pipe.run(k, xs, ps, ds, lam)
        pipe.wait(k)
        assert torch.equal(img, want)
    pipe.capture(xs, ps, ds, lam)
    imgs = [pipe.run(k, xs, ps, ds, lam) for k in (0, 1)]    # both in flight
    for k in (0, 1):
        pipe.wait(k)
    torch.cuda.synchronize()
    assert torch.equal(imgs[0], want) and torch.equal(imgs[1], want)
    xs.mul_(2.0)                                             # new input in the captured tensor: exact power-of-two scale
    img = pipe.run(2, xs, ps, ds, lam)
    pipe.wait(2)
    torch.cuda.synchronize()
    assert torch.equal(img, want * 2.0)


def test_input_scale_leaves_no_trace(torch_cuda):
    """The FP16 operands are scaled by exact powers of two taken from the largest sample of the call and the range-FFT kernel
    undoes them: an aperture cube scaled by 2^80 or 2^-80 must give the image scaled by exactly that, bit for bit."""
    torch = torch_cuda
    s, n, g = 90, 300, 200
    x, pos, dirs, lam = sar_case(s, n, g, seed=33)
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    eng = engine_with_mode(0, s)
    want = eng.sar_image(xs, ps, ds, lam).clone()
    for p2 in (80, -80, 1, -7):
        got = eng.sar_image(xs * (2.0 ** p2), ps, ds, lam)
        assert torch.equal(got, want * (2.0 ** p2)), p2
    zero = eng.sar_image(torch.zeros_like(xs), ps, ds, lam)
    assert torch.count_nonzero(torch.view_as_real(zero)) == 0
