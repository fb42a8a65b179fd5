"""radar_chirp_zoom_static: the allocation-free variant of radar_chirp_zoom (twiddle table in a static device array of the
library instead of the context's cudaMalloc'ed scratch).  It has no counterpart in the reference; it was added
for callers that cannot call cudaMalloc.  The file sorts last on purpose (pytest -x).

Checked through the C ABI: two calls in a row (different frequency lists, the second one shorter so stale entries of the
shared table would show) against radar_chirp_zoom on the same inputs and against a direct fp64 evaluation of the sums,
the size limit, and that the device's free memory does not move across the calls."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-4            # relative L2 against the fp64 sums, as tests/test_gpu_da.py (fp32 device arithmetic, measured ~2e-7)

R, S, CPL, L = 4, 64, 2, 40          # small TDM shape: V = 8 virtual antennas


@pytest.fixture(scope="module")
def setup():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from mmwave_radar_processing_b200 import _native as nat
    from mmwave_radar_processing_b200.engine import RadarEngine
    eng = RadarEngine(R, S, CPL, L, device="cuda:0")
    return torch, nat, eng


def rel_l2(a, b):
    a = np.asarray(a).astype(np.complex128).ravel()
    b = np.asarray(b).astype(np.complex128).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def zoom_fp64(cube, turns, k0, k1):
    """X[f, v, k, j] = sum_l cube[f, v, k0 + k, l] exp(-2 pi i l turns[j]) in fp64."""
    l = np.arange(cube.shape[-1], dtype=np.float64)
    tw = np.exp(-2j * np.pi * np.outer(l, np.asarray(turns, dtype=np.float64)))
    return cube[:, :, k0:k1, :].astype(np.complex128) @ tw


def call(torch, nat, eng, fn, cube_d, turns_d, k0, k1, out_d):
    f, n_ant = cube_d.shape[:2]
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return fn(eng.ctx, nat.ptr(cube_d), f, n_ant, k0, k1, nat.ptr(turns_d), turns_d.numel(), nat.ptr(out_d), stream)


def test_static_chirp_zoom_matches_allocating_entry_point_twice(setup):
    torch, nat, eng = setup
    lib = eng.lib
    assert lib.radar_chirp_zoom_static_capacity() == 65536
    rng = np.random.default_rng(20261105)
    F, n_ant = 3, R * CPL
    cube = (rng.standard_normal((F, n_ant, S, L)) + 1j * rng.standard_normal((F, n_ant, S, L))).astype(np.complex64)
    cube_d = torch.from_numpy(cube).cuda()
    # call 1: L frequencies around zero Doppler over all rows; call 2: fewer, other frequencies, a row window
    cases = [(np.linspace(-0.5, 0.5, L, endpoint=False) / 4.0, 0, S),
             (0.11 + np.arange(17, dtype=np.float64) * 0.0137, 5, 41)]
    bufs = []
    for turns, k0, k1 in cases:
        t_d = torch.from_numpy(np.ascontiguousarray(turns)).cuda()
        shape = (F, n_ant, k1 - k0, t_d.numel())
        bufs.append((t_d, torch.zeros(shape, dtype=torch.complex64, device="cuda"),
                     torch.zeros(shape, dtype=torch.complex64, device="cuda")))
    # set-up only (n_frames = 0): looks the static table's address up, launches nothing
    nat.check(lib.radar_chirp_zoom_static(eng.ctx, None, 0, n_ant, 0, S, None, L, None, None), eng.ctx)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for (turns, k0, k1), (t_d, out_static, _) in zip(cases, bufs):
        nat.check(call(torch, nat, eng, lib.radar_chirp_zoom_static, cube_d, t_d, k0, k1, out_static), eng.ctx)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    assert free_after >= free_before, "the static entry point must not take device memory"
    for (turns, k0, k1), (t_d, out_static, out_ref) in zip(cases, bufs):
        nat.check(call(torch, nat, eng, lib.radar_chirp_zoom, cube_d, t_d, k0, k1, out_ref), eng.ctx)
        torch.cuda.synchronize()
        got, ref = out_static.cpu().numpy(), out_ref.cpu().numpy()
        assert np.isfinite(got.view(np.float32)).all()
        assert rel_l2(got, ref) < TOL                     # same kernels, same table values: equal in practice
        np.testing.assert_array_equal(got, ref)
        assert rel_l2(got, zoom_fp64(cube, turns, k0, k1)) < TOL
    # the engine method is the same call
    t_d, out_static, _ = bufs[1]
    again = torch.zeros_like(out_static)
    eng.chirp_zoom_static(cube_d, t_d, again, cases[1][1], cases[1][2])
    torch.cuda.synchronize()
    np.testing.assert_array_equal(again.cpu().numpy(), out_static.cpu().numpy())


def test_static_chirp_zoom_rejects_sizes_beyond_the_table(setup):
    torch, nat, eng = setup
    lib = eng.lib
    n_out = 65536 // L + 1                                 # L * n_out just above the 65536-entry table
    cube_d = torch.zeros((1, 1, S, L), dtype=torch.complex64, device="cuda")
    t_d = torch.zeros(n_out, dtype=torch.float64, device="cuda")
    out_d = torch.zeros((1, 1, S, n_out), dtype=torch.complex64, device="cuda")
    rc = call(torch, nat, eng, lib.radar_chirp_zoom_static, cube_d, t_d, 0, S, out_d)
    assert rc == -2                                        # RADAR_E_UNSUPPORTED
    msg = lib.radar_last_error(eng.ctx).decode()
    assert "exceeds the static twiddle table" in msg and "65536" in msg
    # the allocating entry point takes the same request
    nat.check(call(torch, nat, eng, lib.radar_chirp_zoom, cube_d, t_d, 0, S, out_d), eng.ctx)
    # the largest size that fits is accepted
    n_fit = 65536 // L
    nat.check(call(torch, nat, eng, lib.radar_chirp_zoom_static, cube_d, t_d[:n_fit].contiguous(), 0, S,
                   torch.zeros((1, 1, S, n_fit), dtype=torch.complex64, device="cuda")), eng.ctx)
    torch.cuda.synchronize()
