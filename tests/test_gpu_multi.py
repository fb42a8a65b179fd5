"""Multi-GPU checks that need >= 2 devices on the box (skipped on a single GPU): launched through torchrun."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import REPO, load_cfg

pytestmark = pytest.mark.gpu


def test_peer_point_exchange_matches_nccl_allgather():
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(REPO, "tests", "multi_gpu_peer_exchange.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert res.returncode == 0 and "PEER_EXCHANGE_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


def test_sar_voxel_sharded_matches_single_gpu():
    """BASELINE.json configs[3] on two GPUs: broadcast + per-rank direction slices + all-gather == one GPU's image."""
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29534", os.path.join(REPO, "tests", "multi_gpu_sar_sharded.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert res.returncode == 0 and "SAR_SHARDED_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


def test_point_mirrors_single_device():
    """radar_ctx_set_point_mirrors on one GPU: two local mirrors receive every packed row below the capacity, nothing
    beyond it; the padded layout ignores the mirrors; switching off restores the plain call."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    cm = load_cfg("6843_RadVel_ods_20Hz", "ods")
    pipe = FramePipeline(cm, detector=DetectorSpec(kind="os", num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5),
                         az_antenna_idxs=(0, 3, 4, 7), el_antenna_idxs=(9, 8, 5, 4), max_dets=256)
    iq = torch.from_numpy(synth_batch_for_cfg(cm, range(10))).cuda()
    out = pipe.run(iq)
    want = out["points"].clone()
    total = int(out["offsets"][10])
    assert total > 100
    cap = total - 37
    m0 = torch.full((cap + 64, 4), -7.0, dtype=torch.float64, device="cuda")
    m1 = torch.full((cap + 64, 4), -7.0, dtype=torch.float64, device="cuda")
    pipe.engine.set_point_mirrors([m0.data_ptr(), m1.data_ptr()], cap)
    try:
        out = pipe.run(iq)
        assert torch.equal(out["points"][:total], want[:total])
        for m in (m0, m1):
            assert torch.equal(m[:cap], want[:cap]) and bool((m[cap:] == -7.0).all())
        with pytest.raises(Exception):
            pipe.engine.set_point_mirrors([m0.data_ptr() + 8], cap)          # misaligned
    finally:
        pipe.engine.set_point_mirrors([], 0)
    m0.fill_(-7.0)
    pipe.run(iq)
    assert bool((m0 == -7.0).all())
