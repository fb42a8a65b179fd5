#!/usr/bin/env python
"""Accumulation-segment length of the tensor-core SAR kernel (RADAR_B200_SAR_SEG, chunks of 32 elements per segment)
against the error it leaves: relative L2 of 256 directions of the throughput grid (S = 90, N = 2000, G = 32768) against a
fp64 evaluation, and the time per image.  One JSON line per setting."""
import json
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from mmwave_radar_processing_b200.engine import RadarEngine  # noqa: E402


def main():
    s, n, g, lam = 90, 2000, 32768, 0.0039
    rng = np.random.default_rng(0)
    x = (rng.integers(-2000, 2000, (s, n)) + 1j * rng.integers(-2000, 2000, (s, n))).astype(np.complex64)
    t = np.arange(n) * 1e-4
    pos = np.stack([0.01 * t, 0.5 * t, np.zeros(n)])
    az, el = rng.uniform(-1, 1, g), rng.uniform(-0.3, 0.3, g)
    dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
    cols = rng.choice(g, 256, replace=False)
    w = np.hamming(n)[:, None] * np.exp(2j * np.pi * (pos.T @ dirs[:, cols]) / lam)
    want = np.fft.fft((x.astype(np.complex128) @ w) * np.hanning(s)[:, None], axis=0)
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    cols_d = torch.from_numpy(cols).cuda()
    for seg in (0, 2, 4, 16, 32, 63):
        os.environ["RADAR_B200_SAR_SEG"] = str(seg)
        eng = RadarEngine(1, s, 1, 1)
        out = torch.empty((s, g), dtype=torch.complex64, device="cuda")
        for _ in range(3):
            eng.sar_image(xs, ps, ds, lam, out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            eng.sar_image(xs, ps, ds, lam, out=out)
        e1.record()
        torch.cuda.synchronize()
        got = out[:, cols_d].cpu().numpy().astype(np.complex128)
        err = float(np.linalg.norm(got - want) / np.linalg.norm(want))
        print(json.dumps({"seg_chunks": seg or "default", "ms_per_image": e0.elapsed_time(e1) / 20, "rel_l2_vs_fp64": err}), flush=True)


if __name__ == "__main__":
    main()
