#!/usr/bin/env python
"""Kernel time of one rank's voxel slice of the SAR image (S = 90, N = 2000, G = 512 x 64) for 1/2/4/8-way sharding, with
and without the split of the aperture over CTA rows (RADAR_B200_SAR_KSPLIT=1 = off).  One GPU.  usage: python tools/bench_sar_slices.py"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mmwave_radar_processing_b200.engine import RadarEngine
s, n, g = 90, 2000, 512 * 64
rng = np.random.default_rng(0)
x = (rng.integers(-2000, 2000, (s, n)) + 1j * rng.integers(-2000, 2000, (s, n))).astype(np.complex64)
t = np.arange(n) * 1e-4
pos = np.stack([0.01 * t, 0.5 * t, np.zeros(n)])
az = np.repeat(np.linspace(-1.0, 1.0, 512), 64); el = np.tile(np.linspace(-0.3, 0.3, 64), 512)
dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
for ks in (1, 0):
    os.environ["RADAR_B200_SAR_KSPLIT"] = str(ks)
    eng = RadarEngine(1, s, 1, 1)
    del os.environ["RADAR_B200_SAR_KSPLIT"]
    for ways in (1, 2, 4, 8):
        hi = g // ways
        for _ in range(5):
            eng.sar_image(xs, ps, ds, 0.0039, 0, hi)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30):
            eng.sar_image(xs, ps, ds, 0.0039, 0, hi)
        e1.record(); torch.cuda.synchronize()
        print(json.dumps({"ksplit": "auto" if ks == 0 else 1, "ways": ways, "directions": hi, "ms_per_slice": e0.elapsed_time(e1) / 30}))
