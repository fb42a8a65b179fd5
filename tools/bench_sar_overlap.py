#!/usr/bin/env python
"""One GPU, consecutive synthetic-array images on two streams with two contexts (each context's scratch serves one image at
a time): does the max / prep / range-FFT work of one image hide under the GEMM of the other?  One JSON line per mode."""
import json
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from mmwave_radar_processing_b200.engine import RadarEngine  # noqa: E402


def main():
    s, n, g = 90, 2000, 32768
    rng = np.random.default_rng(0)
    x = (rng.integers(-2000, 2000, (s, n)) + 1j * rng.integers(-2000, 2000, (s, n))).astype(np.complex64)
    t = np.arange(n) * 1e-4
    pos = np.stack([0.01 * t, 0.5 * t, np.zeros(n)])
    az, el = rng.uniform(-1, 1, g), rng.uniform(-0.3, 0.3, g)
    dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    for n_ctx in (1, 2, 3):
        engs = [RadarEngine(1, s, 1, 1) for _ in range(n_ctx)]
        outs = [torch.empty((s, g), dtype=torch.complex64, device="cuda") for _ in range(n_ctx)]
        streams = [torch.cuda.Stream() for _ in range(n_ctx)]
        for e, o in zip(engs, outs):
            for _ in range(3):
                e.sar_image(xs, ps, ds, 0.0039, out=o)
        torch.cuda.synchronize()
        graphs = []
        for e, o in zip(engs, outs):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                e.sar_image(xs, ps, ds, 0.0039, out=o)
            graphs.append(gr)
        torch.cuda.synchronize()
        reps = 60
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for st in streams:
            st.wait_stream(torch.cuda.current_stream())
        for i in range(reps):
            with torch.cuda.stream(streams[i % n_ctx]):
                graphs[i % n_ctx].replay()
        for st in streams:
            torch.cuda.current_stream().wait_stream(st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        same = all(torch.equal(outs[0], o) for o in outs[1:])
        print(json.dumps({"contexts_and_streams": n_ctx, "ms_per_image": ms, "images_per_s": 1e3 / ms, "images_identical": same}), flush=True)


if __name__ == "__main__":
    main()
