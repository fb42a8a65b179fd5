#!/usr/bin/env python
"""Timing of radar_sar_image on the throughput grid of SURVEY.md section 8d (S = 90, N = 2000, G = 32768) for the
FP32-pipe kernel and the tcgen05 split-FP16 path.  One JSON line per mode: ms per image, useful complex-GEMM TFLOP/s
(8 S N G flop) and, for the tensor-core path, executed FP16 TFLOP/s (3 MMAs per product, padded tile) against the dense
bf16 / fp16 peak of MEASURED_PEAKS.json.  CUDA events on the launching stream, warm-up first."""
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
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    bf16_peak = float(peaks.get("bf16_tflops", peaks.get("tensor_bf16_tflops", 0.0)) or 0.0)
    ref = None
    for mode, dbg, tpr, label in ((1, 0, 0, "fp32_pipe"), (2, 0, 0, "tcgen05_f16x3_1tile"), (3, 0, 0, "tcgen05_f16x3_2tiles"),
                                  (3, 0, 2, "tcgen05_f16x3_2tiles_tpr2"), (2, 0, 1, "tcgen05_f16x3_1tile_tpr1"), (0, 0, 0, "auto"),
                                  (0, 64, 0, "auto_no_drain_stores_INVALID"), (0, 32, 0, "auto_general_phase_path")):
        os.environ["RADAR_B200_SAR"] = str(mode)
        os.environ["RADAR_B200_SAR_DBG"] = str(dbg)
        os.environ["RADAR_B200_SAR_TPR"] = str(tpr)
        eng = RadarEngine(1, s, 1, 1)
        out = torch.empty((s, g), dtype=torch.complex64, device="cuda")
        for _ in range(3):
            eng.sar_image(xs, ps, ds, 0.0039, out=out)
        torch.cuda.synchronize()
        reps = 5 if mode == 1 else 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            eng.sar_image(xs, ps, ds, 0.0039, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        if ref is None:
            ref = out.clone()
        line = {"kernel": label, "S": s, "N": n, "G": g, "ms_per_image": ms,
                "useful_tflops": 8.0 * s * n * g / (ms * 1e-3) / 1e12,
                "rel_l2_vs_fp32_pipe": float((torch.linalg.norm(out - ref) / torch.linalg.norm(ref)).item())}
        if mode != 1:
            executed = 3 * 2.0 * 192 * (2 * 2000) * g        # three MMAs, N padded to 192 columns, K = 2 N real
            line["executed_f16_tflops"] = executed / (ms * 1e-3) / 1e12
            if bf16_peak:
                line["frac_of_measured_bf16_peak"] = line["executed_f16_tflops"] / bf16_peak
                line["measured_bf16_peak_tflops"] = bf16_peak
        print(json.dumps(line), flush=True)

    # more than 128 range samples (RadarHD: S = 512): column tiles of 96 samples on the tensor cores against the FP32-pipe kernel
    s2, g2 = 512, 8192
    x2 = (rng.integers(-2000, 2000, (s2, n)) + 1j * rng.integers(-2000, 2000, (s2, n))).astype(np.complex64)
    x2s, d2s = torch.from_numpy(x2).cuda(), ds[:, :g2].contiguous()
    os.environ["RADAR_B200_SAR_DBG"] = "0"
    os.environ["RADAR_B200_SAR_TPR"] = "0"
    ref2 = None
    for mode, label, reps in ((1, "fp32_pipe_S512", 3), (0, "tcgen05_f16x3_S512_column_tiles", 10)):
        os.environ["RADAR_B200_SAR"] = str(mode)
        eng = RadarEngine(1, s2, 1, 1)
        out = torch.empty((s2, g2), dtype=torch.complex64, device="cuda")
        for _ in range(2):
            eng.sar_image(x2s, ps, d2s, 0.0039, out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            eng.sar_image(x2s, ps, d2s, 0.0039, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        if ref2 is None:
            ref2 = out.clone()
        print(json.dumps({"kernel": label, "S": s2, "N": n, "G": g2, "ms_per_image": ms,
                          "useful_tflops": 8.0 * s2 * n * g2 / (ms * 1e-3) / 1e12,
                          "rel_l2_vs_fp32_pipe": float((torch.linalg.norm(out - ref2) / torch.linalg.norm(ref2)).item())}), flush=True)


if __name__ == "__main__":
    main()
