#!/usr/bin/env python
"""Timing lines for the range-azimuth heat maps (BASELINE.json configs[2]: RadarHD / RadCloud, 64 / 256 / 1024 angle bins)
and the Doppler-beam-sharpened map (RadSAR_1843_official, 100 output bins): frames/s and fraction of the measured HBM peak
(algorithmic bytes of SURVEY.md section 8d: one chirp per TX slot in, 4*S*A out).  One JSON line per case.
usage: python tools/bench_ra_dbs.py [--frames 2048] [--steps 30]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))


def main():
    from bench import hbm_peak, load_cfg
    from oracle import radar_oracle as orc                      # bin selection of the DBS case only (host, fp64, untimed)
    from mmwave_radar_processing_b200.engine import RadarEngine
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=30)
    a = ap.parse_args()
    peak, peak_src = hbm_peak()

    def timed(fn):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / a.steps

    for name in (() if os.environ.get("BENCH_DBS_ONLY") else ("RadarHD", "RadCloud")):
        cm = load_cfg(name, "standard")
        eng = RadarEngine(cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
        iq = torch.from_numpy(np.tile(synth_batch_for_cfg(cm, range(64)), (a.frames // 64, 1, 1, 1, 1))).cuda()
        f = iq.shape[0]
        for bins in (64, 256, 1024):
            out = torch.empty((f, eng.S, bins), dtype=torch.float32, device="cuda")
            ms = timed(lambda: eng.range_angle(iq, 0, None, bins, out=out))
            alg = 4 * eng.R * eng.S * eng.cpl + 4 * eng.S * bins
            print(json.dumps({"kernel": "range_angle_kernel", "workload": f"configs/{name}.cfg range-azimuth map, {bins} angle bins, all {eng.V} antennas",
                              "frames": f, "ms_per_launch": ms, "frames_per_s": f / ms * 1e3, "algorithmic_bytes_per_frame": alg,
                              "achieved_gbs": alg * f / ms / 1e6, "peak_gbs": peak, "peak_source": peak_src, "frac": alg * f / ms / 1e6 / peak}))
            del out
    cm = load_cfg("RadSAR_1843_official", "standard")
    eng = RadarEngine(cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    frames = min(a.frames, 1024)
    iq = torch.from_numpy(np.tile(synth_batch_for_cfg(cm, range(32)), (frames // 32, 1, 1, 1, 1))).cuda()
    _, vb = orc.rd_axes(cm.vel_max_m_s, cm.vel_res_m_s, cm.range_max_m, cm.range_res_m)
    _, coarse = orc.angle_axis(64)
    fine = np.linspace(coarse[0], coarse[-1], 100)
    vbin, abin = orc.dbs_bin_selection(coarse, fine, vb, np.array([0.5, 0.1, 0.0]))
    f = iq.shape[0]
    out = torch.empty((f, eng.S, 100), dtype=torch.float32, device="cuda")
    alg = 4 * eng.R * eng.S * eng.C + 4 * eng.S * 100           # whole int16 cube in, one fp32 map out
    flops = 8.0 * eng.S * eng.V * eng.L * 100                   # GEMM form of the restricted 3-D DFT
    shape = (cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    for mode, chunk, label in (("1", "0", "dbs_kernel (one kernel, shared-memory FFT)"), ("0", "0", "rd_reg_dbs_kernel (sharpening inside the register-resident range-Doppler kernel)"),
                               ("2", "0", "rd_reg_kernel + dbs_combine_kernel, default blocks (1 GiB of scratch)"),
                               ("2", "64", "rd_reg_kernel + dbs_combine_kernel, blocks of 64 frames"), ("2", "113", "rd_reg_kernel + dbs_combine_kernel, blocks of 113 frames (L2-sized)"),
                               ("2", "256", "rd_reg_kernel + dbs_combine_kernel, blocks of 256 frames"),
                               ("2", "1024", "rd_reg_kernel + dbs_combine_kernel, one block")):
        os.environ["RADAR_B200_DBS"] = mode
        os.environ["RADAR_B200_DBS_CHUNK"] = chunk
        eng = RadarEngine(*shape)
        ms = timed(lambda: eng.dbs(iq, vbin.tolist(), abin.tolist(), ant_list=None, num_angle_bins=64, out=out))
        print(json.dumps({"kernel": label, "workload": "configs/RadSAR_1843_official.cfg Doppler beam sharpening, 100 output bins, velocity (0.5, 0.1, 0)",
                          "frames": f, "ms_per_launch": ms, "frames_per_s": f / ms * 1e3, "algorithmic_bytes_per_frame": alg,
                          "achieved_gbs": alg * f / ms / 1e6, "peak_gbs": peak, "frac": alg * f / ms / 1e6 / peak,
                          "useful_tflops": flops * f / ms / 1e9}), flush=True)


if __name__ == "__main__":
    main()
