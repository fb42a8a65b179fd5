#!/usr/bin/env python
"""bench.py - frames/s of the ADC-cube -> point-cloud hot path on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--frames F] [--impl reference]

Default workload (BASELINE.json configs[4], the shape north_star targets): the **6843 ODS synthetic stream of 100 000
frames** (configs/6843_RadVel_ods_20Hz.cfg: 63 samples x 70 loops, 3 TX x 4 RX = 12 virtual antennas).  A *step* is one
pass of the hot path over the whole stream; with N > 1 the stream is cut into N contiguous frame blocks (strong scaling):

    radar_range_doppler (all V antennas, complex RD cube + |RD[0]|)  ->  radar_range_angle (heat map)
      ->  radar_cfar_2d (OS-CFAR-2D + ordered compaction)  ->  radar_det_offsets  ->  radar_point_cloud
      [N > 1: radar_point_cloud also stores every packed row into this rank's slot of a symmetric-memory gathered buffer;
       the slot, the counts and the total go to the peers by copy-engine peer copies over NVLink + a cross-rank signal-pad
       barrier on a side stream, under the kernels of the next step (--exchange peer: the kernel stores to every rank
       itself; --exchange nccl: staging copy + NCCL all-gathers).  NCCL itself only carries the sizing all-reduces.]

``value`` is whole-job frames/s with the inputs resident in HBM; ``e2e`` is the same metric through the host-facing
call (``FramePipeline.run_host``): pinned host int16 frames in, host point clouds out, copies inside the timed region.
``roofline`` is for the dominant kernel (the fused range-Doppler kernel) and lists every stage kernel under ``kernels``;
``cpu_baseline`` is the numpy oracle port timed on this box's host cores (rank 0, N=1 only).  ``extra`` (N=1) carries
the second single-GPU workload (configs/1843_RadVel.cfg, BASELINE.json configs[1]), the CA-CFAR-2D detector timing, the
other BASELINE shapes (RadCloud, RadSAR_1843_official, RadarHD) through the same pipeline and the high-resolution
range-azimuth maps of configs[2] (RadarHD, 256 / 1024 angle bins).
``--impl reference`` times the CPU path on all host cores and prints the same line shape.
``--workload sar``: the voxel-sharded synthetic-array image (BASELINE.json configs[3]), images/s + tensor-pipe roofline.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

METRIC = "frames/s ADC cube->point cloud"
WORKLOADS = {
    # name: (cfg file, array geometry, az antennas, el antennas, RA antennas)
    "1843_RadVel": ("1843_RadVel", "standard", (0, 1, 2, 3), (), (0, 1, 2, 3)),
    "6843_RadVel_ods_20Hz": ("6843_RadVel_ods_20Hz", "ods", (0, 3, 4, 7), (9, 8, 5, 4), (0, 3, 4, 7)),
    "RadSAR_1843_official": ("RadSAR_1843_official", "standard", (0, 1, 2, 3), (), (0, 1, 2, 3)),
    "RadCloud": ("RadCloud", "standard", (0, 1, 2, 3), (), (0, 1, 2, 3)),
    "RadarHD": ("RadarHD", "standard", (0, 1, 2, 3), (), (0, 1, 2, 3)),
}
DEFAULT_WORKLOAD = "6843_RadVel_ods_20Hz"
STREAM_FRAMES = 100_000          # BASELINE.json configs[4]: the whole stream is one step, split over the ranks
# detector of the reference's PointCloudGenerator integration test (tests/verify_processors.py:121-143)
DETECTOR = dict(kind="os", num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5)
DETECTOR_CA = dict(kind="ca", num_train=(4, 4), num_guard=(2, 2), pfa=1e-5)      # verify_processors.py:156-191
MAX_DETS = 256
N_UNIQUE = 256          # distinct synthetic frames per rank, tiled to the batch
E2E_FRAMES_MAX = 16384  # frames per rank of the end-to-end (host buffers) region: bounds the pinned allocation
FALLBACK_HBM_GBS = 6650.0


def load_cfg(name, geometry):
    from mmwave_radar_processing_b200.config_managers.cfgManager import ConfigManager
    cm = ConfigManager()
    cm.load_cfg(os.path.join(REPO, "configs", name + ".cfg"), array_geometry=geometry)
    return cm


def measured_peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh)
    except Exception:  # noqa: BLE001
        return {}


def hbm_peak():
    p = measured_peaks()
    if "hbm_gbs" in p:
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel: str, workload: str):
    """(dram bytes per launch per frame, capture tag) from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(REPO, "profiles", "traffic.json")) as fh:
            d = json.load(fh)
        return d.get(workload, {}).get(kernel), d.get("_capture", "static ncu capture (profiles/traffic.json)")
    except Exception:  # noqa: BLE001
        return None, None


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with NVML while the timed regions run."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int, period_s: float = 0.004):
        super().__init__(daemon=True)
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:  # noqa: BLE001
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._stop_evt.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((mhz, util))
                for bit, name in self.REASONS.items():
                    if bits & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        loaded = [m for m, u in self.samples if u > 0] or [m for m, _ in self.samples]
        return {"sm_mhz": statistics.median(loaded), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples), "samples_under_load": len(loaded)}


# ------------------------------------------------------------------------------------------------ CPU path
def _cpu_frame(args):
    """One frame through the numpy oracle port of the reference path (the checker, timed as the CPU baseline)."""
    from oracle import radar_oracle as orc
    iq, shape, axes, az, el, ra_ants, det = args
    r, s, cpl, l = shape
    cube = orc.virtual_array(orc.iq_to_cube(iq), r, cpl, l)
    rd = orc.range_doppler(cube)
    ra = orc.range_angle(cube, 64, 0, ra_ants, True)
    _, _, mask = orc.cfar_2d(orc.rd_magnitude0(rd), det["kind"], det["num_train"], det["num_guard"],
                             rho=det.get("rho", 0.0), alpha=det.get("alpha", 0.0), pfa=det.get("pfa", 0.0))
    dets = orc.mask_to_dets(mask)
    pts, _, _ = orc.point_cloud(rd, dets, axes[0], axes[1], list(az), list(el))
    return pts.shape[0], float(ra[0, 0])


def _cpu_args(cm, wl, frames_iq):
    from mmwave_radar_processing_b200.pipeline import reference_axes
    rb, vb, _ = reference_axes(cm)
    shape = (cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    _, _, az, el, ra_ants = WORKLOADS[wl]
    return [(frames_iq[i], shape, (rb, vb), az, el, ra_ants, DETECTOR) for i in range(frames_iq.shape[0])]


def cpu_baseline_single(cm, wl, frames_iq, budget_s=12.0):
    """Single-thread oracle port on a bounded sample: as many frames as fit in ~budget_s (at least 16)."""
    try:
        from threadpoolctl import threadpool_limits
        limiter = threadpool_limits(limits=1)
    except Exception:  # noqa: BLE001
        limiter = None
    args = _cpu_args(cm, wl, frames_iq)
    for a in args[:4]:
        _cpu_frame(a)
    t0 = time.perf_counter()
    n = 0
    while True:
        _cpu_frame(args[n % len(args)])
        n += 1
        dt = time.perf_counter() - t0
        if (dt > budget_s and n >= 16) or n >= 20000:
            break
    if limiter is not None:
        limiter.unregister() if hasattr(limiter, "unregister") else None
    return {"value": n / dt, "unit": "frames/s", "cores": 1, "kind": "port",
            "sample": f"{n} frames of {wl} through oracle/radar_oracle.py (numpy fp64, 1 thread), {dt:.1f} s"}


def frames_per_rank(a, world, rank=0):
    """(frames of this rank per step, global frames per step, scaling).  --frames 0: the 100k-frame stream in contiguous
    blocks (strong scaling); --frames F: F frames per GPU (weak scaling)."""
    from mmwave_radar_processing_b200.distributed import shard_range
    if a.frames > 0:
        return a.frames, a.frames * world, "weak"
    lo, hi = shard_range(STREAM_FRAMES, rank, world)
    return hi - lo, STREAM_FRAMES, "strong"


def run_reference_arm(a):
    """--impl reference: the CPU path on all host cores (process pool sharded by frame), rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    cfg_name, geom = WORKLOADS[a.workload][:2]
    cm = load_cfg(cfg_name, geom)
    cores = os.cpu_count() or 1
    uniq = synth_batch_for_cfg(cm, range(64))
    args = _cpu_args(cm, a.workload, uniq)
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        # size a step so that warmup+steps stay within ~2 minutes: probe the per-frame cost first
        t0 = time.perf_counter()
        pool.map(_cpu_frame, args[:cores * 2], chunksize=1)
        per_frame_wall = (time.perf_counter() - t0) / (cores * 2)
        budget = 90.0 / max(1, a.steps + a.warmup)
        frames_per_step = int(max(cores * 2, min(4096, budget / per_frame_wall)))
        batch = [args[i % len(args)] for i in range(frames_per_step)]
        for _ in range(a.warmup):
            pool.map(_cpu_frame, batch, chunksize=max(1, frames_per_step // (cores * 4)))
        t0 = time.perf_counter()
        for _ in range(a.steps):
            pool.map(_cpu_frame, batch, chunksize=max(1, frames_per_step // (cores * 4)))
        dt = time.perf_counter() - t0
    val = a.steps * frames_per_step / dt
    sample = (f"{frames_per_step} frames per timed step (a bounded sample of the workload in config) of {a.workload} through "
              f"oracle/radar_oracle.py (numpy fp64), {cores}-process pool sharded by frame")
    world = max(1, a.gpus)
    f_rank, f_global, scaling = frames_per_rank(a, world)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(cm, a.workload, f_rank, f_global, world, a.exchange),
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


EXCHANGE_TEXT = {
    "copy": " + point rows stored by the point-cloud kernel into this rank's slot of a symmetric-memory gathered buffer, "
            "pushed to the peers by copy-engine peer copies over NVLink + signal-pad barrier (side stream, under the next "
            "step); NCCL carries only the sizing all-reduces",
    "peer": " + point rows stored into every rank's symmetric-memory gathered buffer by the point-cloud kernel (NVLink "
            "peer stores), counts by copy-engine peer copies + signal-pad barrier on a side stream",
    "nccl": " + NCCL all-gather of point clouds (own stream, under the next step)",
}


def workload_config(cm, wl, frames_per_gpu, global_frames, n_gpus, exchange="copy"):
    r, s, cpl, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    stream = global_frames == STREAM_FRAMES
    return {"workload": (f"configs/{WORKLOADS[wl][0]}.cfg synthetic int16 ADC stream of {global_frames} frames" +
                         (" cut into contiguous frame blocks, one per GPU" if stream else " (batched cubes)")),
            "num_rx": r, "chirps_per_loop": cpl, "virtual_antennas": r * cpl, "samples": s, "loops": l,
            "frames_per_gpu_per_step": frames_per_gpu, "global_frames_per_step": global_frames,
            "detector": "os_cfar_2d num_train=(8,4) num_guard=(2,1) rho=0.75 alpha=3.5",
            "stages": "range_doppler+range_angle+cfar_2d+det_offsets+point_cloud",
            "parallelism": f"frame-sharded x{n_gpus}" + (EXCHANGE_TEXT[exchange] if n_gpus > 1 else ""),
            "l2_policy": "per-step working set (int16 in + complex RD out) exceeds the 126 MB L2; no explicit flush"}


def stage_rooflines(cm, wl, F, stage_ms, ms_per_step, points, peak, rd_kernel="rd_reg_kernel"):
    """Algorithmic bytes per frame (DESIGN.md section 4 / SURVEY.md section 8d) and achieved fraction of the HBM peak for
    every stage kernel of one step."""
    R, S, cpl, L = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    V, C = R * cpl, cpl * L
    _, _, az, el, ra_ants = WORKLOADS[wl]
    ppf = points / max(1, F)
    bytes_pf = {
        "range_doppler": 4 * R * S * C + 8 * V * S * L + 4 * S * L,
        "range_angle": 4 * R * S * cpl + 4 * S * 64,
        "cfar": 4 * S * L + 8 * ppf + 4,
        "point_cloud": (8 * (len(az) + len(el)) + 8 + 32) * ppf,
    }
    names = {"range_doppler": rd_kernel, "range_angle": "range_angle_kernel", "cfar": "cfar_kernel",
             "point_cloud": "point_cloud_thread_kernel"}
    out = []
    for st, ms in stage_ms.items():
        b = bytes_pf[st]
        ach = b * F / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        out.append({"stage": st, "kernel": names[st], "algorithmic_bytes_per_frame": b, "ms_per_launch": ms,
                    "achieved_gbs": ach, "frac": ach / peak, "share_of_step": ms / ms_per_step})
    step_bytes = sum(bytes_pf[st] for st in stage_ms)
    return out, bytes_pf, step_bytes


def time_stage_loop(torch, pipe, iq_dev, steps, warmup):
    """Plain single-GPU loop with per-stage events -> (ms_total, stage_ms, last out)."""
    for _ in range(warmup):
        pipe.run(iq_dev)
    torch.cuda.synchronize()
    marks = []

    def mark(name):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        marks.append((name, ev))
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    out = None
    for _ in range(steps):
        mark("begin")
        out = pipe.run(iq_dev, mark)
    t1.record()
    torch.cuda.synchronize()
    stage = {}
    for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
        if n1 != "begin":
            stage[n1] = stage.get(n1, 0.0) + e0.elapsed_time(e1)
    return t0.elapsed_time(t1), {k: v / steps for k, v in stage.items()}, out


def extra_single_gpu_lines(torch, dev, peak):
    """N=1 extras: (a) configs/1843_RadVel.cfg, 4096 frames per step (BASELINE.json configs[1]; round-1 headline, kept for
    continuity); (b) the CA-CFAR-2D detector of tests/verify_processors.py:156-191 on both shapes; (c) the other BASELINE
    shapes through the pipeline, the high-resolution range-azimuth maps, Doppler beam sharpening and the synthetic-array image."""
    import numpy as np
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    extras = []
    for wl, det_name, det in (("1843_RadVel", "os_cfar_2d", DETECTOR), ("1843_RadVel", "ca_cfar_2d", DETECTOR_CA),
                              (DEFAULT_WORKLOAD, "ca_cfar_2d", DETECTOR_CA),
                              # the other BASELINE shapes (configs[2], configs[3]) through the same pipeline
                              ("RadCloud", "os_cfar_2d", DETECTOR), ("RadSAR_1843_official", "os_cfar_2d", DETECTOR),
                              ("RadarHD", "os_cfar_2d", DETECTOR)):
        cfg_name, geom, az, el, ra_ants = WORKLOADS[wl]
        cm = load_cfg(cfg_name, geom)
        F = 4096 if wl in ("1843_RadVel", DEFAULT_WORKLOAD, "RadCloud") else 2048
        pipe = FramePipeline(cm, detector=DetectorSpec(**det), az_antenna_idxs=az, el_antenna_idxs=el, max_dets=MAX_DETS,
                             with_range_angle=True, ra_antennas=ra_ants, device=dev)
        uniq = synth_batch_for_cfg(cm, range(N_UNIQUE))
        iq_dev = torch.from_numpy(np.concatenate([uniq] * (F // N_UNIQUE))).to(dev)
        steps = 50 if wl in ("1843_RadVel", DEFAULT_WORKLOAD) else 20
        ms_total, stage_ms, out = time_stage_loop(torch, pipe, iq_dev, steps, 5)
        pts = int(out["offsets"][F].item())
        kernels, bytes_pf, step_bytes = stage_rooflines(cm, wl, F, stage_ms, ms_total / steps, pts, peak, pipe.engine.last_rd_kernel())
        if det_name == "ca_cfar_2d":
            extras.append({"workload": f"configs/{cfg_name}.cfg", "detector": "ca_cfar_2d num_train=(4,4) num_guard=(2,2) pfa=1e-5",
                           "frames_per_step": F, "cfar_ms_per_launch": stage_ms["cfar"],
                           "cfar_frac_of_hbm": [k for k in kernels if k["stage"] == "cfar"][0]["frac"],
                           "points_per_step": pts})
        else:
            extras.append({"workload": f"configs/{cfg_name}.cfg batched synthetic int16 ADC cubes", "frames_per_step": F,
                           "steps": steps, "value": F * steps / (ms_total * 1e-3), "unit": "frames/s",
                           "ms_per_step": ms_total / steps, "stage_ms_per_step": stage_ms,
                           "roofline": {"kernel": kernels[0]["kernel"], "frac": kernels[0]["frac"],
                                        "achieved": kernels[0]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                                        "algorithmic_bytes_per_frame": bytes_pf["range_doppler"], "kernels": kernels,
                                        "step_hbm_frac": step_bytes * F / (ms_total / steps * 1e-3) / 1e9 / peak},
                           "points_per_step": pts})
        del pipe, iq_dev, out
        torch.cuda.empty_cache()
    # BASELINE.json configs[2]: high-resolution range-azimuth maps (RadarHD, all 8 antennas)
    from mmwave_radar_processing_b200.engine import RadarEngine
    cm = load_cfg("RadarHD", "standard")
    eng = RadarEngine(cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops, device=dev)
    F = 2048
    iq_dev = torch.from_numpy(np.concatenate([synth_batch_for_cfg(cm, range(64))] * (F // 64))).to(dev)
    for bins in (256, 1024):
        out = torch.empty((F, eng.S, bins), dtype=torch.float32, device=dev)
        for _ in range(5):
            eng.range_angle(iq_dev, 0, None, bins, out=out)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            eng.range_angle(iq_dev, 0, None, bins, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        alg = 4 * eng.R * eng.S * eng.cpl + 4 * eng.S * bins
        extras.append({"workload": f"configs/RadarHD.cfg range-azimuth map, {bins} angle bins, {eng.V} antennas", "kernel": "range_angle_kernel",
                       "frames_per_step": F, "ms_per_launch": ms, "value": F / ms * 1e3, "unit": "frames/s",
                       "algorithmic_bytes_per_frame": alg, "frac_of_hbm": alg * F / ms / 1e6 / peak})
        del out
    del iq_dev
    torch.cuda.empty_cache()
    # Doppler beam sharpening (K5b) on RadSAR_1843_official: the reference-facing class picks the bins (fp64 argmins on the
    # host), radar_dbs does the rest
    from mmwave_radar_processing_b200.processors.range_angle_resp_dbs_enhanced import RangeAngleProcessorDBSEnhanced
    cm = load_cfg("RadSAR_1843_official", "standard")
    proc = RangeAngleProcessorDBSEnhanced(cm, num_angle_bins_dbs_enhanced_response=100)
    vbin, abin = proc.select_bins(np.array([0.5, 0.1, 0.0]))
    eng = RadarEngine(cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops, device=dev)
    F = 1024
    iq_dev = torch.from_numpy(np.concatenate([synth_batch_for_cfg(cm, range(32))] * (F // 32))).to(dev)
    out = torch.empty((F, eng.S, len(vbin)), dtype=torch.float32, device=dev)
    vl, al = [int(v) for v in vbin], [int(v) for v in abin]
    for _ in range(5):
        eng.dbs(iq_dev, vl, al, ant_list=None, num_angle_bins=int(proc.num_angle_bins), out=out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        eng.dbs(iq_dev, vl, al, ant_list=None, num_angle_bins=int(proc.num_angle_bins), out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    alg = 4 * eng.R * eng.S * eng.C + 4 * eng.S * len(vl)
    extras.append({"workload": f"configs/RadSAR_1843_official.cfg Doppler beam sharpening, {len(vl)} output angles, velocity (0.5, 0.1, 0), all {eng.V} antennas",
                   "kernel": "rd_reg_dbs_kernel", "frames_per_step": F, "ms_per_launch": ms, "value": F / ms * 1e3, "unit": "frames/s",
                   "algorithmic_bytes_per_frame": alg, "frac_of_hbm": alg * F / ms / 1e6 / peak})
    del iq_dev, out
    torch.cuda.empty_cache()
    # the synthetic-array image (K5a), BASELINE.json configs[3] shape: the whole line is `bench.py --workload sar`
    s_, n_, g_ = 90, 2000, 512 * 64
    rng = np.random.default_rng(0)
    x = (rng.integers(-2000, 2000, (s_, n_)) + 1j * rng.integers(-2000, 2000, (s_, n_))).astype(np.complex64)
    t = np.arange(n_) * 1e-4
    pos = np.stack([0.01 * t, 0.5 * t, np.zeros(n_)])
    az = np.repeat(np.linspace(-1.0, 1.0, 512), 64)
    el = np.tile(np.linspace(-0.3, 0.3, 64), 512)
    dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
    eng = RadarEngine(1, s_, 1, 1, device=dev)
    xs, ps, ds = torch.from_numpy(x).to(dev), torch.from_numpy(pos).to(dev), torch.from_numpy(dirs).to(dev)
    img = torch.empty((s_, g_), dtype=torch.complex64, device=dev)
    for _ in range(5):
        eng.sar_image(xs, ps, ds, 0.0039, out=img)
    e0.record()
    for _ in range(20):
        eng.sar_image(xs, ps, ds, 0.0039, out=img)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    bf16 = float(measured_peaks().get("bf16_tflops", 0.0) or 0.0)
    executed = 3 * 2.0 * 192 * (2 * n_) * g_ / (ms * 1e-3) / 1e12
    extras.append({"workload": "configs/RadSAR_1843_official.cfg synthetic-array image: S=90 range bins, N=2000 aperture elements, G=512x64 directions",
                   "kernel": "sar_tc_kernel (+ max / prep / range FFT)", "ms_per_image": ms, "value": 1e3 / ms, "unit": "images/s",
                   "useful_tflops": 8.0 * s_ * n_ * g_ / (ms * 1e-3) / 1e12, "executed_f16_tflops": executed,
                   "frac_of_measured_bf16_peak": executed / bf16 if bf16 else None})
    del xs, ps, ds, img
    torch.cuda.empty_cache()
    return extras


# ------------------------------------------------------------------------------------------------ SAR workload
def run_sar(a):
    """--workload sar: BASELINE.json configs[3] - one synthetic-array image (configs/RadSAR_1843_official.cfg: S = 90 range
    bins, N = 2000 aperture elements, G = 512 x 64 directions), the direction grid sharded over the ranks after a broadcast
    of the aperture cube + geometry; every rank ends with the full image (all-gather of the tiles).  A step = one image."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from mmwave_radar_processing_b200 import distributed as rdist
    from mmwave_radar_processing_b200.engine import RadarEngine

    rank, local_rank, world = rdist.init_from_env()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    s, n, g = 90, 2000, 512 * 64
    rng = np.random.default_rng(0)
    x = (rng.integers(-2000, 2000, (s, n)) + 1j * rng.integers(-2000, 2000, (s, n))).astype(np.complex64)
    t = np.arange(n) * 1e-4
    pos = np.stack([0.01 * t, 0.5 * t, np.zeros(n)])
    az = np.repeat(np.linspace(-1.0, 1.0, 512), 64)
    el = np.tile(np.linspace(-0.3, 0.3, 64), 512)
    dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
    lam = 0.0039
    eng = RadarEngine(1, s, 1, 1, device=dev)
    if rank == 0:
        xs, ps = torch.from_numpy(x).to(dev), torch.from_numpy(pos).to(dev)
    else:
        xs = torch.empty((s, n), dtype=torch.complex64, device=dev)
        ps = torch.empty((3, n), dtype=torch.float64, device=dev)
    ds = torch.from_numpy(dirs).to(dev)

    # N > 1: the exchange is fused into the kernels (distributed.PeerImageExchange: the range-FFT kernel stores its tile
    # into every rank's image over NVLink peer memory, the kernels read rank 0's aperture from its symmetric region);
    # RADAR_B200_SAR_EXCHANGE=nccl keeps the collectives (broadcast + all-gather) for comparison.
    fused = world > 1 and os.environ.get("RADAR_B200_SAR_EXCHANGE", "peer") != "nccl"
    # RADAR_B200_SAR_PIPELINE=0: one image at a time; default: the two image slots run on their own streams and contexts, so the
    # peer stores and barriers of image k overlap the GEMM of image k + 1
    want_pipe = os.environ.get("RADAR_B200_SAR_PIPELINE", "1") != "0"
    pipelined = fused and want_pipe
    ex = rdist.PeerImageExchange(eng, s, n, g, depth=2, src=0, pipelined=pipelined) if fused else None
    # N = 1: the same two slots on one GPU (LocalImagePipeline): the small kernels of one image run under the GEMM of the next
    local = rdist.LocalImagePipeline(eng, g, depth=2) if (world == 1 and want_pipe) else None
    step_no = [0]

    def step():
        if local is not None:
            k = step_no[0]
            step_no[0] += 1
            return local.run(k, xs, ps, ds, lam)
        if ex is not None:
            k = step_no[0]
            step_no[0] += 1
            return ex.run(k, ds, lam, xs if rank == 0 else None, ps if rank == 0 else None)
        # broadcast of the aperture cube + geometry from rank 0, this rank's direction slice, all-gather of the tiles
        img, _ = rdist.sar_image_sharded(lambda ap, po, di, lo, hi: eng.sar_image(ap, po, di, lam, lo, hi),
                                         xs, ps, ds, src=0, gather=True)
        return img

    for _ in range(max(3, a.warmup)):
        img = step()
    torch.cuda.synchronize()
    graphed = False
    if local is not None and os.environ.get("RADAR_B200_SAR_GRAPH", "1") != "0":
        local.capture(xs, ps, ds, lam)
        graphed = True
        for _ in range(4):
            img = step()
        torch.cuda.synchronize()
    if ex is not None and os.environ.get("RADAR_B200_SAR_GRAPH", "1") != "0":
        # one image is ten short launches per rank: replay them as a CUDA graph per image slot so the host does not set the pace
        ok = torch.ones(1, device=dev)
        try:
            replay = ex.capture(ds, lam, xs if rank == 0 else None, ps if rank == 0 else None)
        except Exception as exc:  # noqa: BLE001 - capture is an optimisation; the eager path below is the same work
            replay = None
            ok.zero_()
            print(f"[bench] rank {rank}: graph capture failed ({exc}); eager launches", file=sys.stderr)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok.item()) > 0:
            graphed = True

            def step():  # noqa: F811
                k = step_no[0]
                step_no[0] += 1
                return replay(k)
            for _ in range(3):
                img = step()
            torch.cuda.synchronize()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        img = step()
    if ex is not None or local is not None:
        for kk in range(max(0, step_no[0] - 2), step_no[0]):     # pipelined: the last images are still on their own streams
            (ex or local).wait(kk)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms_total = e0.elapsed_time(e1)
    # kernel-only time of this rank's slice (no collectives, no mirrors)
    if ex is not None:
        ex.disarm()
    lo, hi = rdist.shard_range(g, rank, world)
    kx, kp = ex.inputs(0) if ex is not None else (xs, ps)     # the fused variant: this rank's pulled copy of rank 0's cube
    kt = torch.empty((s, hi - lo), dtype=torch.complex64, device=dev)
    eng.sar_image(kx, kp, ds, lam, lo, hi, out=kt)
    torch.cuda.synchronize()
    try:                                                       # the launches of one image as a graph: device time, not the host's pace
        gk = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gk):
            eng.sar_image(kx, kp, ds, lam, lo, hi, out=kt)
        one_image = gk.replay
    except Exception:  # noqa: BLE001
        one_image = lambda: eng.sar_image(kx, kp, ds, lam, lo, hi, out=kt)   # noqa: E731
    for _ in range(3):
        one_image()
    torch.cuda.synchronize()
    k0.record()
    for _ in range(a.steps):
        one_image()
    k1.record()
    torch.cuda.synchronize()
    k_ms = k0.elapsed_time(k1) / a.steps
    if world > 1:
        tt = torch.tensor([ms_total, k_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total, k_ms = float(tt[0].item()), float(tt[1].item())
    # end to end: host aperture + geometry in, host image out (rank 0 feeds, every rank reads its copy back)
    xh, ph = torch.from_numpy(x).pin_memory(), torch.from_numpy(pos).pin_memory()
    ih = torch.empty((s, g), dtype=torch.complex64).pin_memory()
    e2e_steps = min(a.steps, 10)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(e2e_steps):
        if rank == 0:
            xs.copy_(xh, non_blocking=True)
            ps.copy_(ph, non_blocking=True)
        img = step()
        if ex is not None or local is not None:
            (ex or local).wait(step_no[0] - 1)                   # the image's stream -> the stream of the read-back
        ih.copy_(img, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1)
    if world > 1:
        tt = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item())
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        peaks = measured_peaks()
        bf16 = float(peaks.get("bf16_tflops", 0.0) or 0.0)
        per = ms_total / a.steps
        g_rank = hi - lo
        sp2 = 2 * ((s + 7) // 8 * 8)
        useful = 8.0 * s * n * g_rank / (k_ms * 1e-3) / 1e12
        executed = 3 * 2.0 * sp2 * (2 * n) * ((g_rank + 127) // 128 * 128) / (k_ms * 1e-3) / 1e12
        line = {"metric": "images/s synthetic-array (SAR) beamformer, direction grid sharded by voxel", "value": 1e3 / per,
                "unit": "images/s", "n_gpus": world, "steps": a.steps, "warmup": max(3, a.warmup), "ms_per_step": per,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16x3 (split fp32, hi+lo FP16 pairs) -> f32",
                "data": "synthetic aperture cube (seeded integers), linear platform track",
                "config": {"workload": "configs/RadSAR_1843_official.cfg multi-frame synthetic array: S=90 range bins, N=2000 "
                                       "aperture elements, G=512x64 directions", "S": s, "N": n, "G": g,
                           "parallelism": (f"voxel-sharded x{world}: radar_sar_image(g_begin, g_end) per rank reads rank 0's aperture cube + "
                                           "geometry over NVLink peer memory and its range-FFT kernel stores the tile into every rank's "
                                           "image (symmetric memory, radar_ctx_set_image_mirrors); two device barriers per image, no collective"
                                           + ("; two images in flight (one stream + one context per image slot)" if pipelined else "")
                                           if fused else
                                           "one GPU, two images in flight (one stream + one context per image slot, CUDA-graph replay): the max / prep / "
                                           "range-FFT kernels of one image run under the GEMM of the next" if local is not None else
                                           f"voxel-sharded x{world}: NCCL broadcast of aperture cube + geometry, "
                                           "radar_sar_image(g_begin, g_end) per rank, NCCL all-gather of the image tiles"),
                           "l2_policy": "operands (1.4 MB) are L2-resident by design; output tiles are written once"},
                "roofline": {"kernel": "sar_tc_kernel", "bound": "tensor", "achieved": executed,
                             "peak": bf16 if bf16 else None, "unit": "TFLOP/s",
                             "frac": executed / bf16 if bf16 else None, "traffic": None,
                             "peak_source": "measured dense bf16 peak (kind::f16 runs at the same rate), MEASURED_PEAKS.json",
                             "useful_tflops": useful, "kernel_ms_per_launch": k_ms,
                             "executed_flops_note": "3 tcgen05.mma kind::f16 per product (hi*hi + hi*lo + lo*hi of FP16 pairs), K = 2N, 192 columns"},
                "e2e": {"value": 1e3 * e2e_steps / e2e_ms, "unit": "images/s",
                        "h2d_bytes_per_step": int(x.nbytes + pos.nbytes), "d2h_bytes_per_step": int(s * g * 8),
                        "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                        "api": "distributed.sar_image_sharded over RadarEngine.sar_image (pinned host aperture in, host image out)"},
                "gpu_launches": a.steps * 4, "cuda_graph": graphed, "clocks": clocks}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


# ------------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS) + ["sar"])
    ap.add_argument("--frames", type=int, default=0,
                    help="frames per GPU per step (weak scaling); 0 = the 100k-frame stream split over the GPUs (strong)")
    ap.add_argument("--chunk", type=int, default=512, help="frames per H2D/D2H chunk of the end-to-end path")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 -> min(steps, 5)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the N=1 extra workloads (1843_RadVel, CA-CFAR)")
    ap.add_argument("--no-range-angle", action="store_true")
    ap.add_argument("--exchange", default="copy", choices=["copy", "peer", "nccl"],
                    help="N > 1 exchange of the point clouds: copy = kernel stores into this rank's gathered slot + copy-engine "
                         "pushes to the peers over NVLink (default); peer = the kernel stores into every rank's slot itself; "
                         "nccl = staging copy + NCCL all-gathers")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    if a.impl == "reference":
        if a.workload == "sar":
            a.workload = DEFAULT_WORKLOAD
        return run_reference_arm(a)
    if a.workload == "sar":
        return run_sar(a)

    import numpy as np
    import torch
    import torch.distributed as dist
    from mmwave_radar_processing_b200 import distributed as rdist
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg

    rank, local_rank, world = rdist.init_from_env()
    if world != a.gpus and world > 1:
        a.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)

    cfg_name, geom, az, el, ra_ants = WORKLOADS[a.workload]
    cm = load_cfg(cfg_name, geom)
    R, S, cpl, L = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    F, F_global, scaling = frames_per_rank(a, world, rank)
    if world > 1 and a.frames == 0 and STREAM_FRAMES % world != 0:
        raise SystemExit("the stream must split evenly over the ranks (the gathered buffer has equal slots)")
    pipe = FramePipeline(cm, detector=DetectorSpec(**DETECTOR), az_antenna_idxs=az, el_antenna_idxs=el,
                         max_dets=MAX_DETS, with_range_angle=not a.no_range_angle, ra_antennas=ra_ants, device=dev)

    # synthetic frames: N_UNIQUE distinct seeded frames per rank, tiled to the rank's block ON THE DEVICE
    uniq = synth_batch_for_cfg(cm, range(rank * N_UNIQUE, rank * N_UNIQUE + min(N_UNIQUE, F)))
    reps = (F + uniq.shape[0] - 1) // uniq.shape[0]
    uniq_dev = torch.from_numpy(uniq).to(dev)
    iq_dev = uniq_dev.repeat((reps, 1, 1, 1, 1))[:F].contiguous()
    del uniq_dev
    Fe = min(F, E2E_FRAMES_MAX)
    iq_host, counts_host, points_host = pipe.host_buffers(Fe, points_per_frame=MAX_DETS // 2)
    iq_host.copy_(iq_dev[:Fe])

    # ---- N > 1: the exchange of the point clouds (PeerPointExchange: one symmetric allocation, 3 gathered slots per rank).
    # "copy" (default): the point-cloud kernel stores every packed row into this rank's own slot (radar_ctx_set_point_mirrors)
    # and the slot, the counts and the total go to the peers by copy-engine peer copies over NVLink, then the cross-rank
    # barrier - all on a side stream under the kernels of the next step, no SM time.  "peer": the kernel stores the rows into
    # every rank's slot itself (best at N = 2, loses at N = 8: 32-byte remote stores per thread).  "nccl": staging copy +
    # NCCL all-gathers on the side stream (the NCCL kernels take SMs from the next step's range-Doppler / range-angle).
    # The fixed capacity is sized from a first, synchronous step (largest per-rank total + 25 %) and checked at the end;
    # the copy-engine pushes carry only the rows the previous steps produced (+10 %, rounded up to 256 rows).
    cap_rows = 0
    push_rows = None
    comm = torch.cuda.Stream(device=dev) if world > 1 else None
    staging, gathered, done_ev = [None, None, None], [None, None, None], [None, None, None]
    step_no = 0
    peer_ex = None
    DEPTH = 3                                                    # gathered slots in flight (peer exchange)

    def step(mark=None):
        nonlocal step_no
        if world > 1 and peer_ex is not None:
            k = step_no % DEPTH
            main = torch.cuda.current_stream()
            if step_no >= 2 and done_ev[(step_no - 2) % DEPTH] is not None:
                main.wait_event(done_ev[(step_no - 2) % DEPTH])   # every rank is past step i-2: slot k (step i-3) is free
            peer_ex.arm(k, remote=a.exchange == "peer")
            out = pipe.run(iq_dev, mark)
            ready = torch.cuda.Event()
            ready.record(main)
            with torch.cuda.stream(comm):
                comm.wait_event(ready)
                peer_ex.publish(k, out["counts"], out["offsets"][F:F + 1], push_points=a.exchange == "copy", rows=push_rows)
                done_ev[k] = torch.cuda.Event()
                done_ev[k].record(comm)
            gathered[k] = peer_ex.gathered(k)
            step_no += 1
            return out
        out = pipe.run(iq_dev, mark)
        if world > 1:
            k = step_no & 1
            main = torch.cuda.current_stream()
            if done_ev[k] is not None:
                main.wait_event(done_ev[k])                      # the gather that last read staging[k] has finished
            if staging[k] is None:
                staging[k] = (torch.empty((cap_rows, 4), dtype=torch.float64, device=dev),
                              torch.empty((F,), dtype=torch.int32, device=dev), torch.empty((1,), dtype=torch.int32, device=dev))
            pts, cnt, tot = staging[k]
            pts.copy_(out["points"][:cap_rows]); cnt.copy_(out["counts"]); tot.copy_(out["offsets"][F:F + 1])
            ready = torch.cuda.Event()
            ready.record(main)
            with torch.cuda.stream(comm):
                comm.wait_event(ready)
                gathered[k] = rdist.allgather_point_clouds(pts, cnt, tot, cap_rows, *(gathered[k] or (None, None, None)))
                done_ev[k] = torch.cuda.Event()
                done_ev[k].record(comm)
            step_no += 1
        return out

    def drain():
        if world > 1:
            torch.cuda.current_stream().wait_stream(comm)

    if world > 1:
        first = pipe.run(iq_dev)
        t = first["offsets"][F:F + 1].to(torch.int64).clone()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        largest = int(t.item())
        cap_rows = int((largest * 5 // 4 + 1023) // 1024 * 1024)
        push_rows = min(cap_rows, int((largest * 11 // 10 + 255) // 256 * 256))
        if a.exchange in ("peer", "copy"):
            # symmetric memory needs peer access between all GPUs of the node; every rank must take the same branch
            ok = torch.ones((1,), dtype=torch.int32, device=dev)
            try:
                peer_ex = rdist.PeerPointExchange(pipe.engine, F, cap_rows, depth=DEPTH)
            except Exception as exc:  # noqa: BLE001
                print(f"[bench] rank {rank}: peer exchange unavailable ({exc}); using NCCL all-gathers", file=sys.stderr)
                ok.zero_()
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 0:
                if peer_ex is not None:
                    peer_ex.disarm()
                peer_ex = None
                a.exchange = "nccl"

    stage_names = ["range_doppler"] + ([] if a.no_range_angle else ["range_angle"]) + ["cfar", "point_cloud"]

    def barrier():
        if world > 1:
            dist.barrier()

    for _ in range(a.warmup):
        step()
    drain()
    torch.cuda.synchronize()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()

    # ---- device-resident timed region: exactly K steps, stage events recorded between the launches
    marks = []

    def mark(name):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        marks.append((name, ev))

    barrier()
    torch.cuda.synchronize()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for _ in range(a.steps):
        mark("begin")
        out = step(mark)
    drain()
    t_end.record()
    torch.cuda.synchronize()
    barrier()
    ms_total = t_start.elapsed_time(t_end)
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        for g in gathered:
            if g is not None:
                rdist.check_capacity(g[2], push_rows if (peer_ex is not None and a.exchange == "copy") else cap_rows)
        if peer_ex is not None:
            peer_ex.disarm()                                     # the end-to-end region below is per rank
    stage_ms = {n: 0.0 for n in stage_names}
    for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
        if n1 != "begin":
            stage_ms[n1] += e0.elapsed_time(e1)
    stage_ms = {k: v / a.steps for k, v in stage_ms.items()}
    total_points = int(out["offsets"][F].item())

    # ---- end-to-end timed region through the host-facing call
    e2e_steps = a.e2e_steps or min(a.steps, 5)
    for _ in range(2):
        pipe.run_host(iq_host, counts_host, points_host, chunk_frames=a.chunk)
    barrier()
    torch.cuda.synchronize()
    e_start = torch.cuda.Event(enable_timing=True)
    e_end = torch.cuda.Event(enable_timing=True)
    e_start.record()
    for _ in range(e2e_steps):
        totals, regions, h2d, d2h = pipe.run_host(iq_host, counts_host, points_host, chunk_frames=a.chunk)
    e_end.record()
    torch.cuda.synchronize()
    barrier()
    e2e_ms = e_start.elapsed_time(e_end)
    if world > 1:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    # plain pinned-host -> device copy of the same bytes, every rank at once: the ceiling of the end-to-end number
    h2d_dst = pipe._chunks[0].iq
    barrier()
    torch.cuda.synchronize()
    e_start.record()
    for _ in range(e2e_steps):
        for lo in range(0, Fe, h2d_dst.shape[0]):
            n = min(h2d_dst.shape[0], Fe - lo)
            h2d_dst[:n].copy_(iq_host[lo:lo + n], non_blocking=True)
    e_end.record()
    torch.cuda.synchronize()
    barrier()
    h2d_ms = e_start.elapsed_time(e_end)
    if world > 1:
        t = torch.tensor([h2d_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        h2d_ms = float(t.item())
    clocks = sampler.stop() if sampler else None

    if rank == 0:
        ms_per_step = ms_total / a.steps
        value = F_global * a.steps / (ms_total * 1e-3)
        peak, peak_src = hbm_peak()
        rd_kernel = pipe.engine.last_rd_kernel()
        kernels, bytes_pf, step_bytes = stage_rooflines(cm, a.workload, F, stage_ms, ms_per_step, total_points, peak, rd_kernel)
        rd_ms = stage_ms["range_doppler"]
        achieved = bytes_pf["range_doppler"] * F / (rd_ms * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(rd_kernel, a.workload)
        e2e_value = world * Fe * e2e_steps / (e2e_ms * 1e-3)
        h2d_gbs = world * h2d * e2e_steps / (h2d_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f32", "data": f"synthetic ({min(N_UNIQUE, F)} distinct seeded frames per rank tiled to the rank's block)",
            "config": workload_config(cm, a.workload, F, F_global, world, a.exchange),
            "roofline": {"kernel": rd_kernel, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": (traffic * F if traffic else None), "traffic_source": traffic_src,
                         "peak_source": peak_src, "algorithmic_bytes_per_frame": bytes_pf["range_doppler"],
                         "kernel_ms_per_launch": rd_ms, "kernels": kernels,
                         "step_algorithmic_bytes_per_frame": step_bytes,
                         "step_hbm_frac": step_bytes * F / (ms_per_step * 1e-3) / 1e9 / peak},
            "stage_ms_per_step": stage_ms,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                    "frames_per_gpu_per_step": Fe, "chunk_frames": a.chunk,
                    "api": "FramePipeline.run_host (pinned host int16 -> host point clouds)",
                    "h2d_ceiling": {"aggregate_gbs": h2d_gbs, "frames_per_s": world * Fe * e2e_steps / (h2d_ms * 1e-3),
                                    "frac": e2e_value / (world * Fe * e2e_steps / (h2d_ms * 1e-3)),
                                    "what": "the same pinned host buffers copied to the device by every rank at once, no kernels"}},
            "gpu_launches": a.steps * pipe.launches_per_batch,
            "points_per_step_rank0": total_points,
            "clocks": clocks,
        }
        if world > 1:
            line["exchange"] = {"variant": a.exchange, "capacity_rows": cap_rows,
                                "pushed_rows_per_peer": push_rows if a.exchange == "copy" else None,
                                "bytes_pushed_per_rank_per_step": (push_rows * 32 + 4 * F + 4) * (world - 1) if a.exchange == "copy" else None}
        if world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_single(cm, a.workload, uniq[:64])
        if world == 1 and not a.no_extra:
            del iq_dev, out
            pipe._main = None
            pipe._chunks = []
            torch.cuda.empty_cache()
            try:
                line["extra"] = extra_single_gpu_lines(torch, dev, peak)
            except Exception as exc:  # noqa: BLE001
                line["extra"] = [{"error": f"{type(exc).__name__}: {exc}"}]
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
