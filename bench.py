#!/usr/bin/env python
"""bench.py - frames/s of the ADC-cube -> point-cloud hot path on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--frames F] [--impl reference]

A *step* is one pass of the hot path over one batch of F synthetic int16 ADC frames per GPU:

    radar_range_doppler (all V antennas, complex RD cube + |RD[0]|)  ->  radar_range_angle (heat map)
      ->  radar_cfar_2d (OS-CFAR-2D + ordered compaction)  ->  radar_det_offsets  ->  radar_point_cloud
      [N > 1: radar_point_cloud also stores every packed row into this rank's slot of a symmetric gathered buffer; the slot,
       the counts and the total are pushed to the peers by copy-engine peer copies over NVLink + a cross-rank barrier on a
       side stream, under the kernels of the next step (--exchange peer: the kernel stores to every rank itself;
       --exchange nccl: staging copy + NCCL all-gathers)]

``value`` is whole-job frames/s with the inputs resident in HBM; ``e2e`` is the same metric through the host-facing
call (``FramePipeline.run_host``): pinned host int16 frames in, host point clouds out, copies inside the timed
region.  ``roofline`` is for the dominant kernel (the fused range-Doppler kernel), ``cpu_baseline`` is the numpy
oracle port timed on this box's host cores (rank 0, N=1 only).  ``--impl reference`` times that CPU path on all
host cores and prints the same line shape.
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
# detector of the reference's PointCloudGenerator integration test (tests/verify_processors.py:121-143)
DETECTOR = dict(kind="os", num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5)
MAX_DETS = 256
N_UNIQUE = 256          # distinct synthetic frames per rank, tiled to the batch
FALLBACK_HBM_GBS = 6650.0


def load_cfg(name, geometry):
    from mmwave_radar_processing_b200.config_managers.cfgManager import ConfigManager
    cm = ConfigManager()
    cm.load_cfg(os.path.join(REPO, "configs", name + ".cfg"), array_geometry=geometry)
    return cm


def hbm_peak():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:  # noqa: BLE001
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel: str, workload: str):
    """dram bytes per launch per frame from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(REPO, "profiles", "traffic.json")) as fh:
            return json.load(fh).get(workload, {}).get(kernel)
    except Exception:  # noqa: BLE001
        return None


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with NVML while the timed regions run."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int, period_s: float = 0.01):
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
                "samples": len(self.samples)}


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
    sample = (f"{frames_per_step} frames/step of {a.workload} through oracle/radar_oracle.py (numpy fp64), "
              f"{cores}-process pool sharded by frame")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(cm, a.workload, frames_per_step, 1),
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_config(cm, wl, frames_per_gpu, n_gpus, exchange="peer"):
    r, s, cpl, l = cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops
    return {"workload": f"configs/{WORKLOADS[wl][0]}.cfg batched synthetic int16 ADC cubes", "num_rx": r,
            "chirps_per_loop": cpl, "virtual_antennas": r * cpl, "samples": s, "loops": l,
            "frames_per_gpu_per_step": frames_per_gpu, "global_frames_per_step": frames_per_gpu * n_gpus,
            "detector": "os_cfar_2d num_train=(8,4) num_guard=(2,1) rho=0.75 alpha=3.5",
            "stages": "range_doppler+range_angle+cfar_2d+det_offsets+point_cloud",
            "parallelism": f"frame-sharded x{n_gpus}" + ({"copy": " + point rows stored into a symmetric gathered buffer by the point-cloud kernel, pushed to the peers by copy-engine peer copies over NVLink + barrier (side stream, under the next step)", "peer": " + point rows stored into every rank's gathered buffer by the point-cloud kernel (NVLink peer memory), counts by peer copies + barrier on a side stream", "nccl": " + NCCL all-gather of point clouds (own stream, under the next step)"}[exchange] if n_gpus > 1 else ""),
            "l2_policy": "per-step working set (int16 in + complex RD out) exceeds the 126 MB L2; no explicit flush"}


# ------------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="1843_RadVel", choices=sorted(WORKLOADS))
    ap.add_argument("--frames", type=int, default=4096, help="frames per GPU per step")
    ap.add_argument("--chunk", type=int, default=512, help="frames per H2D/D2H chunk of the end-to-end path")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 -> min(steps, 10)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-range-angle", action="store_true")
    ap.add_argument("--exchange", default="copy", choices=["copy", "peer", "nccl"],
                    help="N > 1 exchange of the point clouds: copy = kernel stores into this rank's gathered slot + copy-engine "
                         "pushes to the peers over NVLink (default); peer = the kernel stores into every rank's slot itself; "
                         "nccl = staging copy + NCCL all-gathers")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    if a.impl == "reference":
        return run_reference_arm(a)

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
    V, C = R * cpl, cpl * L
    F = a.frames
    pipe = FramePipeline(cm, detector=DetectorSpec(**DETECTOR), az_antenna_idxs=az, el_antenna_idxs=el,
                         max_dets=MAX_DETS, with_range_angle=not a.no_range_angle, ra_antennas=ra_ants, device=dev)

    # synthetic frames: N_UNIQUE distinct seeded frames per rank, tiled to the batch
    uniq = synth_batch_for_cfg(cm, range(rank * N_UNIQUE, rank * N_UNIQUE + min(N_UNIQUE, F)))
    reps = (F + uniq.shape[0] - 1) // uniq.shape[0]
    iq_host_np = np.concatenate([uniq] * reps)[:F]
    iq_host, counts_host, points_host = pipe.host_buffers(F, points_per_frame=MAX_DETS // 2)
    iq_host.copy_(torch.from_numpy(iq_host_np))
    iq_dev = iq_host.to(dev)

    # ---- N > 1: the exchange of the point clouds (PeerPointExchange: one symmetric allocation, 3 gathered slots per rank).
    # "copy" (default): the point-cloud kernel stores every packed row into this rank's own slot (radar_ctx_set_point_mirrors)
    # and the slot, the counts and the total go to the peers by copy-engine peer copies over NVLink, then the cross-rank
    # barrier - all on a side stream under the kernels of the next step, no SM time.  "peer": the kernel stores the rows into
    # every rank's slot itself (best at N = 2, loses at N = 8: 32-byte remote stores per thread).  "nccl": staging copy +
    # NCCL all-gathers on the side stream (the NCCL kernels take SMs from the next step's range-Doppler / range-angle).
    # The fixed capacity is sized from a first, synchronous step (largest per-rank total + 25 %) and checked at the end.
    cap_rows = 0
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
                peer_ex.publish(k, out["counts"], out["offsets"][F:F + 1], push_points=a.exchange == "copy")
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
        cap_rows = int((int(t.item()) * 5 // 4 + 1023) // 1024 * 1024)
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
                rdist.check_capacity(g[2], cap_rows)
        if peer_ex is not None:
            peer_ex.disarm()                                     # the end-to-end region below is per rank
    stage_ms = {n: 0.0 for n in stage_names}
    for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
        if n1 != "begin":
            stage_ms[n1] += e0.elapsed_time(e1)
    stage_ms = {k: v / a.steps for k, v in stage_ms.items()}
    total_points = int(out["offsets"][F].item())

    # ---- end-to-end timed region through the host-facing call
    e2e_steps = a.e2e_steps or min(a.steps, 10)
    for _ in range(3):
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
    clocks = sampler.stop() if sampler else None

    if rank == 0:
        ms_per_step = ms_total / a.steps
        value = world * F * a.steps / (ms_total * 1e-3)
        peak, peak_src = hbm_peak()
        rd_bytes_per_frame = 4 * R * S * C + 8 * V * S * L + 4 * S * L
        rd_ms = stage_ms["range_doppler"]
        achieved = rd_bytes_per_frame * F / (rd_ms * 1e-3) / 1e9
        traffic = ncu_traffic("rd_stream_kernel", a.workload)
        step_bytes_per_frame = rd_bytes_per_frame + 4 * S * L + S * L
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": f"synthetic ({min(N_UNIQUE, F)} distinct seeded frames per rank tiled to the batch)",
            "config": workload_config(cm, a.workload, F, world, a.exchange),
            "roofline": {"kernel": "rd_stream_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": (traffic * F if traffic else None),
                         "peak_source": peak_src, "algorithmic_bytes_per_frame": rd_bytes_per_frame,
                         "kernel_ms_per_launch": rd_ms,
                         "step_hbm_frac": step_bytes_per_frame * F / (ms_per_step * 1e-3) / 1e9 / peak},
            "stage_ms_per_step": stage_ms,
            "e2e": {"value": world * F * e2e_steps / (e2e_ms * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                    "chunk_frames": a.chunk, "api": "FramePipeline.run_host (pinned host int16 -> host point clouds)"},
            "gpu_launches": a.steps * pipe.launches_per_batch,
            "points_per_step_rank0": total_points,
            "clocks": clocks,
        }
        if world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_single(cm, a.workload, uniq[:64])
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
