"""Batched streaming front-end (our extension of the reference API): int16 ADC frames -> point clouds.

``FramePipeline`` chains the three device stages of ``PointCloudGenerator.process`` for a whole batch of
frames with preallocated buffers:

    radar_range_doppler (K1+K2)  ->  radar_cfar_2d / radar_cfar_1d_rows (K4, K4c)  ->  radar_point_cloud (K3b)

Frames are independent, so a batch is just the leading axis of every buffer and the
multi-GPU path shards that axis (``distributed.py``).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np
import torch

from .engine import RadarEngine, cfar_params


def reference_axes(cfg, num_angle_bins: int = 64):
    """range_bins, vel_bins, angle_bins with the reference's numpy expressions
    (processors/range_doppler_resp.py:37-47, processors/point_cloud_generator.py:93-105)."""
    vel_bins = np.arange(start=-1 * cfg.vel_max_m_s, stop=cfg.vel_max_m_s - cfg.vel_res_m_s + 1e-3, step=cfg.vel_res_m_s)
    range_bins = np.arange(start=0, step=cfg.range_res_m, stop=cfg.range_max_m - cfg.range_res_m / 2 + 1e-3)
    a = num_angle_bins
    phase = np.arange(start=np.pi, stop=-np.pi - 2 * np.pi / (a - 1), step=-2 * np.pi / (a - 1))
    phase[-1] = -1 * np.pi
    with np.errstate(invalid="ignore"):
        angle_bins = np.arcsin(phase / np.pi)
    return range_bins, vel_bins, angle_bins


@dataclass
class DetectorSpec:
    """kind: 'ca' | 'os' (2-D on the RD map of antenna 0); parameters as in the reference's CFAR classes."""
    kind: str = "ca"
    num_train: Sequence[int] = (4, 4)
    num_guard: Sequence[int] = (2, 2)
    pfa: float = 1e-5
    rho: float = 0.75
    alpha: float = 3.5

    def c_params(self):
        if self.kind == "ca":
            return cfar_params("ca", self.num_train, self.num_guard, pfa=self.pfa)
        if self.kind == "os":
            return cfar_params("os", self.num_train, self.num_guard, rho=self.rho, alpha=self.alpha)
        raise ValueError(f"unknown detector kind {self.kind}")


@dataclass
class FramePipeline:
    cfg: object
    detector: DetectorSpec = field(default_factory=DetectorSpec)
    az_antenna_idxs: Sequence[int] = (0, 1, 2, 3)
    el_antenna_idxs: Sequence[int] = ()
    shift_az_resp: bool = True
    shift_el_resp: bool = False
    num_angle_bins: int = 64
    max_dets: int = 256
    device: Optional[object] = None

    def __post_init__(self):
        self.engine = RadarEngine.for_config(self.cfg, device=self.device)
        rb, vb, ab = reference_axes(self.cfg, self.num_angle_bins)
        self.range_bins, self.vel_bins, self.angle_bins = rb, vb, ab
        self.engine.set_axes(rb, vb)
        self.engine.set_angle_bins(ab)
        self._params = self.detector.c_params()
        self._cap = 0
        self.launches_per_batch = 3          # rd_fused_kernel, cfar_kernel, point_cloud_kernel

    def _ensure(self, n_frames: int):
        if n_frames <= self._cap:
            return
        e, dev = self.engine, self.engine.device
        self.rd = torch.empty((n_frames, e.V, e.S, e.L), dtype=torch.complex64, device=dev)
        self.mag = torch.empty((n_frames, e.S, e.L), dtype=torch.float32, device=dev)
        self.dets = torch.zeros((n_frames, self.max_dets, 2), dtype=torch.int32, device=dev)
        self.counts = torch.zeros((n_frames,), dtype=torch.int32, device=dev)
        self.points = torch.zeros((n_frames, self.max_dets, 4), dtype=torch.float64, device=dev)
        self._cap = n_frames

    def run(self, iq_dev: torch.Tensor):
        """iq_dev: int16 [F][chirp][rx][sample][2] on the device -> dict of device tensors (views of the
        pipeline's buffers): points f64 [F,max_dets,4], counts int32 [F], dets int32 [F,max_dets,2]."""
        f = iq_dev.shape[0]
        self._ensure(f)
        e = self.engine
        rd, mag = self.rd[:f], self.mag[:f]
        e.range_doppler(iq_dev, want_rd=True, mag_antenna=0, rd_out=rd, mag_out=mag)
        res = e.cfar(mag, self._params, one_d=False, max_dets=self.max_dets, dets_out=self.dets[:f], count_out=self.counts[:f])
        pts, _, _ = e.point_cloud(rd, res["dets"], res["counts"], list(self.az_antenna_idxs), list(self.el_antenna_idxs),
                                  shift_az=self.shift_az_resp, shift_el=self.shift_el_resp,
                                  num_angle_bins=self.num_angle_bins, points_out=self.points[:f], want_bins=False)
        return dict(points=pts, counts=res["counts"], dets=res["dets"], rd=rd, mag=mag)

    def run_host(self, iq_host: torch.Tensor, points_host: torch.Tensor, counts_host: torch.Tensor):
        """End-to-end step: pinned int16 host batch -> device -> pipeline -> pinned host results.
        Everything is enqueued on the current stream; the caller synchronises."""
        f = iq_host.shape[0]
        if not hasattr(self, "_iq_dev") or self._iq_dev.shape[0] < f:
            self._iq_dev = torch.empty(iq_host.shape, dtype=torch.int16, device=self.engine.device)
        iq_dev = self._iq_dev[:f]
        iq_dev.copy_(iq_host, non_blocking=True)
        out = self.run(iq_dev)
        points_host[:f].copy_(out["points"], non_blocking=True)
        counts_host[:f].copy_(out["counts"], non_blocking=True)
        return out

    def gather_host(self, out) -> list:
        """List of per-frame (N_i, 4) float64 numpy point clouds (what the reference returns per frame)."""
        counts = out["counts"].cpu().numpy()
        pts = out["points"].cpu().numpy()
        return [pts[i, :min(int(c), self.max_dets)].copy() for i, c in enumerate(counts)]
