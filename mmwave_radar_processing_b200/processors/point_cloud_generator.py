"""ADC cube -> (x, y, z, velocity) point cloud (reference: processors/point_cloud_generator.py:10-256).

detector (RD + CFAR) -> per-detection zero-padded angle FFT over the azimuth / elevation antenna sets
-> first-max argmax -> ``angle_bins[idx]`` -> FLU cartesian.  The whole chain stays on the device; the
host path downloads the ``(N, 4)`` fp64 array the reference returns.
"""
from typing import Dict, List, Union

import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ..logging.logger import get_logger
from ._device import engine_for_cube_shape
from ._processor import _Processor
from .range_doppler_detection.registry import get_range_doppler_detector_registry


def _as_index_array(value, name):
    if value is None:
        return np.array([], dtype=int)
    if isinstance(value, list):
        return np.array(value, dtype=int)
    if isinstance(value, np.ndarray):
        return value.astype(int)
    raise ValueError(f"{name} must be a list or numpy array")


class PointCloudGenerator(_Processor):

    def __init__(self, config_manager: ConfigManager, az_antenna_idxs: Union[List[int], np.ndarray],
                 el_antenna_idxs: Union[List[int], np.ndarray], detector_type: str = "range_doppler_detector_2d",
                 detector_params: Dict = {}, shift_az_resp: bool = True, shift_el_resp: bool = False,
                 num_angle_bins: int = 64, **kwargs):
        self.logger = get_logger(__name__)
        self.shift_az_resp = shift_az_resp
        self.shift_el_resp = shift_el_resp
        self.az_antenna_idxs = _as_index_array(az_antenna_idxs, "az_antenna_idxs")
        self.el_antenna_idxs = _as_index_array(el_antenna_idxs, "el_antenna_idxs")
        self.num_angle_bins = num_angle_bins
        self.phase_shifts: np.ndarray = None
        self.angle_bins: np.ndarray = None
        registry = get_range_doppler_detector_registry()
        if detector_type not in registry:
            raise ValueError(f"Unknown detector type: {detector_type}. Available: {list(registry.keys())}")
        self.detector = registry[detector_type](config_manager, **detector_params)
        self.logger.info(f"PointCloudGenerator initialized with detector: {detector_type}")
        super().__init__(config_manager)

    def configure(self):
        self.detector.configure()
        a = self.num_angle_bins
        # identical numpy expression to the reference: the axis has A or A+1 entries depending on rounding (ref :93-105)
        self.phase_shifts = np.arange(start=np.pi, stop=-np.pi - 2 * np.pi / (a - 1), step=-2 * np.pi / (a - 1))
        self.phase_shifts[-1] = -1 * np.pi
        with np.errstate(invalid="ignore"):
            self.angle_bins = np.arcsin(self.phase_shifts / np.pi)

    # ------------------------------------------------------------------ device chain
    def _device_points(self, eng, rd_dev, dets_dev, counts_dev):
        eng.set_axes(self.detector.range_bins, self.detector.vel_bins)
        eng.set_angle_bins(self.angle_bins)
        return eng.point_cloud(rd_dev, dets_dev, counts_dev, self.az_antenna_idxs.tolist(),
                               self.el_antenna_idxs.tolist(), shift_az=self.shift_az_resp,
                               shift_el=self.shift_el_resp, num_angle_bins=int(self.num_angle_bins))

    def process(self, adc_cube, **kwargs):
        dets = self.detector.process(adc_cube, **kwargs)
        det = self.detector
        if det._host_mode:
            if dets.shape[0] == 0:
                return np.empty((0, 4))
            pts, self._az_bin_dev, self._el_bin_dev = self._device_points(det._eng, det._rd_dev, det._dets_dev, det._counts_dev)
            return pts[0, :dets.shape[0]].cpu().numpy()
        # device batch: (points [F,max_dets,4] f64, counts [F]) stay on the device
        pts, self._az_bin_dev, self._el_bin_dev = self._device_points(det._eng, det._rd_dev, det._dets_dev, det._counts_dev)
        return pts, det._counts_dev

    def _compute_angle_estimation(self, rng_dop_resp_raw: np.ndarray, det_range_idxs: np.ndarray,
                                  det_velocity_idxs: np.ndarray):
        """Angles of the given detections (reference :143-214), evaluated by the device kernel."""
        import torch
        raw = np.asarray(rng_dop_resp_raw)
        n = len(det_range_idxs)
        if n == 0:
            return np.zeros(0), np.zeros(0)
        v, s, l = raw.shape
        # numpy fancy indexing semantics of the reference's gather (:180-186): negative indices wrap, out-of-range ones raise
        ri, di = np.asarray(det_range_idxs).astype(np.int64), np.asarray(det_velocity_idxs).astype(np.int64)
        if ri.shape != di.shape:
            raise IndexError("shape mismatch: range and velocity index arrays differ in length")
        for idx, size, axis in ((ri, s, 1), (di, l, 2)):
            if idx.size and (idx.min() < -size or idx.max() >= size):
                bad = int(idx[(idx < -size) | (idx >= size)][0])
                raise IndexError(f"index {bad} is out of bounds for axis {axis} with size {size}")
        det_range_idxs, det_velocity_idxs = np.where(ri < 0, ri + s, ri), np.where(di < 0, di + l, di)
        eng = engine_for_cube_shape(v, s, l)
        eng.set_axes(np.zeros(s), np.zeros(l))
        eng.set_angle_bins(self.angle_bins)
        rd = torch.from_numpy(np.ascontiguousarray(raw, dtype=np.complex64)[None]).to(eng.device)
        dets = torch.from_numpy(np.stack((np.asarray(det_range_idxs), np.asarray(det_velocity_idxs)), axis=1)
                                .astype(np.int32)[None]).to(eng.device).contiguous()
        counts = torch.tensor([n], dtype=torch.int32, device=eng.device)
        _, az_bin, el_bin = eng.point_cloud(rd, dets, counts, self.az_antenna_idxs.tolist(), self.el_antenna_idxs.tolist(),
                                            shift_az=self.shift_az_resp, shift_el=self.shift_el_resp,
                                            num_angle_bins=int(self.num_angle_bins))
        az = self.angle_bins[az_bin[0].cpu().numpy()] if self.az_antenna_idxs.size else np.zeros(n)
        el = self.angle_bins[el_bin[0].cpu().numpy()] if self.el_antenna_idxs.size else np.zeros(n)
        return az, el

    def _convert_to_cartesian(self, ranges, az_angles, el_angles, velocities) -> np.ndarray:
        x = ranges * np.cos(el_angles) * np.cos(az_angles)
        y = ranges * np.cos(el_angles) * np.sin(az_angles)
        z = ranges * np.sin(el_angles)
        return np.column_stack((x, y, z, velocities))

    def reset(self):
        self.detector.reset()
        return super().reset()
