"""Base class of the Range-Doppler detectors (reference: processors/range_doppler_detection/
range_doppler_detector.py:11-116).

``process`` forms the complex RD cube of ALL antennas (``rng_dop_resp_raw``) and the magnitude of virtual
antenna 0 (``rng_dop_resp`` - the only map that drives CFAR, ref :72-80) in ONE launch of the fused
kernel, then runs the detection strategy on the device.  The device copies are kept
(``_rd_dev``, ``_dets_dev``, ``_counts_dev``) so PointCloudGenerator can chain without a host round trip.
"""
from abc import ABC, abstractmethod
from typing import Optional, Tuple

import numpy as np

from ...config_managers.cfgManager import ConfigManager
from ...logging.logger import get_logger
from .._device import stage_cube, to_host_c128, to_host_f64
from ..range_doppler_resp import RangeDopplerProcessor


class RangeDopplerDetector(RangeDopplerProcessor, ABC):

    def __init__(self, config_manager: ConfigManager, **kwargs):
        super().__init__(config_manager)
        self.logger = get_logger(__name__)
        self.rng_dop_resp_raw: Optional[np.ndarray] = None
        self.rng_dop_resp: Optional[np.ndarray] = None
        self.dets: Optional[np.ndarray] = None
        self._eng = None
        self._adc_dev = None
        self._rd_dev = None
        self._mag_dev = None
        self._dets_dev = None
        self._counts_dev = None
        self._host_mode = True

    def configure(self):
        super().configure()

    def reset(self):
        super().reset()
        self.rng_dop_resp_raw = None
        self.rng_dop_resp = None
        self.dets = None
        self._rd_dev = self._mag_dev = self._dets_dev = self._counts_dev = None

    def process(self, adc_cube, **kwargs):
        self._compute_range_doppler_response(adc_cube)
        self.dets = self._detect(adc_cube, self.rng_dop_resp, **kwargs)
        return self.dets

    def _compute_range_doppler_response(self, adc_cube) -> Tuple[np.ndarray, np.ndarray]:
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        self._eng, self._adc_dev, self._host_mode = eng, dev, squeeze
        self._rd_dev, self._mag_dev = eng.range_doppler(dev, want_rd=True, mag_antenna=0)
        if squeeze:
            self.rng_dop_resp_raw = to_host_c128(self._rd_dev[0])
            self.rng_dop_resp = to_host_f64(self._mag_dev[0])
        else:
            self.rng_dop_resp_raw, self.rng_dop_resp = self._rd_dev, self._mag_dev
        return self.rng_dop_resp_raw, self.rng_dop_resp

    def _finish_dets(self, res):
        """Keep the device lists; in host mode return the reference's (N,2) int array."""
        self._dets_dev, self._counts_dev = res["dets"], res["counts"]
        if not self._host_mode:
            return self._dets_dev
        n = int(self._counts_dev[0].item())
        if n == 0:
            return np.empty((0, 2), dtype=int)
        return self._dets_dev[0, :n].cpu().numpy().astype(int)

    @abstractmethod
    def _detect(self, adc_cube, rng_dop_resp, **kwargs):
        pass

    def _map_detections_to_bins(self, dets: np.ndarray):
        if dets is None or dets.size == 0:
            return np.array([]), np.array([]), np.array([]), np.array([])
        r_idx = dets[:, 0].astype(int)
        v_idx = dets[:, 1].astype(int)
        return self.range_bins[r_idx], self.vel_bins[v_idx], r_idx, v_idx
