"""Ground-return detector (reference: processors/range_doppler_detection/range_doppler_ground_detector.py:13-121):
altimeter -> range band [altitude, min(max range, altitude/cos 60deg)] -> 1-D Doppler CFAR per range bin.

The CFAR over the band runs on the device (row-selected 1-D kernel) and so do the altimeter's profiles and peak picking
(processors/altimeter.py); a device batch gets one band per frame from the altitude recurrence."""
from typing import Dict

import numpy as np

from ...config_managers.cfgManager import ConfigManager
from ...detectors.detector_registry import get_detector_registry
from ..altimeter import Altimeter
from .range_doppler_detector import RangeDopplerDetector


class RangeDopplerGroundDetector(RangeDopplerDetector):

    def __init__(self, config_manager: ConfigManager, vel_cfar_type: str = "os_cfar_1d", vel_cfar_params: Dict = {},
                 altimeter_params: Dict = {}, **kwargs):
        super().__init__(config_manager, **kwargs)
        registry = get_detector_registry()
        if vel_cfar_type not in registry:
            raise ValueError(f"Unknown CFAR type: {vel_cfar_type}. Available: {list(registry.keys())}")
        self.vel_detector = registry[vel_cfar_type](**vel_cfar_params)
        self.logger.info(f"RangeDopplerGroundDetector initialized params: {vel_cfar_params}")
        self.altimeter_params = altimeter_params
        self.altimeter = Altimeter(config_manager, **altimeter_params)
        self.logger.info(f"RangeDopplerGroundDetector initialized with Velocity CFAR: {vel_cfar_type}")

    def reset(self):
        self.altimeter.reset()
        return super().reset()

    def _detect(self, adc_cube, rng_dop_resp, **kwargs):
        raise NotImplementedError("RangeDopplerGroundDetector uses custom process logic.")

    def range_band(self, altitude_m: float):
        """Inclusive range-bin band for an altitude (ref :88-103)."""
        lo = int(np.argmin(np.abs(self.range_bins - altitude_m)))
        max_rng = min(np.max(self.range_bins), altitude_m / np.cos(np.deg2rad(60)))
        hi = int(np.argmin(np.abs(self.range_bins - max_rng)))
        return (lo, lo) if hi == lo else (lo, hi)

    def process(self, adc_cube, **kwargs):
        import torch
        altitudes = np.atleast_1d(self.altimeter.process(adc_cube=adc_cube, **self.altimeter_params))
        self._compute_range_doppler_response(adc_cube)
        eng = self._eng
        f, rows = self._mag_dev.shape[0], self._mag_dev.shape[1]
        sel = np.zeros((f, rows), dtype=np.uint8)
        for i in range(f):
            lo, hi = self.range_band(float(altitudes[min(i, altitudes.shape[0] - 1)]))
            if hi >= lo:
                sel[i, lo:hi + 1] = 1
        rsel = torch.from_numpy(sel).to(eng.device)
        vp, _ = self.vel_detector._c_params()
        res = eng.cfar(self._mag_dev, vp, one_d=True, row_select=rsel)
        self.dets = self._finish_dets(res)
        return self.dets
