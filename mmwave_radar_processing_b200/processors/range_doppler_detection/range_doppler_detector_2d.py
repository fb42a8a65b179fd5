"""RD map + 2-D CFAR (reference: processors/range_doppler_detection/range_doppler_detector_2d.py:12-66)."""
from typing import Dict

import numpy as np

from ...config_managers.cfgManager import ConfigManager
from ...detectors.detector_registry import get_detector_registry
from .range_doppler_detector import RangeDopplerDetector


class RangeDopplerDetector2D(RangeDopplerDetector):

    def __init__(self, config_manager: ConfigManager, cfar_type: str = "ca_cfar_2d", cfar_params: Dict = {}, **kwargs):
        super().__init__(config_manager, **kwargs)
        registry = get_detector_registry()
        if cfar_type not in registry:
            raise ValueError(f"Unknown CFAR type: {cfar_type}. Available: {list(registry.keys())}")
        self.detector = registry[cfar_type](**cfar_params)
        self.logger.info(f"RangeDopplerDetector initialized with {cfar_type} and params {cfar_params}")

    def _detect(self, adc_cube, rng_dop_resp, **kwargs):
        det = self.detector
        if not hasattr(det, "num_train") or not isinstance(det.num_train, (tuple, list, np.ndarray)):
            raise ValueError("Input X must be a 2D array.")      # a 1-D detector on a 2-D map (reference raises too)
        params, alpha = det._c_params()
        res = self._eng.cfar(self._mag_dev, params, one_d=False, want_noise=self._host_mode, want_mask=self._host_mode)
        if self._host_mode:
            rows, cols = self._mag_dev.shape[1:]
            tr, td, gr, gd = det._window()
            cr, cd = tr + gr, td + gd
            noise = res["noise"][0].cpu().numpy()
            thr = np.full((rows, cols), np.inf)
            if rows - 2 * cr > 0 and cols - 2 * cd > 0:
                thr[cr:rows - cr, cd:cols - cd] = alpha * noise[cr:rows - cr, cd:cols - cd]
            det.thresholds, det.noise_estimates = thr, noise
            det.detections = res["mask"][0].cpu().numpy().astype(bool)
        return self._finish_dets(res)
