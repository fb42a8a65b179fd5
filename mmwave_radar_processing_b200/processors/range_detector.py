"""Range profile + 1-D CFAR (reference: processors/range_detector.py:11-97)."""
from typing import Optional

import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ..detectors.detector_registry import get_detector_registry
from .range_resp import RangeProcessor


class RangeDetector(RangeProcessor):

    def __init__(self, config_manager: ConfigManager, cfar_type: str = "os_cfar_1d", cfar_params: dict = {}, **kwargs):
        registry = get_detector_registry()
        if cfar_type not in registry:
            raise ValueError(f"Unknown CFAR type: {cfar_type}. Available: {list(registry.keys())}")
        self.cfar_detector = registry[cfar_type](**cfar_params)
        self.dets: Optional[np.ndarray] = None
        self.thresholds: Optional[np.ndarray] = None
        self.range_resp: Optional[np.ndarray] = None
        super().__init__(config_manager)
        self.logger.info(f"RangeDetector initialized with CFAR type: {cfar_type}")

    def process(self, adc_cube, **kwargs) -> np.ndarray:
        self.range_resp = super().process(adc_cube, chirp_idx=0)
        idxs = self.cfar_detector.detect(self.range_resp)
        self.thresholds = self.cfar_detector.thresholds
        self.dets = np.array(idxs) if len(idxs) > 0 else np.array([])
        return self.dets

    def _map_detections_to_bins(self, dets: np.ndarray) -> np.ndarray:
        if self.range_bins is None:
            raise ValueError("Range bins are not configured.")
        return self.range_bins[dets]
