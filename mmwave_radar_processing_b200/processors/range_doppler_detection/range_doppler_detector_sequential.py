"""Range CFAR on the coarse profile, then Doppler CFAR on the detected range rows
(reference: processors/range_doppler_detection/range_doppler_detector_sequential.py:12-108).

Device flow: range profile kernel -> 1-D CFAR (mask = row selection) -> RD kernel -> 1-D CFAR along
Doppler restricted to the selected rows -> ordered compaction.  The reference's per-row Python list
comprehension disappears; the row-major order of the result is the same."""
from typing import Dict

import numpy as np

from ...config_managers.cfgManager import ConfigManager
from ...detectors.detector_registry import get_detector_registry
from ..range_resp import RangeProcessor
from .range_doppler_detector import RangeDopplerDetector


class RangeDopplerDetectorSequential(RangeDopplerDetector):

    def __init__(self, config_manager: ConfigManager, rng_cfar_type: str = "os_cfar_1d", rng_cfar_params: Dict = {},
                 vel_cfar_type: str = "os_cfar_1d", vel_cfar_params: Dict = {}, **kwargs):
        super().__init__(config_manager, **kwargs)
        registry = get_detector_registry()
        if rng_cfar_type not in registry:
            raise ValueError(f"Unknown CFAR type: {rng_cfar_type}. Available: {list(registry.keys())}")
        self.rng_detector = registry[rng_cfar_type](**rng_cfar_params)
        if vel_cfar_type not in registry:
            raise ValueError(f"Unknown CFAR type: {vel_cfar_type}. Available: {list(registry.keys())}")
        self.vel_detector = registry[vel_cfar_type](**vel_cfar_params)
        self.range_processor = RangeProcessor(config_manager)
        self.logger.info(f"RangeDopplerDetectorSequential initialized with Range CFAR: {rng_cfar_type}, "
                         f"Velocity CFAR: {vel_cfar_type}")

    def _detect(self, adc_cube, rng_dop_resp, **kwargs):
        raise NotImplementedError("RangeDopplerDetectorSequential uses custom process logic.")

    def process(self, adc_cube, **kwargs):
        self._compute_range_doppler_response(adc_cube)
        eng = self._eng
        prof = eng.range_profile(self._adc_dev, loop_idx=0)                      # [F, S]
        rp, _ = self.rng_detector._c_params()
        rsel = eng.cfar(prof[:, None, :].contiguous(), rp, one_d=True, want_mask=True, max_dets=1)["mask"]
        vp, _ = self.vel_detector._c_params()
        res = eng.cfar(self._mag_dev, vp, one_d=True, row_select=rsel[:, 0, :].contiguous())
        self.dets = self._finish_dets(res)
        return self.dets
