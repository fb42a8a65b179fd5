"""Altimeter (reference: processors/altimeter.py:6-139).  SURVEY section 8f next-1.

Everything that touches samples runs on the device: the coarse range profile (``radar_range_profile``), the zoom
profile around the ground peak (``radar_range_zoom``) and the prominence peak picking (``radar_find_peaks``, which also
takes the 20 log10 in fp64).  What stays in Python is the reference's own control flow on at most three numbers per
frame: the ground-peak gate and the altitude recurrence, which is sequential across frames by definition (the zoom
window of frame f depends on the altitude accepted at frame f-1).  For a device batch the coarse peaks of ALL frames are
picked in one launch; only when ``precise_est_enabled`` does every frame need its own small zoom launch."""
from typing import Optional

import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import is_device_tensor
from .range_resp import RangeProcessor


class Altimeter(RangeProcessor):

    def __init__(self, config_manager: ConfigManager, min_altitude_m: float, zoom_search_region_m: float,
                 altitude_search_limit_m: float, range_bias: float = 0.0, **kwargs) -> None:
        super().__init__(config_manager)
        self.min_altitude_m = float(min_altitude_m)
        self.zoom_search_region_m = float(zoom_search_region_m)
        self.altitude_search_limit_m = float(altitude_search_limit_m)
        self.range_bias = float(range_bias)
        self.coarse_fft_data: Optional[np.ndarray] = None
        self.current_altitude_measured_m = self.min_altitude_m
        self.current_altitude_corrected_m = self.min_altitude_m

    def reset(self):
        self.current_altitude_measured_m = self.min_altitude_m
        return super().reset()

    def find_ground_peak(self, detected_peaks_m: np.ndarray):
        if detected_peaks_m.size > 0:
            ok = detected_peaks_m[(detected_peaks_m >= self.min_altitude_m) &
                                  (np.abs(detected_peaks_m - self.current_altitude_measured_m) <= self.altitude_search_limit_m)]
            if ok.size > 0:
                return np.min(ok)
        return -1.0

    def _perform_coarse_fft(self, adc_cube):
        return self.coarse_fft(adc_cube=adc_cube, chirp_idx=0)

    def _peaks_m(self, profiles_dev, bins_rows, max_peaks):
        """profiles_dev float32 [n, S] linear magnitude -> list of n arrays of peak ranges (strongest first)."""
        eng = self._any_engine()
        idx, _, cnt = eng.find_peaks(profiles_dev.contiguous(), in_db=False, prominence_db=6.0, threshold_db=20.0,
                                     max_peaks=max_peaks)
        idx, cnt = idx.cpu().numpy(), cnt.cpu().numpy()
        return [np.asarray(bins_rows[i])[idx[i, :cnt[i]]] for i in range(idx.shape[0])]

    def _refine(self, frame_dev, ground_peak: float) -> float:
        lo = max(1e-6, ground_peak - self.zoom_search_region_m)
        hi = min(np.max(self.range_bins) - 1e-6, ground_peak + self.zoom_search_region_m)
        zoom, _ = self.zoom_fft(frame_dev, lo, hi, chirp_idx=0)           # device batch of one frame -> [1, S]
        peaks = self._peaks_m(zoom, [np.linspace(lo, hi, zoom.shape[1])], max_peaks=2)[0]
        return self.find_ground_peak(detected_peaks_m=peaks) if peaks.size > 0 else -1.0

    def _step(self, peaks_m: np.ndarray, frame_dev, precise_est_enabled: bool):
        if peaks_m.size == 0:
            return self.current_altitude_corrected_m
        ground = self.find_ground_peak(detected_peaks_m=peaks_m)
        if ground < 0:
            return self.current_altitude_corrected_m
        if not precise_est_enabled:
            self.current_altitude_measured_m = ground
            self.current_altitude_corrected_m = ground + self.range_bias
            return self.current_altitude_corrected_m
        refined = self._refine(frame_dev, ground)
        if refined > 0:
            self.current_altitude_measured_m = refined
            self.current_altitude_corrected_m = refined + self.range_bias
        return self.current_altitude_corrected_m

    def process(self, adc_cube, precise_est_enabled: bool = True, **kwargs):
        """Host cube (V,S,L): the altitude after this frame, as the reference.  Device batch (extension): the frames are
        followed in order and a float64 numpy array with the altitude after EVERY frame is returned."""
        from ._device import stage_cube
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        prof = eng.range_profile(dev, loop_idx=0)                          # float32 [F, S]
        self.coarse_fft_data = prof[0].detach().cpu().numpy().astype(np.float64) if squeeze else prof
        peaks = self._peaks_m(prof, [self.range_bins] * prof.shape[0], max_peaks=3)
        alts = np.empty(prof.shape[0], dtype=np.float64)
        for f in range(prof.shape[0]):
            alts[f] = self._step(peaks[f], dev[f:f + 1], precise_est_enabled)
        return float(alts[0]) if squeeze else alts
