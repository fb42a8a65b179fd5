"""Altimeter (reference: processors/altimeter.py:6-139).  SURVEY section 8f next-1.

The coarse range profile comes from the device kernel; the peak picking / ground-peak gating state machine
works on that one profile (S floats) on the host, exactly as the reference orders it.  The optional ZoomFFT
refinement is host-side scipy until the zoom-DFT kernel exists."""
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
        prof = self.coarse_fft(adc_cube=adc_cube, chirp_idx=0)
        if is_device_tensor(prof):          # batched device input: the state machine follows frame 0..F-1 on host
            prof = prof.detach().cpu().numpy().astype(np.float64)
        return prof

    def _get_coarse_peaks(self, coarse_fft: np.ndarray) -> np.ndarray:
        with np.errstate(divide="ignore"):
            db = 20 * np.log10(coarse_fft)
        peaks, _ = self.find_peaks(rng_resp_db=db, rng_bins=self.range_bins, max_peaks=3)
        return peaks

    def _refine_altitude_estimate(self, adc_cube, ground_peak: float) -> float:
        lo = max(1e-6, ground_peak - self.zoom_search_region_m)
        hi = min(np.max(self.range_bins) - 1e-6, ground_peak + self.zoom_search_region_m)
        zoom_avg, zoom_bins = self.zoom_fft(adc_cube=adc_cube, range_start_m=lo, range_stop_m=hi, chirp_idx=0)
        peaks, _ = self.find_peaks(rng_resp_db=20 * np.log10(zoom_avg), rng_bins=zoom_bins, max_peaks=2)
        return self.find_ground_peak(detected_peaks_m=peaks) if peaks.size > 0 else -1.0

    def _step(self, profile: np.ndarray, adc_cube, precise_est_enabled: bool):
        self.coarse_fft_data = profile
        peaks = self._get_coarse_peaks(profile)
        if peaks.size == 0:
            return self.current_altitude_corrected_m
        ground = self.find_ground_peak(detected_peaks_m=peaks)
        if ground < 0:
            return self.current_altitude_corrected_m
        if not precise_est_enabled:
            self.current_altitude_measured_m = ground
            self.current_altitude_corrected_m = ground + self.range_bias
            return self.current_altitude_corrected_m
        refined = self._refine_altitude_estimate(adc_cube, ground)
        if refined > 0:
            self.current_altitude_measured_m = refined
            self.current_altitude_corrected_m = refined + self.range_bias
        return self.current_altitude_corrected_m

    def process(self, adc_cube, precise_est_enabled: bool = True, **kwargs):
        prof = self._perform_coarse_fft(adc_cube)
        if prof.ndim == 2:                   # device batch: altitude after the LAST frame of the batch
            if precise_est_enabled:
                raise NotImplementedError("precise (ZoomFFT) altitude on device batches is SURVEY 8f next-1")
            alt = self.current_altitude_corrected_m
            for row in prof:
                alt = self._step(row, None, False)
            return alt
        return self._step(prof, adc_cube, precise_est_enabled)
