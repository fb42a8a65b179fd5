"""Coarse range profile (reference: processors/range_resp.py:8-164).

``coarse_fft`` (one chirp, Hann over samples, |FFT|, mean over antennas) runs on the device
(``radar_range_profile``).  ``zoom_fft`` / ``find_peaks`` are SURVEY section 8f "next-1": they act on one
63..512-sample profile and stay host-side (scipy) until the chirp-Z kernel lands; they are not on the
measured hot path.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import stage_cube, to_host_f64
from ._processor import _Processor


class RangeProcessor(_Processor):

    def __init__(self, config_manager: ConfigManager, **kwargs):
        self.num_range_bins = None
        self.range_bins = None
        super().__init__(config_manager)

    def configure(self):
        cm = self.config_manager
        self.num_range_bins = cm.get_num_adc_samples(profile_idx=0)
        self.range_bins = np.arange(start=0, step=cm.range_res_m, stop=cm.range_max_m - cm.range_res_m / 2)

    def coarse_fft(self, adc_cube, chirp_idx: int = 0):
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        loop = int(chirp_idx) % eng.L if int(chirp_idx) < 0 else int(chirp_idx)
        prof = eng.range_profile(dev, loop_idx=loop)
        self._profile_dev = prof
        if not squeeze:
            return prof
        return to_host_f64(prof[0])

    def zoom_fft(self, adc_cube: np.ndarray, range_start_m: float, range_stop_m: float, chirp_idx: int = 0):
        from scipy.signal import ZoomFFT          # next-1: host side
        x = np.asarray(adc_cube)[:, :, chirp_idx]
        n = x.shape[1]
        xw = x * np.hanning(n)
        fs = 1 / self.config_manager.range_res_m
        f0 = range_start_m * fs / self.config_manager.range_max_m
        f1 = range_stop_m * fs / self.config_manager.range_max_m
        zoom = ZoomFFT(n, [f0, f1], fs=fs)
        avg = np.mean(np.abs(zoom(xw, axis=1)), axis=0)
        return avg, np.linspace(range_start_m, range_stop_m, len(avg))

    def find_peaks(self, rng_resp_db: np.ndarray, rng_bins: np.ndarray, max_peaks: int = 3, threshold_dB: int = 20):
        from scipy.signal import find_peaks        # next-1: host side
        peaks, _ = find_peaks(rng_resp_db, prominence=6)
        if len(peaks) == 0:
            return np.array([]), np.array([])
        vals = rng_resp_db[peaks]
        keep = vals >= (np.max(vals) - threshold_dB)
        peaks, vals = peaks[keep], vals[keep]
        top = peaks[np.argsort(vals)[::-1]][:max_peaks]
        return rng_bins[top], rng_resp_db[top]

    def process(self, adc_cube, chirp_idx: int = 0, **kwargs):
        return self.coarse_fft(adc_cube, chirp_idx)
