"""Range profiles (reference: processors/range_resp.py:8-164).

``coarse_fft`` (one chirp, Hann over samples, |FFT|, mean over antennas) is ``radar_range_profile``; ``zoom_fft``
(scipy ZoomFFT in the reference) is ``radar_range_zoom`` - the direct evaluation of the same sum; ``find_peaks``
(scipy ``find_peaks(prominence=6)`` + 20 dB gate + strongest-N) is ``radar_find_peaks``.  All three run on the device;
host numpy inputs are staged through the compatibility path like every other processor.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import is_device_tensor, stage_cube, to_host_f64
from ._processor import _Processor


def _to_numpy(x):
    return x.detach().cpu().numpy() if is_device_tensor(x) else x


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

    def zoom_fft(self, adc_cube, range_start_m, range_stop_m, chirp_idx: int = 0):
        """-> (zoom magnitude averaged over antennas, zoom range bins).  Host cube in: fp64 (S,) arrays out, as the
        reference.  Device batch in (extension): float32 [F, S] tensor out, ``range_start_m`` / ``range_stop_m`` may be
        per-frame sequences and the bins are returned as a float64 [F, S] numpy array."""
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        loop = int(chirp_idx) % eng.L if int(chirp_idx) < 0 else int(chirp_idx)
        zoom = eng.range_zoom(dev, range_start_m, range_stop_m, self.config_manager.range_max_m, loop_idx=loop)
        if squeeze:
            return to_host_f64(zoom[0]), np.linspace(float(range_start_m), float(range_stop_m), eng.S)
        lo = np.broadcast_to(np.asarray(_to_numpy(range_start_m), dtype=np.float64).reshape(-1), (dev.shape[0],))
        hi = np.broadcast_to(np.asarray(_to_numpy(range_stop_m), dtype=np.float64).reshape(-1), (dev.shape[0],))
        return zoom, np.linspace(lo, hi, eng.S, axis=1)

    def find_peaks(self, rng_resp_db, rng_bins, max_peaks: int = 3, threshold_dB: int = 20):
        """Strongest prominent peaks of a range response in dB -> (range values, dB values), strongest first."""
        import torch
        eng = self._any_engine()
        if is_device_tensor(rng_resp_db):
            prof = rng_resp_db.to(device=eng.device, dtype=torch.float32)
        else:
            prof = torch.from_numpy(np.ascontiguousarray(rng_resp_db, dtype=np.float32)).to(eng.device)
        batch = prof.dim() == 2
        idx, db, cnt = eng.find_peaks(prof.reshape(-1, prof.shape[-1]).contiguous(), in_db=True, prominence_db=6.0,
                                      threshold_db=float(threshold_dB), max_peaks=int(max_peaks))
        if batch:
            return idx, db, cnt
        n = int(cnt[0].item())
        if n == 0:
            return np.array([]), np.array([])
        sel = idx[0, :n].cpu().numpy()
        return np.asarray(rng_bins)[sel], np.asarray(_to_numpy(rng_resp_db), dtype=np.float64)[sel]

    def _any_engine(self):
        from ..engine import RadarEngine
        return RadarEngine.for_config(self.config_manager)

    def process(self, adc_cube, chirp_idx: int = 0, **kwargs):
        return self.coarse_fft(adc_cube, chirp_idx)
