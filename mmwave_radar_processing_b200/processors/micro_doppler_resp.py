"""Micro-Doppler spectrogram (reference: processors/micro_doppler_resp.py:6-114) on the CUDA kernels.

Per frame: UN-windowed range-Doppler magnitude of one antenna (``radar_range_doppler`` with window_flags = 0), maximum
over the range rows inside ``target_ranges`` (``radar_row_reduce``), pushed as the newest column of a
(velocity x frames) history.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import stage_cube, to_host_f64
from ._processor import _Processor


class MicroDopplerProcessor(_Processor):

    def __init__(self, config_manager: ConfigManager, target_ranges=[0, 1.0], num_frames_history: int = 20,
                 **kwargs) -> None:
        if isinstance(target_ranges, list):
            target_ranges = np.array(target_ranges)
        self.vel_bins: np.ndarray = None
        self.range_bins: np.ndarray = None
        self.target_ranges = target_ranges
        self.range_bin_idxs_to_keep: np.ndarray = None
        self.num_frames_history = num_frames_history
        self.time_bins: np.ndarray = None
        self.micro_doppler_resp: np.ndarray = None
        super().__init__(config_manager)

    def reset(self):
        self.micro_doppler_resp = np.zeros(shape=(self.vel_bins.shape[0], self.num_frames_history))
        super().reset()

    def configure(self):
        cm = self.config_manager
        self.vel_bins = np.arange(start=-1 * cm.vel_max_m_s, stop=cm.vel_max_m_s - cm.vel_res_m_s + 1e-3,
                                  step=cm.vel_res_m_s)
        self.range_bins = np.arange(start=0, step=cm.range_res_m, stop=cm.range_max_m - cm.range_res_m / 2 + 1e-3)
        self.range_bin_idxs_to_keep = np.logical_and(self.range_bins >= self.target_ranges[0],
                                                     self.range_bins <= self.target_ranges[1]).astype(np.bool_)
        self.micro_doppler_resp = np.zeros(shape=(self.vel_bins.shape[0], self.num_frames_history))
        frame_period = cm.frameCfg_periodicity_ms * 1e-3
        self.time_bins = np.linspace(start=0, stop=self.num_frames_history * frame_period, num=self.num_frames_history)

    def process(self, adc_cube, rx_idx=0, **kwargs) -> np.ndarray:
        """Host cube in: the updated (velocity x history) array, newest frame in column 0, as the reference.
        Device batch in (extension): the per-frame columns as a float32 [F, L] tensor; the history is not touched."""
        if not np.any(self.range_bin_idxs_to_keep):
            raise ValueError("zero-size array to reduction operation maximum which has no identity")
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        _, mag = eng.range_doppler(dev, window_flags=0, want_rd=False, mag_antenna=int(rx_idx) % eng.V)
        cols = eng.row_reduce(mag, row_select=self.range_bin_idxs_to_keep, op="max")
        if not squeeze:
            return cols
        self.micro_doppler_resp[:, 1:] = self.micro_doppler_resp[:, 0:-1]
        self.micro_doppler_resp[:, 0] = to_host_f64(cols[0])
        return self.micro_doppler_resp
