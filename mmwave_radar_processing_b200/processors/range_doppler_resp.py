"""Range-Doppler response (reference: processors/range_doppler_resp.py:8-110) on the fused CUDA kernel.

Hann(S) x Hann(L), unnormalised forward 2-D DFT over (samples, chirps), ``fftshift`` on the Doppler axis
only; ``rx_idx >= 0`` selects one antenna, ``-1`` returns all; magnitude or complex.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import stage_cube, to_host_c128, to_host_f64
from ._processor import _Processor


class RangeDopplerProcessor(_Processor):

    def __init__(self, config_manager: ConfigManager, **kwargs) -> None:
        self.vel_bins: np.ndarray = None
        self.range_bins: np.ndarray = None
        super().__init__(config_manager)

    def configure(self):
        cm = self.config_manager
        # keep the reference's numpy expressions: float-arange lengths must match (ref :37-47)
        self.vel_bins = np.arange(start=-1 * cm.vel_max_m_s, stop=cm.vel_max_m_s - cm.vel_res_m_s + 1e-3,
                                  step=cm.vel_res_m_s)
        self.range_bins = np.arange(start=0, step=cm.range_res_m, stop=cm.range_max_m - cm.range_res_m / 2 + 1e-3)

    def process(self, adc_cube, rx_idx: int = 0, return_magnitude: bool = True, **kwargs):
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        rx_idx = int(rx_idx)
        if return_magnitude:
            _, mag = eng.range_doppler(dev, want_rd=False, mag_antenna=rx_idx if rx_idx >= 0 else -1)
            out = mag
        else:
            rd, _ = eng.range_doppler(dev, ant_list=[rx_idx] if rx_idx >= 0 else None, want_rd=True)
            out = rd[:, 0] if rx_idx >= 0 else rd
        if not squeeze:
            return out
        return to_host_f64(out[0]) if return_magnitude else to_host_c128(out[0])
