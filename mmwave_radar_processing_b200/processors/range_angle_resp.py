"""Range-angle heat map (reference: processors/range_angle_resp.py:6-122) on the CUDA kernel.

Hann over samples and over ALL antennas of the cube (applied before the ``rx_antennas`` subset), one
chirp, zero-padded angle FFT, ``fftshift`` on the angle axis, magnitude.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import stage_cube, to_host_f64
from ._processor import _Processor


class RangeAngleProcessor(_Processor):

    def __init__(self, config_manager: ConfigManager, num_angle_bins: int = 64, **kwargs) -> None:
        self.num_angle_bins = num_angle_bins
        self.phase_shifts: np.ndarray = None
        self.angle_bins: np.ndarray = None
        self.num_range_bins: int = None
        self.range_bins: np.ndarray = None
        super().__init__(config_manager)

    def _configure_axes(self):
        cm = self.config_manager
        self.num_range_bins = cm.get_num_adc_samples(profile_idx=0)
        # NOTE the +1e-3 is added to the VALUES here (ref :31-34), unlike RangeDopplerProcessor
        self.range_bins = np.arange(start=0, step=cm.range_res_m, stop=cm.range_max_m - cm.range_res_m / 2) + 1e-3
        self.num_rx_antennas = cm.num_rx_antennas
        a = self.num_angle_bins
        self.phase_shifts = np.arange(start=np.pi, stop=-np.pi - 2 * np.pi / (a - 1), step=-2 * np.pi / (a - 1))
        self.phase_shifts[-1] = -1 * np.pi
        with np.errstate(invalid="ignore"):
            return np.arcsin(self.phase_shifts / np.pi)

    def _mesh(self, angle_bins):
        self.thetas, self.rhos = np.meshgrid(angle_bins, self.range_bins)
        self.x_s = np.multiply(self.rhos, np.cos(self.thetas))
        self.y_s = np.multiply(self.rhos, np.sin(self.thetas))

    def configure(self):
        self.angle_bins = self._configure_axes()
        self._mesh(self.angle_bins)

    def process(self, adc_cube, chirp_idx=0, rx_antennas: np.ndarray = np.array([]),
                perform_windowing: bool = True, **kwargs):
        if isinstance(rx_antennas, list):
            rx_antennas = np.array(rx_antennas)
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        ants = [int(a) % eng.V for a in np.asarray(rx_antennas).ravel()] if np.asarray(rx_antennas).size > 0 else None
        loop = int(chirp_idx) % eng.L
        out = eng.range_angle(dev, loop_idx=loop, ant_list=ants, num_angle_bins=int(self.num_angle_bins),
                              windowing=bool(perform_windowing))
        if not squeeze:
            return out
        return to_host_f64(out[0])
