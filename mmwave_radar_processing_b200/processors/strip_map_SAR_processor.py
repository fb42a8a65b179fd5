"""Strip-map SAR response (reference: processors/strip_map_SAR_processor.py:8-196) on the CUDA kernel.

The image is the UN-windowed 2-D spectrum of one receiver over (samples, chirps) with the chirp axis shifted - the
range-Doppler kernel with window_flags = 0 - cut to the valid range rows and cross-range (angle) columns.  The array
geometry, the slices and the ground-plane mesh are host fp64 with the reference's expressions.
"""
import numpy as np

from ..config_managers.cfgManager import SPEED_OF_LIGHT_M_S
from ._device import is_device_tensor, stage_cube, to_host_c128
from ._processor import _Processor
from .virtual_array_reformater import VirtualArrayReformatter


class StripMapSARProcessor(_Processor):

    def __init__(self, config_manager, az_angle_range_rad=np.deg2rad(np.array([-30, 30])), **kwargs):
        if isinstance(az_angle_range_rad, list):
            az_angle_range_rad = np.array(az_angle_range_rad)
        if config_manager.virtual_antennas_enabled:
            self.virtual_array_reformatter = VirtualArrayReformatter(config_manager)
        self.chirps_per_frame: int = None
        self.chirp_period_us: float = None
        self.chirp_tx_masks: np.ndarray = None
        self.chirp_rx_positions_m: np.ndarray = None
        self.lambda_m: float = None
        self.phase_shifts: np.ndarray = None
        self.angle_bins_rad: np.ndarray = None
        self.az_angle_range_rad: np.ndarray = az_angle_range_rad
        self.num_range_bins: int = None
        self.radar_range_bins: np.ndarray = None
        self.valid_ranges_slice: slice = None
        self.valid_angles_slice: slice = None
        self.ground_range_bins: np.ndarray = None
        self.ground_az_bins_rad: np.ndarray = None
        self.thetas: np.ndarray = None
        self.rhos: np.ndarray = None
        self.x_s: np.ndarray = None
        self.y_s: np.ndarray = None
        super().__init__(config_manager)

    def configure(self):
        self._compute_key_radar_parameters()

    def _compute_key_radar_parameters(self):
        cm = self.config_manager
        self.num_range_bins = cm.get_num_adc_samples(profile_idx=0)
        self.range_bins = np.linspace(start=0, stop=cm.range_max_m, num=self.num_range_bins)
        start_freq_GHz = float(cm.profile_cfgs[0]["startFreq_GHz"])
        self.lambda_m = SPEED_OF_LIGHT_M_S / (start_freq_GHz * 1e9)
        chirp_cfgs_per_loop = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
        self.chirps_per_frame = cm.frameCfg_loops * chirp_cfgs_per_loop
        self.phase_shifts = np.linspace(start=np.pi, stop=-np.pi, num=cm.frameCfg_loops)
        self.chirp_period_us = cm.profile_cfgs[0]["idleTime_us"] + cm.profile_cfgs[0]["rampEndTime_us"]

    def configure_array_geometry(self, vel_m_per_s: float, sensor_height_m: float, max_SAR_distance: float):
        """ref :103-158 (kept expression for expression, including arcsin(phase*lambda)/(2 pi d_rx))."""
        d_rx = 2 * self.chirp_period_us * 1e-6 * vel_m_per_s
        self.angle_bins_rad = np.arcsin(self.phase_shifts * self.lambda_m) / (2 * np.pi * d_rx)
        min_angle_idx = np.argmin(np.abs(self.angle_bins_rad - np.min(self.az_angle_range_rad)))
        max_angle_idx = np.argmin(np.abs(self.angle_bins_rad - np.max(self.az_angle_range_rad)))
        self.valid_angles_slice = slice(np.min([min_angle_idx, max_angle_idx]), np.max([min_angle_idx, max_angle_idx]))
        self.ground_az_bins_rad = self.angle_bins_rad[self.valid_angles_slice]
        min_rng_idx = np.nonzero(self.range_bins > sensor_height_m)[0][0]
        max_rng_idx = np.nonzero(self.range_bins < max_SAR_distance)[0][-1]
        self.valid_ranges_slice = slice(min_rng_idx, max_rng_idx)
        self.ground_range_bins = np.sqrt(np.power(self.range_bins[self.valid_ranges_slice], 2) - np.power(d_rx, 2))
        self.thetas, self.rhos = np.meshgrid(self.ground_az_bins_rad, self.ground_range_bins, indexing='xy')
        self.x_s = np.multiply(self.rhos, np.cos(self.thetas))
        self.y_s = np.multiply(self.rhos, np.sin(self.thetas))

    def process(self, adc_cube, vel_m_per_s: float, sensor_height_m: float = 0.24, rx_index=0,
                max_SAR_distance: float = 1.5, **kwargs):
        """adc_cube: the RAW (rx, samples, chirps) host cube as in the reference (the virtual array is formed here), or a
        device int16 IQ batch (extension; complex64 device cubes must already be virtual-array cubes)."""
        if self.config_manager.virtual_antennas_enabled and not is_device_tensor(adc_cube):
            adc_cube = self.virtual_array_reformatter.process(adc_cube=adc_cube)
        self.configure_array_geometry(vel_m_per_s=vel_m_per_s, sensor_height_m=sensor_height_m,
                                      max_SAR_distance=max_SAR_distance)
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        rd, _ = eng.range_doppler(dev, ant_list=[int(rx_index) % eng.V], window_flags=0, want_rd=True)
        resp = rd[:, 0, self.valid_ranges_slice, self.valid_angles_slice]
        if not squeeze:
            return resp
        return to_host_c128(resp[0])
