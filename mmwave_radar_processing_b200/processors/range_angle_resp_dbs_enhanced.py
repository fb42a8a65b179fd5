"""Range-angle map with Doppler beam sharpening (reference: processors/range_angle_resp_dbs_enhanced.py:7-342).

Below the speed gate it is the plain range-angle map.  Above it, the host picks - per output angle - the
Doppler bin nearest to ``-r_hat(theta).v`` and the nearest coarse angle bin (fp64 argmins, ref :239-258) and
the device evaluates the windowed range/Doppler/angle transform at exactly those cells (``radar_dbs``).
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import stage_cube, to_host_f64
from .range_angle_resp import RangeAngleProcessor


class RangeAngleProcessorDBSEnhanced(RangeAngleProcessor):

    def __init__(self, config_manager: ConfigManager, num_angle_bins_range_angle_response: int = 64,
                 num_angle_bins_dbs_enhanced_response: int = 64, min_x_y_vel_dbs: float = 0.25, **kwargs) -> None:
        self.angle_bins_no_dbs_enhancement: np.ndarray = None
        self.num_angle_bins_dbs_enhanced_response = num_angle_bins_dbs_enhanced_response
        self.angle_bins_dbs_enhanced: np.ndarray = None
        self.phase_shifts_dbs_enhanced: np.ndarray = None
        self.min_vel_dbs = min_x_y_vel_dbs
        self.vel_bins: np.ndarray = None
        super().__init__(config_manager=config_manager, num_angle_bins=num_angle_bins_range_angle_response, **kwargs)

    def configure(self) -> None:
        cm = self.config_manager
        self.angle_bins_no_dbs_enhancement = self._configure_axes()
        self.vel_bins = np.arange(start=-1 * cm.vel_max_m_s, stop=cm.vel_max_m_s - cm.vel_res_m_s + 1e-3,
                                  step=cm.vel_res_m_s)
        self.angle_bins_dbs_enhanced = np.linspace(start=self.angle_bins_no_dbs_enhancement[0],
                                                   stop=self.angle_bins_no_dbs_enhancement[-1],
                                                   num=self.num_angle_bins_dbs_enhanced_response)
        self.compute_mesh_grid()

    def compute_mesh_grid_dbs_enhanced(self) -> None:
        self.angle_bins = self.angle_bins_dbs_enhanced
        self._mesh(self.angle_bins_dbs_enhanced)

    def compute_mesh_grid(self) -> None:
        self.angle_bins = self.angle_bins_no_dbs_enhancement
        self._mesh(self.angle_bins)

    def process_no_dbs(self, adc_cube, chirp_idx=0, rx_antennas: np.ndarray = np.array([]), **kwargs):
        self.compute_mesh_grid()
        return super().process(adc_cube=adc_cube, chirp_idx=chirp_idx, rx_antennas=rx_antennas, **kwargs)

    def get_dop_vel(self, angle: float, ego_vel: np.ndarray) -> float:
        r = np.array([np.cos(angle), np.sin(angle), 0])
        return -1 * np.dot((r / np.linalg.norm(r)), ego_vel)

    def select_bins(self, velocity_ned: np.ndarray):
        """(Doppler bin, coarse angle bin) per output angle - same expressions as ref :239-258."""
        n = self.angle_bins_dbs_enhanced.shape[0]
        vel_bin = np.empty(n, dtype=np.int64)
        ang_bin = np.empty(n, dtype=np.int64)
        for i in range(n):
            dop = self.get_dop_vel(angle=self.angle_bins_dbs_enhanced[i], ego_vel=velocity_ned)
            vel_bin[i] = np.argmin(np.abs(self.vel_bins - dop))
            ang_bin[i] = np.argmin(np.abs(self.angle_bins_no_dbs_enhancement - self.angle_bins_dbs_enhanced[i]))
        return vel_bin, ang_bin

    def process_dbs_enhanced(self, adc_cube, velocity_ned: np.ndarray, rx_antennas: np.ndarray = np.array([]), **kwargs):
        self.compute_mesh_grid_dbs_enhanced()
        if isinstance(rx_antennas, list):
            rx_antennas = np.array(rx_antennas)
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        ants = [int(a) % eng.V for a in np.asarray(rx_antennas).ravel()] if np.asarray(rx_antennas).size > 0 else None
        vel_bin, ang_bin = self.select_bins(np.asarray(velocity_ned, dtype=float))
        out = eng.dbs(dev, vel_bin.tolist(), ang_bin.tolist(), ant_list=ants, num_angle_bins=int(self.num_angle_bins))
        if not squeeze:
            return out
        return to_host_f64(out[0])

    def process(self, adc_cube, velocity_ned: np.ndarray, rx_antennas: np.ndarray = np.array([]),
                chirp_idx: int = 0, **kwargs):
        if np.linalg.norm(np.asarray(velocity_ned)[0:2]) < self.min_vel_dbs:
            return self.process_no_dbs(adc_cube=adc_cube, chirp_idx=chirp_idx, rx_antennas=rx_antennas, **kwargs)
        return self.process_dbs_enhanced(adc_cube=adc_cube, velocity_ned=velocity_ned, rx_antennas=rx_antennas, **kwargs)
