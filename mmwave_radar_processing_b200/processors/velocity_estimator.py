"""ADC-domain ego-velocity estimator (reference: processors/velocity_estimator.py:7-865) on the CUDA kernels.

A ``DopplerAzimuthProcessor`` subclass: Doppler-azimuth (and, for ODS arrays, Doppler-elevation) responses over a range
window around the altitude, the zero-azimuth peak gives vx, the per-row peaks feed a 1-D least-squares fit inside RANSAC
for the lateral components.  The responses come from ``radar_doppler_azimuth`` / ``radar_chirp_zoom``, the peak picking from
``radar_find_peaks`` and every RANSAC fit (scikit-learn's ``RANSACRegressor(random_state=42)`` in the reference) from
``radar_ego_velocity`` with the host-tabulated subsets.  scikit-learn and scipy are not imported.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ..point_cloud_processing.ransac_subsets import SubsetTable
from .doppler_azimuth_resp import DopplerAzimuthProcessor

_FAIL = (0.0, 0.0, 0.0)


class VelocityEstimator(DopplerAzimuthProcessor):
    """Coordinate system: x forward, y left."""

    def __init__(self, config_manager: ConfigManager, lower_range_bound: float, upper_range_bound: float,
                 precise_vel_bound: float = 0.25,
                 valid_angle_range: np.ndarray = np.array([np.deg2rad(-70), np.deg2rad(70)]),
                 peak_threshold_dB: float = 30.0, x_measurement_only: bool = False, min_R2_threshold: float = 0.6,
                 min_inlier_percent: float = 0.75, subset_policy: str = "permutation", **kwargs) -> None:
        if isinstance(valid_angle_range, list):
            valid_angle_range = np.array(valid_angle_range)
        super().__init__(config_manager=config_manager, num_angle_bins=64, valid_angle_range=valid_angle_range)
        self.upper_range_bound = upper_range_bound
        self.lower_range_bound = lower_range_bound
        self.precise_vel_bound = precise_vel_bound
        self.azimuth_response_mag = np.empty(shape=0)
        self.az_resps_mag = []
        self.elevation_response_mag = np.empty(shape=0)
        self.el_resps_mag = []
        self.precise_azimuth_response_mag = np.empty(shape=0)
        self.precise_elevation_response_mag = np.empty(shape=0)
        self.peak_threshold_dB = peak_threshold_dB
        self.azimuth_peaks: np.ndarray = np.empty(shape=0)
        self.azimuth_peak_zero_az: np.ndarray = np.empty(shape=0)
        self.elevation_peaks: np.ndarray = np.empty(shape=0)
        self.elevation_peak_zero_az: np.ndarray = np.empty(shape=0)
        self.x_measurement_only = x_measurement_only
        self.min_R2_threshold = min_R2_threshold
        self.min_inlier_percent = min_inlier_percent
        self.ego_vx_estimate: float = -1.0
        self.azimuth_ego_vy_estimate: float = 0.0
        self.azimuth_estimate_R2: float = 0.0
        self.azimuth_inlier_percent: float = 0.0
        self.elevation_ego_vy_estimate: float = 0.0
        self.elevation_estimate_R2: float = 0.0
        self.elevation_inlier_percent: float = 0.0
        self.proposed_velocity_estimate: np.ndarray = np.empty(shape=0)
        self.current_velocity_estimate: np.ndarray = np.array([0.0, 0.0, 0.0])
        self.history_R2_statistics = []
        self.history_inlier_statistics = []
        self._table = SubsetTable(min_samples=10, max_trials=20, seed=42, policy=subset_policy)

    def reset(self):
        self.history_R2_statistics = []
        self.history_inlier_statistics = []
        return super().reset()

    def update_history(self, estimated: np.ndarray = np.empty(0), ground_truth: np.ndarray = np.empty(0)) -> None:
        geom = self.config_manager.array_geometry
        if geom == "ods":
            self.history_R2_statistics.append(np.array([self.azimuth_estimate_R2, self.elevation_estimate_R2]))
            self.history_inlier_statistics.append(np.array([self.azimuth_inlier_percent, self.elevation_inlier_percent]))
        elif geom == "standard":
            self.history_R2_statistics.append(np.array([self.azimuth_estimate_R2]))
            self.history_inlier_statistics.append(np.array([self.azimuth_inlier_percent]))
        return super().update_history(estimated=estimated, ground_truth=ground_truth)

    # ------------------------------------------------------------------ responses (ref :129-264)
    def get_range_window(self, altitude: float = 0.0, sensing_direction: str = "down") -> np.ndarray:
        if sensing_direction == "down":
            return np.array([max(0, altitude - self.lower_range_bound),
                             min(self.config_manager.range_max_m, altitude + self.upper_range_bound)])
        elif sensing_direction == "out":
            return np.array([1.0, self.config_manager.range_max_m])

    def _pair_response(self, adc_cube, sets, range_window, shift_angle, use_precise_fft, center_vel):
        vel_range = np.array([center_vel - self.precise_vel_bound, center_vel + self.precise_vel_bound])
        return [DopplerAzimuthProcessor.process(self, adc_cube=adc_cube, rx_antennas=s, range_window=range_window,
                                                shift_angle=shift_angle, use_precise_fft=use_precise_fft,
                                                precise_vel_range=vel_range)      # clamped in place, as upstream
                for s in sets]

    def compute_azimuth_response(self, adc_cube, range_window, use_precise_fft: bool = False,
                                 precise_fft_center_vel: float = 0.0):
        cm = self.config_manager
        if cm.array_geometry == "standard":
            sets = [np.arange(8) if cm.virtual_antennas_enabled else np.arange(4)]
            resp = self._pair_response(adc_cube, sets, range_window, True, use_precise_fft, precise_fft_center_vel)[0]
        elif cm.array_geometry == "ods":
            sets = [np.array([0, 3, 4, 7]), np.array([1, 2, 5, 6])] if cm.virtual_antennas_enabled else \
                [np.array([0, 3]), np.array([1, 2])]
            r1, r2 = self._pair_response(adc_cube, sets, range_window, True, use_precise_fft, precise_fft_center_vel)
            self.az_resps_mag = [r1, r2]
            resp = (r1 + r2) / 2
        if use_precise_fft:
            self.precise_azimuth_response_mag = resp
        else:
            self.azimuth_response_mag = resp

    def compute_elevation_response(self, adc_cube, range_window, use_precise_fft: bool = False,
                                   precise_fft_center_vel: float = 0.0):
        cm = self.config_manager
        if cm.array_geometry == "standard":
            raise NotImplementedError("Elevation response computation is not implemented for standard array geometry.")
        sets = [np.array([10, 11, 6, 7]), np.array([9, 8, 5, 4])] if cm.virtual_antennas_enabled else \
            [np.array([1, 0]), np.array([3, 4])]
        r1, r2 = self._pair_response(adc_cube, sets, range_window, False, use_precise_fft, precise_fft_center_vel)
        self.el_resps_mag = [r1, r2]
        resp = (r1 + r2) / 2
        if use_precise_fft:
            self.precise_elevation_response_mag = resp
        else:
            self.elevation_response_mag = resp

    # ------------------------------------------------------------------ peaks (ref :266-344)
    def _responses(self, precise):
        if precise:
            return self.precise_azimuth_response_mag, self.precise_elevation_response_mag, self.zoomed_vel_bins
        return self.azimuth_response_mag, self.elevation_response_mag, self.vel_bins

    def detect_vel_row_peaks(self, use_precise_response: bool = False):
        az, el, bins = self._responses(use_precise_response)
        if az.shape[0] > 0:
            self.azimuth_peaks = self.detect_peaks_rows(az, bins, self.peak_threshold_dB)
        if el.shape[0] > 0:
            self.elevation_peaks = self.detect_peaks_rows(el, bins, self.peak_threshold_dB)

    def detect_vel_zero_az_peaks(self, use_precise_response: bool = False):
        az, el, bins = self._responses(use_precise_response)
        if az.shape[0] > 0:
            self.azimuth_peak_zero_az = self.detect_peak_zero_az(az, bins, self.peak_threshold_dB)
        if el.shape[0] > 0:
            self.elevation_peak_zero_az = self.detect_peak_zero_az(el, bins, self.peak_threshold_dB)

    # ------------------------------------------------------------------ RANSAC fits (ref :344-507) on radar_ego_velocity
    def _ransac(self, rows: np.ndarray, dims: int, residual_threshold: float):
        """rows float64 [N, 4]: point rows (x, y, z, v) for dims 2 / 3, raw design rows (h0, h1, h2, y) for dims < 0.
        -> (coef, r2 over the inliers, inlier share), or None where sklearn raises ValueError (fewer rows than
        min_samples, non-finite input, no consensus set)."""
        import torch
        from ..engine import RadarEngine
        eng = RadarEngine.for_config(self.config_manager)
        rows = np.ascontiguousarray(rows, dtype=np.float64)
        n = rows.shape[0]
        self._table.ensure([n])
        off = torch.tensor([0, n], dtype=torch.int32, device=eng.device)
        vel, r2, frac, status = eng.ego_velocity(torch.from_numpy(rows).to(eng.device), off, dims, self._table,
                                                 residual_threshold=residual_threshold)
        if int(status[0].item()) <= 0:
            return None
        return vel[0].cpu().numpy(), float(r2[0].item()), float(frac[0].item())

    @staticmethod
    def _design_rows(h: np.ndarray, y: np.ndarray) -> np.ndarray:
        rows = np.zeros((h.shape[0], 4), dtype=np.float64)
        rows[:, 0] = h
        rows[:, 3] = y
        return rows

    def lsq_fit_ego_vy_ransac(self, peaks: np.ndarray):
        if self.ego_vx_estimate >= 0.1:
            return self.lsq_fit_ego_vy_ransac_standard(peaks=peaks)
        return self.lsq_fit_ego_vy_ransac_small_vx(peaks=peaks)

    def lsq_fit_ego_vel_ransac_points(self, points: np.ndarray = np.empty(shape=0)):
        """Nx4 (x, y, z, velocity) -> ([vx, vy], r2, inlier share)."""
        if points.shape[0] == 0:
            return np.array([0.0, 0.0])
        fit = self._ransac(points[:, 0:4], 2, 0.15)
        if fit is None:
            return (np.array([0.0, 0.0]),) + _FAIL[1:]
        return fit[0][:2], fit[1], fit[2]

    def lsq_fit_ego_vy_ransac_standard(self, peaks: np.ndarray):
        """vd = -vx cos(theta) - vy sin(theta) with vx known: regress -vd - vx cos(theta) on sin(theta)."""
        if peaks is None or len(peaks) == 0:
            return 0.0
        fit = self._ransac(self._design_rows(np.sin(peaks[:, 0]),
                                             -1 * peaks[:, 1] - self.ego_vx_estimate * np.cos(peaks[:, 0])), -1, 0.15)
        if fit is None:
            return _FAIL
        return fit[0][0], fit[1], fit[2]

    def lsq_fit_ego_vy_ransac_small_vx(self, peaks: np.ndarray):
        """Small vx: theta = -(1/vy) (vd - vx), regress the angle on (vd - vx)."""
        if peaks is None or len(peaks) == 0:
            return 0.0, 0.0
        fit = self._ransac(self._design_rows(peaks[:, 1] - self.ego_vx_estimate, peaks[:, 0]), -1, 0.20)
        if fit is None:
            return _FAIL
        a = fit[0][0]
        return (-1 / a if a != 0 else 0.0), fit[1], fit[2]

    def lsq_fit_ego_vy(self, peaks: np.ndarray):
        """Plain least squares of the standard model (one unknown: closed form)."""
        if peaks is None or len(peaks) == 0:
            return 0.0
        h = np.sin(peaks[:, 0])
        y = -1 * peaks[:, 1] - (self.ego_vx_estimate * np.cos(peaks[:, 0]))
        hh = float(h @ h)
        return float(h @ y) / hh if hh > 0 else 0.0

    def lsq_predict_velocity_measurement(self, v: np.ndarray, angles_to_predict: np.ndarray = np.empty(shape=0)):
        if angles_to_predict.size == 0:
            angles_to_predict = self.valid_angle_bins
        return -1 * np.dot(np.stack([np.cos(angles_to_predict), np.sin(angles_to_predict)], axis=-1), v)

    # ------------------------------------------------------------------ statistics / gating (ref :546-638)
    @staticmethod
    def _r2(measured, predicted):
        rss = np.sum((measured - predicted) ** 2)
        tss = np.sum((measured - np.mean(measured)) ** 2)
        return 1 - rss / tss if tss > 0 else 0

    def compute_R2_statistics(self):
        geom = self.config_manager.array_geometry
        v = self.proposed_velocity_estimate
        az, el = self.azimuth_peaks, self.elevation_peaks
        if az.shape[0] > 0:
            if geom == "ods":
                pred = self.lsq_predict_velocity_measurement(v=np.array([v[2], v[0]]), angles_to_predict=az[:, 0])
            elif geom == "standard":
                pred = self.lsq_predict_velocity_measurement(v=np.array([v[0], v[1]]), angles_to_predict=az[:, 0])
            self.azimuth_estimate_R2 = self._r2(az[:, 1], pred)
        if el.shape[0] > 0:
            if geom != "ods":
                raise NotImplementedError("Elevation R2 computation is only implemented for ODS array geometry.")
            pred = self.lsq_predict_velocity_measurement(v=np.array([v[2], v[1]]), angles_to_predict=el[:, 0])
            self.elevation_estimate_R2 = self._r2(el[:, 1], pred)

    def update_and_check_current_vel_measurements(self):
        cur, prop = self.current_velocity_estimate, self.proposed_velocity_estimate
        geom = self.config_manager.array_geometry
        if self.x_measurement_only:
            cur[0] = prop[0]
        elif geom == "ods":
            az_ok = self.azimuth_estimate_R2 >= self.min_R2_threshold and \
                self.azimuth_inlier_percent >= self.min_inlier_percent
            el_ok = self.elevation_estimate_R2 >= self.min_R2_threshold and \
                self.elevation_inlier_percent >= self.min_inlier_percent
            cur[0] = prop[0] if az_ok else 0.0
            cur[1] = prop[1] if el_ok else 0.0
            cur[2] = prop[2]
        elif geom == "standard":
            good = self.azimuth_estimate_R2 >= self.min_R2_threshold
            if self.ego_vx_estimate < 0.0:          # the point-cloud branch left ego_vx at its initial -1
                self.current_velocity_estimate = prop if good else np.array([0.0, 0.0, 0.0])
            else:
                cur[1] = prop[1] if good else 0.0
                cur[0] = prop[0]
                cur[2] = 0.0

    # ------------------------------------------------------------------ estimates (ref :640-723)
    def estimate_ego_vx_velocity(self):
        """vx = minus the Doppler of the zero-azimuth (/ zero-elevation) peak, averaged when both exist."""
        found = [p[1] for p in (self.azimuth_peak_zero_az, self.elevation_peak_zero_az) if p.shape[0] > 0]
        if len(found) == 2:
            self.ego_vx_estimate = -1 * (found[0] + found[1]) / 2.0
        elif len(found) == 1:
            self.ego_vx_estimate = -1 * found[0]
        else:
            self.ego_vx_estimate = 0.0
        return self.ego_vx_estimate

    def estimate_ego_velocity_adc_data(self):
        if self.x_measurement_only:
            self.proposed_velocity_estimate = np.array([self.ego_vx_estimate])
            return
        if self.azimuth_peaks.shape[0] > 0:
            self.azimuth_ego_vy_estimate, self.azimuth_estimate_R2, self.azimuth_inlier_percent = \
                self.lsq_fit_ego_vy_ransac(peaks=self.azimuth_peaks)
        if self.elevation_peaks.shape[0] > 0:
            self.elevation_ego_vy_estimate, self.elevation_estimate_R2, self.elevation_inlier_percent = \
                self.lsq_fit_ego_vy_ransac(peaks=self.elevation_peaks)
        geom = self.config_manager.array_geometry
        if geom == "ods":
            self.proposed_velocity_estimate = np.array([self.azimuth_ego_vy_estimate, self.elevation_ego_vy_estimate,
                                                        self.ego_vx_estimate])
        elif geom == "standard":
            self.proposed_velocity_estimate = np.array([self.ego_vx_estimate, self.azimuth_ego_vy_estimate, 0.0])

    def estimate_ego_velocity_points(self, points: np.ndarray = np.empty(shape=0)):
        if points.shape[0] < 0:
            return
        if self.config_manager.array_geometry == "ods":
            raise NotImplementedError(
                "Velocity estimation from points is not implemented for ODS array geometry currently.")
        vel_est, self.azimuth_estimate_R2, self.azimuth_inlier_percent = self.lsq_fit_ego_vel_ransac_points(points=points)
        if self.x_measurement_only:
            self.proposed_velocity_estimate = np.array([vel_est[0]])
        else:
            self.proposed_velocity_estimate = np.array([vel_est[0], vel_est[1], 0.0])

    def _predict_pair(self, vel, direction, sign_z):
        geom = self.config_manager.array_geometry
        if direction not in ("azimuth", "elevation"):
            raise ValueError("Direction must be either 'azimuth' or 'elevation'")
        if geom == "ods":
            return np.array([sign_z * vel[2], vel[1] if (direction == "azimuth") == (sign_z < 0) else vel[0]])
        if direction == "elevation":
            raise NotImplementedError("Elevation prediction is not implemented for standard array geometry.")
        return np.array([vel[0], vel[1]])

    def get_gt_velocity_measurement_predictions(self, direction: str = "azimuth"):
        """Doppler each valid angle bin would show under the latest ground-truth velocity (ods: v = [-gt_z, gt_y] for
        azimuth, [-gt_z, gt_x] for elevation)."""
        if len(self.history_gt) == 0:
            return np.empty(shape=(0))
        return self.lsq_predict_velocity_measurement(v=self._predict_pair(self.history_gt[-1], direction, -1))

    def get_estimated_velocity_measurement_predictions(self, direction: str = "azimuth"):
        """Same under the current estimate (ods: v = [est_z, est_x] for azimuth, [est_z, est_y] for elevation)."""
        if len(self.proposed_velocity_estimate) == 0:
            return np.empty(shape=(0))
        return self.lsq_predict_velocity_measurement(v=self._predict_pair(self.current_velocity_estimate, direction, 1))

    # ------------------------------------------------------------------ entry point (ref :783-865)
    def process(self, adc_cube: np.ndarray = np.empty(shape=0), points: np.ndarray = np.empty(shape=0),
                altitude: float = 0.0, enable_precise_responses: bool = False, **kwargs) -> np.ndarray:
        """adc_cube indexed [rx, sample, chirp] (host complex array, or the device int16 frame) and / or an Nx4 point
        cloud -> velocity estimate [vx, vy, vz]."""
        if isinstance(points, list):
            points = np.array(points)
        ods = self.config_manager.array_geometry == "ods"
        if adc_cube.shape[0] > 0:
            window = self.get_range_window(altitude=altitude, sensing_direction=self.config_manager.array_direction)
            passes = [(False, 0.0)]
            if enable_precise_responses:
                passes.append((True, None))
            for precise, center in passes:
                center = -1 * self.ego_vx_estimate if center is None else center
                self.compute_azimuth_response(adc_cube=adc_cube, range_window=window, use_precise_fft=precise,
                                              precise_fft_center_vel=center)
                if ods:
                    self.compute_elevation_response(adc_cube=adc_cube, range_window=window, use_precise_fft=precise,
                                                    precise_fft_center_vel=center)
                self.detect_vel_zero_az_peaks(use_precise_response=precise)
                self.estimate_ego_vx_velocity()
            if not self.x_measurement_only:
                self.detect_vel_row_peaks(use_precise_response=enable_precise_responses)
            self.estimate_ego_velocity_adc_data()
        elif points.shape[0] > 0:
            self.estimate_ego_velocity_points(points=points)
        self.update_and_check_current_vel_measurements()
        return self.current_velocity_estimate
