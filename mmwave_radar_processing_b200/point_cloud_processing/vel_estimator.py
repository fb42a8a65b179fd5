"""Ego velocity from a radar point cloud (reference: point_cloud_processing/vel_estimator.py:7-243) on the CUDA kernel.

``process(points)``: least squares of -v against the unit direction of every point inside RANSAC (sklearn's
RANSACRegressor in the reference: residual_threshold 0.15, random_state 42, max_trials 20, min_samples 10), 2-D for
"standard" arrays and 3-D for "ods"; the proposal replaces the current estimate when R^2 and the inlier share clear their
thresholds.  The fit runs in ``radar_ego_velocity`` (one warp per cloud, fp64); the random subsets sklearn would draw are
tabulated on the host with numpy's RandomState (``ransac_subsets``).  scikit-learn is not imported.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ..processors._device import is_device_tensor
from ..processors._processor import _Processor
from .ransac_subsets import SubsetTable


class VelocityEstimator(_Processor):
    """Coordinate system: x forward, y left."""

    def __init__(self, config_manager: ConfigManager, min_R2_threshold: float = 0.6, min_inlier_percent: float = 0.75,
                 subset_policy: str = "permutation", **kwargs) -> None:
        super().__init__(config_manager=config_manager)
        self.min_R2_threshold = min_R2_threshold
        self.min_inlier_percent = min_inlier_percent
        self.estimated_R2: float = 0.0
        self.inlier_percent: float = 0.0
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
        self.history_R2_statistics.append(self.estimated_R2)
        self.history_inlier_statistics.append(self.inlier_percent)
        return super().update_history(estimated=estimated, ground_truth=ground_truth)

    # ------------------------------------------------------------------ device call
    def _engine(self):
        from ..engine import RadarEngine
        return RadarEngine.for_config(self.config_manager)

    def estimate_batch(self, points, offsets, dims=None):
        """Extension: packed device point clouds (float64 [rows, 4], FramePipeline's layout) + offsets int32 [F+1] ->
        (vel [F,3], r2 [F], inlier fraction [F], status [F]) device tensors.  One small download of ``offsets`` decides
        which subset blocks must exist."""
        eng = self._engine()
        dims = (2 if self.config_manager.array_geometry == "standard" else 3) if dims is None else dims
        counts = np.diff(offsets.detach().cpu().numpy())
        self._table.ensure(counts)
        return eng.ego_velocity(points, offsets, dims, self._table)

    def _fit(self, points: np.ndarray, dims: int):
        """-> (coef [dims] or zeros(2) when sklearn would have raised, r2, inlier share)."""
        import torch
        eng = self._engine()
        pts = np.ascontiguousarray(points[:, 0:4], dtype=np.float64)
        n = pts.shape[0]
        self._table.ensure([n])
        dev = torch.from_numpy(pts).to(eng.device)
        off = torch.tensor([0, n], dtype=torch.int32, device=eng.device)
        vel, r2, frac, status = eng.ego_velocity(dev, off, dims, self._table)
        if int(status[0].item()) <= 0:
            return np.array([0.0, 0.0]), 0.0, 0.0          # the reference's `except ValueError` branch
        return vel[0, :dims].cpu().numpy(), float(r2[0].item()), float(frac[0].item())

    def lsq_fit_ego_vel_ransac_points_2D(self, points: np.ndarray = np.empty(shape=0)):
        if points.shape[0] == 0:
            return np.array([0.0, 0.0])
        return self._fit(points, 2)

    def lsq_fit_ego_vel_ransac_points_3D(self, points: np.ndarray = np.empty(shape=0)):
        if points.shape[0] == 0:
            return np.array([0.0, 0.0, 0.0]), 0.0, 0.0
        return self._fit(points, 3)

    def update_and_check_current_vel_measurements(self):
        if self.estimated_R2 >= self.min_R2_threshold and self.inlier_percent >= self.min_inlier_percent:
            self.current_velocity_estimate = self.proposed_velocity_estimate

    def estimate_ego_velocity_points(self, points: np.ndarray = np.empty(shape=0)):
        if points.shape[0] >= 0:
            if self.config_manager.array_geometry == "standard":
                vel_est, self.estimated_R2, self.inlier_percent = self.lsq_fit_ego_vel_ransac_points_2D(points=points)
                self.proposed_velocity_estimate = np.array([vel_est[0], vel_est[1], 0.0])
            elif self.config_manager.array_geometry == "ods":
                vel_est, self.estimated_R2, self.inlier_percent = self.lsq_fit_ego_vel_ransac_points_3D(points=points)
                self.proposed_velocity_estimate = vel_est      # a 2-vector when the fit failed, as in the reference

    def process(self, points: np.ndarray = np.empty(shape=0), **kwargs) -> np.ndarray:
        """-> velocity estimate [vx, vy, vz]"""
        if isinstance(points, list):
            points = np.array(points)
        if is_device_tensor(points):
            points = points.detach().cpu().numpy()
        if points.shape[0] > 0:
            self.estimate_ego_velocity_points(points=points)
        self.update_and_check_current_vel_measurements()
        return self.current_velocity_estimate
