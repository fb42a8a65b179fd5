"""Multi-frame synthetic-array (SAR) beamformer
(reference: processors/simple_synthetic_array_beamformer_processor_multiFrame.py:10-872).

Host side (fp64, same ordering as the reference): valid-chirp selection, velocity-history validity gate,
element positions chained backwards over frames, pointing vectors, mesh grids.  Device side
(``radar_sar_image``): Hamming-weighted steering + beam sum + Hann + range FFT for every (azimuth,
elevation) voxel; the ring buffer of valid chirps lives in device memory as complex64.

``enable_calibration`` (ref :673-816) runs in ``radar_sar_calibrate`` (fp64 on the device, from the history ring
and the image already there), the polar->cartesian ``griddata`` resampling (ref :587-613) in
``radar_polar_to_cart``.  scipy is not imported.
"""
import numpy as np

from ..config_managers.cfgManager import SPEED_OF_LIGHT_M_S, ConfigManager
from ._processor import _Processor


def _arr(v):
    return np.array(v) if isinstance(v, list) else v


class SyntheticArrayBeamformerProcessor(_Processor):

    def __init__(self, config_manager: ConfigManager, receiver_idx: int = 0, chirp_cfg_idx: int = 0,
                 num_frames: int = 2, stride: int = 1,
                 az_angle_bins_rad=np.deg2rad(np.linspace(start=-30, stop=30, num=60)),
                 el_angle_bins_rad=np.array([0]),
                 min_vel: np.ndarray = np.array([0.17, 0.0, 0.0]),
                 max_vel: np.ndarray = np.array([0.25, 0.05, 0.05]),
                 max_vel_stdev: np.ndarray = np.array([0.1, 0.1, 0.1]),
                 enable_calibration: bool = False, num_calibration_iters: int = 1,
                 interpolated_grid_resolution_m=0.1, **kwargs) -> None:
        self.receiver_idx = receiver_idx
        self.chirp_cfg_idx = chirp_cfg_idx
        self.num_frames = num_frames
        self.stride = stride
        self.num_range_bins: int = None
        self.range_bins: np.ndarray = None
        self.chirps_per_frame: int = None
        self.frame_period_ms: float = None
        self.chirp_period_us: float = None
        self.chirp_cfg_idxs: np.ndarray = None
        self.valid_chirps_mask: np.ndarray = None
        self.lambda_m: float = None
        self.chirp_start_times_us: np.ndarray = None
        self.history_avg_vel: np.ndarray = None
        self.array_geometry: np.ndarray = np.empty(shape=0)
        self.array_geometry_calibrated: np.ndarray = np.empty(shape=0)
        self.az_angle_bins_rad = _arr(az_angle_bins_rad)
        self.el_angle_bins_rad = _arr(el_angle_bins_rad)
        self.d: np.ndarray = None
        self.beamformed_resp: np.ndarray = None
        self.interpolated_beamformed_resp: np.ndarray = np.empty(shape=0)
        self.rhos = self.thetas = self.phis = None
        self.x_s = self.y_s = self.z_s = None
        self.orig_grid_points: np.ndarray = None
        self.interpolated_grid_resolution_m = interpolated_grid_resolution_m
        self.interp_x_s = self.interp_y_s = self.interp_grid_points = None
        self.min_vel = _arr(min_vel)
        self.max_vel = _arr(max_vel)
        self.max_vel_stdev = _arr(max_vel_stdev)
        self.enable_calibration = enable_calibration
        self._resp_dev = None
        self.calibration_targets = []
        self.num_calibration_iters = num_calibration_iters
        self.array_geometry_valid = False
        self._hist_dev = None            # complex64 [F, S, Lv] on the device, oldest -> newest
        self._eng = None
        super().__init__(config_manager)

    # ------------------------------------------------------------------ configuration (host, fp64)
    def configure(self):
        self._compute_key_radar_parameters()
        self._compute_mesh_grids()
        self._compute_beam_stearing_vectors()
        self._init_out_resp()
        self._reset_vel_history()

    def _reset_vel_history(self):
        self.history_avg_vel = np.zeros(shape=(self.num_frames, 3), dtype=float)
        self.array_geometry_valid = False

    def _compute_key_radar_parameters(self):
        cm = self.config_manager
        self.num_range_bins = cm.get_num_adc_samples(profile_idx=0)
        self.range_bins = np.linspace(start=0, stop=cm.range_max_m, num=self.num_range_bins)   # NB: linspace here (ref :182-186)
        self.lambda_m = SPEED_OF_LIGHT_M_S / (float(cm.profile_cfgs[0]["startFreq_GHz"]) * 1e9)
        cpl = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
        self.chirps_per_frame = cm.frameCfg_loops * cpl
        self.chirp_period_us = cm.profile_cfgs[0]["idleTime_us"] + cm.profile_cfgs[0]["rampEndTime_us"]
        self.frame_period_ms = cm.frameCfg_periodicity_ms
        slots = np.arange(start=cm.frameCfg_start_index, stop=cm.frameCfg_end_index + 1)
        self.chirp_cfg_idxs = np.tile(A=slots, reps=cm.frameCfg_loops).flatten()
        picked = np.where(self.chirp_cfg_idxs == self.chirp_cfg_idx)[0][::self.stride]
        self.valid_chirps_mask = np.zeros_like(self.chirp_cfg_idxs, dtype=bool)
        self.valid_chirps_mask[picked] = True
        # chirp start times are <= 0: the LAST chirp of the frame sits at t = 0 (ref :226-234)
        all_times = np.arange(start=self.chirps_per_frame - 1, stop=-1, step=-1) * -self.chirp_period_us
        self.chirp_start_times_us = all_times[self.valid_chirps_mask]
        self._num_valid = int(np.sum(self.valid_chirps_mask))
        self._hist_dev = None

    @property
    def history_acd_cube_valid_chirps(self) -> np.ndarray:
        """Ring buffer [frame, sample, chirp] (downloaded view of the device copy)."""
        if self._hist_dev is None:
            return np.zeros((self.num_frames, self.num_range_bins, self._num_valid), dtype=complex)
        return self._hist_dev.cpu().numpy().astype(np.complex128)

    def _compute_mesh_grids(self):
        self.rhos, self.thetas, self.phis = np.meshgrid(self.range_bins, self.az_angle_bins_rad,
                                                        self.el_angle_bins_rad, indexing="ij")
        self.x_s = np.multiply(np.multiply(self.rhos, np.cos(self.thetas)), np.cos(self.phis))
        self.y_s = np.multiply(np.multiply(self.rhos, np.sin(self.thetas)), np.cos(self.phis))
        self.z_s = np.multiply(self.rhos, np.sin(self.phis))
        x_flat, y_flat = self.x_s.flatten(), self.y_s.flatten()
        x_lin = np.arange(np.min(x_flat), np.max(x_flat), self.interpolated_grid_resolution_m)
        y_lin = np.arange(np.min(y_flat), np.max(y_flat), self.interpolated_grid_resolution_m)
        self.interp_x_s, self.interp_y_s = np.meshgrid(x_lin, y_lin, indexing="ij")
        self.orig_grid_points = np.column_stack((x_flat, y_flat))
        self.interp_grid_points = np.column_stack((self.interp_x_s.flatten(), self.interp_y_s.flatten()))

    def _update_vel_history(self, current_vel: np.ndarray = np.array([0, 0, 0])) -> bool:
        """Validity gate (ref :352-375): bounds on every stored velocity, per-axis std, direction cosines."""
        self.history_avg_vel[0:-1, :] = self.history_avg_vel[1:, :]
        self.history_avg_vel[-1, :] = current_vel
        h = self.history_avg_vel
        within = np.all((self.min_vel <= np.abs(h)) & (np.abs(h) <= self.max_vel))
        stdev_ok = np.all(np.std(h, axis=0) <= self.max_vel_stdev)
        dirs = h / (np.linalg.norm(h, axis=1, keepdims=True) + 1e-6)
        consistent = np.all(np.dot(dirs, dirs.T) > 0.95)
        self.array_geometry_valid = within and stdev_ok and consistent
        return self.array_geometry_valid

    def _generate_frame_array_geometries(self, frame_vel=np.array([0, 0, 0]), frame_start_pose=np.array([0, 0, 0])):
        t = self.chirp_start_times_us
        frame_p = np.array([2 * t * 1e-6 * frame_vel[0] + frame_start_pose[0],
                            2 * t * 1e-6 * frame_vel[1] + frame_start_pose[1],
                            2 * t * 1e-6 * frame_vel[2] + frame_start_pose[2]])
        frame_end_pose = frame_start_pose + 2 * frame_vel * (-1 * self.frame_period_ms * 1e-3)
        return frame_p, frame_end_pose

    def _update_array_geometries(self, current_vel: np.ndarray = np.array([0, 0, 0])):
        self._update_vel_history(current_vel=current_vel)
        if self.array_geometry.shape[0] == 0:
            self.array_geometry = np.zeros(shape=(self.num_frames, 3, self._num_valid))
        pose = np.array([0, 0, 0])
        for frame_idx in range(self.num_frames - 1, -1, -1):
            frame_p, pose_next = self._generate_frame_array_geometries(
                frame_vel=self.history_avg_vel[frame_idx, :], frame_start_pose=pose)
            self.array_geometry[frame_idx, :, :] = frame_p
            pose = pose_next

    def _compute_beam_stearing_vectors(self):
        thetas, phis = np.meshgrid(self.az_angle_bins_rad, self.el_angle_bins_rad, indexing="ij")
        self.d = np.array([np.multiply(np.cos(thetas), np.cos(phis)),
                           np.multiply(np.sin(thetas), np.cos(phis)),
                           np.sin(phis)])

    def _init_out_resp(self):
        self.beamformed_resp = np.zeros(shape=(self.range_bins.shape[0], self.az_angle_bins_rad.shape[0],
                                               self.el_angle_bins_rad.shape[0]), dtype=complex)

    # ------------------------------------------------------------------ device work
    def _engine(self):
        if self._eng is None:
            from ._device import engine_for_cube_shape
            self._eng = engine_for_cube_shape(1, self.num_range_bins, 1)
        return self._eng

    def _push_frame(self, adc_cube):
        """Select receiver + valid chirps and append to the device ring buffer (ref :837-847)."""
        import torch
        eng = self._engine()
        sel = np.asarray(adc_cube)[self.receiver_idx][:, self.valid_chirps_mask]           # [S, Lv]
        new = torch.from_numpy(np.ascontiguousarray(sel, dtype=np.complex64)).to(eng.device)
        if self._hist_dev is None:
            self._hist_dev = torch.zeros((self.num_frames, self.num_range_bins, self._num_valid),
                                         dtype=torch.complex64, device=eng.device)
        self._hist_dev = torch.cat((self._hist_dev[1:], new[None]), dim=0)

    def compute_synthetic_response(self, array_geometry, g_begin: int = 0, g_end=None) -> np.ndarray:
        """Beamformed response indexed [range_bin, azimuth, elevation] (ref :543-585).
        ``g_begin/g_end`` restrict the flat voxel range (multi-GPU sharding); the host array keeps zeros elsewhere."""
        import torch
        eng = self._engine()
        if self._hist_dev is None:
            self._hist_dev = torch.zeros((self.num_frames, self.num_range_bins, self._num_valid),
                                         dtype=torch.complex64, device=eng.device)
        s = self.num_range_bins
        aperture = self._hist_dev.permute(1, 0, 2).reshape(s, -1).contiguous()            # X[s, f*Lv + c]
        pos = np.ascontiguousarray(np.asarray(array_geometry).transpose((1, 0, 2)).reshape(3, -1), dtype=np.float64)
        dirs = np.ascontiguousarray(self.d.reshape(3, -1), dtype=np.float64)              # g = a*E + e
        g_total = dirs.shape[1]
        g_end = g_total if g_end is None else g_end
        out = eng.sar_image(aperture, torch.from_numpy(pos).to(eng.device), torch.from_numpy(dirs).to(eng.device),
                            self.lambda_m, g_begin, g_end)
        self._resp_dev = out
        flat = self.beamformed_resp.reshape(s, g_total)
        flat[:, g_begin:g_end] = out.cpu().numpy().astype(np.complex128)
        return self.beamformed_resp

    def get_interpolated_response_cart(self, diagonal: int = 0) -> np.ndarray:
        """Polar -> cartesian resampling at el index 0 (ref :587-613: scipy griddata, method="linear", fill 0) in
        ``radar_polar_to_cart``.  Every polar cell is a cyclic trapezoid, so the Delaunay split scipy/Qhull picks per cell
        is arbitrary; ``diagonal`` fixes ours (each reference pixel equals one of the two splits, tests/test_gpu_sarcal.py)."""
        import torch
        if self.el_angle_bins_rad.shape[0] != 1:
            # the reference hands griddata S*A*E points but S*A values here
            raise ValueError("different number of values and points")
        eng = self._engine()
        if self._resp_dev is None or self._resp_dev.shape[1] != self.az_angle_bins_rad.shape[0]:
            image = torch.from_numpy(np.ascontiguousarray(self.beamformed_resp[:, :, 0], dtype=np.complex64)).to(eng.device)
        else:
            image = self._resp_dev
        az = np.asarray(self.az_angle_bins_rad, dtype=np.float64)
        order = np.argsort(az, kind="stable")
        if np.any(np.diff(az[order]) <= 0) or np.any(np.diff(self.range_bins) <= 0):
            raise ValueError("range and azimuth bins must be distinct")
        nx, ny = self.interp_x_s.shape
        x0, y0 = (self.interp_x_s[0, 0], self.interp_y_s[0, 0]) if nx and ny else (0.0, 0.0)
        res = self.interpolated_grid_resolution_m
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(eng.device)
        out = eng.polar_to_cart(image, 1, up(np.asarray(self.range_bins, dtype=np.float64)), up(az[order]), x0, res, nx, y0,
                                res, ny, diagonal=diagonal, az_order=up(order.astype(np.int32)))
        self.interpolated_beamformed_resp = out.cpu().numpy().astype(np.complex128)
        return self.interpolated_beamformed_resp

    def perform_array_calibration(self):
        """Self-calibration of the synthetic-array geometry against three strong scatterers (ref :673-816) in
        ``radar_sar_calibrate``, from the device history and the image of the last ``compute_synthetic_response``.
        -> corrected geometry (3, N) (also stored per frame in ``array_geometry_calibrated``), or an empty array when
        fewer than three targets qualify."""
        import torch
        eng = self._engine()
        s = self.num_range_bins
        aperture = self._hist_dev.permute(1, 0, 2).reshape(s, -1).contiguous()
        n = aperture.shape[1]
        pos = np.ascontiguousarray(np.asarray(self.array_geometry).transpose((1, 0, 2)).reshape(3, -1), dtype=np.float64)
        dirs = np.ascontiguousarray(self.d.reshape(3, -1), dtype=np.float64)
        if self._resp_dev is None or self._resp_dev.shape[1] != dirs.shape[1]:
            raise ValueError("calibration needs the full beamformed response (compute_synthetic_response over all voxels)")
        up = lambda a: torch.from_numpy(a).to(eng.device)
        new, targets = eng.sar_calibrate(aperture, up(np.hamming(n)), up(pos), self._resp_dev, up(dirs),
                                         self.el_angle_bins_rad.shape[0], self.lambda_m)
        targets = targets.cpu().numpy()
        self.calibration_targets = [(int(targets[2 * t]), int(targets[2 * t + 1])) for t in range(3) if targets[2 * t] >= 0]
        if targets[6] < 0:
            raise IndexError("fewer than three range peaks for calibration")      # the reference indexes past its peak list
        if targets[6] < 3:
            print("No suitable calibration targets found.")
            return np.empty(shape=0)
        new = new.cpu().numpy()
        self.array_geometry_calibrated = new.reshape(3, self.num_frames, -1).transpose(1, 0, 2)
        return new

    def process(self, adc_cube: np.ndarray, current_vel) -> np.ndarray:
        if isinstance(current_vel, list):
            current_vel = np.array(current_vel)
        self._push_frame(adc_cube)                       # the frame is appended BEFORE the gate (ref :845-853)
        self._update_array_geometries(current_vel=current_vel)
        if not self.array_geometry_valid:
            return np.empty(shape=0)
        resp = self.compute_synthetic_response(self.array_geometry)
        if self.enable_calibration:
            for _ in range(self.num_calibration_iters):
                if self.perform_array_calibration().shape[0] != 0:
                    resp = self.compute_synthetic_response(self.array_geometry_calibrated)
                else:
                    self.array_geometry_calibrated = self.array_geometry
                    break
        self.get_interpolated_response_cart()
        return resp

    def reset(self):
        self._hist_dev = None
        self._reset_vel_history()
        return super().reset()
