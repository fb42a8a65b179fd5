"""Doppler-azimuth response (reference: processors/doppler_azimuth_resp.py:7-491) on the CUDA kernels.

``process``: antenna subset, Hann over samples / chirps (and over the subset's antennas for standard arrays with
TDM-MIMO), range FFT, range-window rows, then per range row the zero-padded (chirp x antenna) 2-D spectrum, magnitude,
valid-angle columns, mean over range.  The coarse variant starts from the windowed range-Doppler cube
(``radar_range_doppler``) and runs ``radar_doppler_azimuth``; the precise variant replaces the Doppler FFT by the zoom
transform (scipy ZoomFFT in the reference = ``radar_chirp_zoom``, a direct evaluation of the same sums) for the negative
and positive velocity halves.  All bin axes and the half/zero decisions are host fp64 with the reference's expressions.
"""
from typing import Union

import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import is_device_tensor, stage_cube, to_host_f64
from ._processor import _Processor


class DopplerAzimuthProcessor(_Processor):
    """Coordinate system: x forward, y left."""

    def __init__(self, config_manager: ConfigManager, num_angle_bins: int = 64,
                 valid_angle_range: np.ndarray = np.array([np.deg2rad(-60), np.deg2rad(60)]),
                 min_zoom_fft_vel_span=0.1, **kwargs) -> None:
        self.num_range_bins = None
        self.range_bins = None
        self.vel_bins: np.ndarray = None
        self.zoomed_vel_bins: np.ndarray = None
        self.min_zoom_fft_vel_span = min_zoom_fft_vel_span
        self.num_angle_bins = num_angle_bins
        self.phase_shifts: np.ndarray = None
        self.angle_bins: np.ndarray = None
        self.valid_angle_range = valid_angle_range
        self.valid_angle_mask = None
        self.valid_angle_bins = None
        super().__init__(config_manager)

    def configure(self):
        cm = self.config_manager
        self.num_range_bins = cm.get_num_adc_samples(profile_idx=0)
        # the reference's numpy expressions, float-arange quirks included (ref :50-76)
        self.range_bins = np.arange(start=0, step=cm.range_res_m, stop=cm.range_max_m - cm.range_res_m / 2 + 1e-3)
        self.vel_bins = np.arange(start=-1 * cm.vel_max_m_s, stop=cm.vel_max_m_s - cm.vel_res_m_s + 1e-3,
                                  step=cm.vel_res_m_s)
        self.num_rx_antennas = cm.num_rx_antennas
        a = self.num_angle_bins
        self.phase_shifts = np.arange(start=np.pi, stop=-np.pi - 2 * np.pi / (a - 1), step=-2 * np.pi / (a - 1))
        self.phase_shifts[-1] = -1 * np.pi
        with np.errstate(invalid="ignore"):
            self.angle_bins = np.arcsin(self.phase_shifts / np.pi)
        self.valid_angle_mask = (self.angle_bins >= self.valid_angle_range[0]) & \
            (self.angle_bins <= self.valid_angle_range[1])
        self.valid_angle_bins = self.angle_bins[self.valid_angle_mask]

    # ------------------------------------------------------------------ host logic of the precise variant
    def set_zoomed_fft_vel_bins(self, vel_range):
        """ref :164-205"""
        num_vel_bins = self.vel_bins.size
        neg = np.linspace(start=vel_range[0], stop=min(-1e-4, vel_range[1]),
                          num=num_vel_bins if vel_range[0] <= 0 else 0, endpoint=False)
        pos = np.linspace(start=max(1e-4, vel_range[0]), stop=vel_range[1],
                          num=num_vel_bins if vel_range[1] > 0 else 0, endpoint=False)
        if neg.size > 0 and pos.size > 0:
            self.zoomed_vel_bins = np.concatenate((neg, pos))
        elif neg.size > 0:
            self.zoomed_vel_bins = neg
        elif pos.size > 0:
            self.zoomed_vel_bins = pos
        else:
            self.zoomed_vel_bins = np.array([])
        return self.zoomed_vel_bins

    def _zoom_turns(self, vel_range, num_samples):
        """Turns per chirp of ZoomFFT(num_samples, [f0, f1], fs=2*fs) as the reference sets it up (ref :146-156)."""
        cm = self.config_manager
        fs = 1 / (cm.vel_res_m_s)
        freq_start = vel_range[0] * fs / cm.vel_max_m_s
        freq_stop = vel_range[1] * fs / cm.vel_max_m_s
        return (freq_start + np.arange(num_samples) * (freq_stop - freq_start) / num_samples) / (fs * 2)

    def _precise_plan(self, vel_range):
        """Clamp / widen the velocity range IN PLACE (the reference mutates the caller's array, ref :243-258), set the
        zoomed bins and return per half (n_bins, turns or None)."""
        cm = self.config_manager
        vel_range[0] = max(vel_range[0], -1 * cm.vel_max_m_s)
        vel_range[1] = min(vel_range[1], cm.vel_max_m_s)
        vel_spread = 2 * self.min_zoom_fft_vel_span
        if (vel_range[1] - vel_range[0]) < vel_spread:
            dist_to_max_bound = abs(vel_range[1] - cm.vel_max_m_s)
            dist_to_min_bound = abs(vel_range[0] - (-1 * cm.vel_max_m_s))
            if dist_to_max_bound > dist_to_min_bound:
                vel_range[1] = vel_range[0] + vel_spread
            elif dist_to_min_bound > dist_to_max_bound:
                vel_range[0] = vel_range[1] - vel_spread
        self.set_zoomed_fft_vel_bins(vel_range)
        halves = []
        neg = self.zoomed_vel_bins[self.zoomed_vel_bins <= 0]
        if len(neg) > 0 and np.abs(max(neg) - min(neg)) > self.min_zoom_fft_vel_span:
            rng = np.array([np.min(neg), np.max(neg)]) + 2 * cm.vel_max_m_s
            halves.append((len(neg), self._zoom_turns(rng, len(neg))))
        else:
            halves.append((len(neg), None))
        pos = self.zoomed_vel_bins[self.zoomed_vel_bins > 0]
        if len(pos) > 0 and np.abs(max(pos) - min(pos)) > self.min_zoom_fft_vel_span:
            rng = np.array([np.min(pos), np.max(pos)])
            halves.append((len(pos), self._zoom_turns(rng, len(pos))))
        else:
            halves.append((len(pos), None))
        return halves

    # ------------------------------------------------------------------ device path
    def _range_rows(self, range_window):
        if range_window.size == 0:
            range_window = np.array([0, self.config_manager.range_max_m])
        mask = (self.range_bins >= range_window[0]) & (self.range_bins <= range_window[1])
        idx = np.nonzero(mask)[0]
        if idx.size == 0:
            return 0, 0
        return int(idx[0]), int(idx[-1]) + 1      # the mask of an ascending axis is one contiguous run

    def process(self, adc_cube, rx_antennas: Union[np.ndarray, list] = [], range_window: Union[np.ndarray, list] = [],
                shift_angle: bool = True, use_precise_fft: bool = False,
                precise_vel_range: Union[np.ndarray, list] = np.array([-0.25, 0.25]), **kwargs):
        """-> Doppler-azimuth response indexed [vel, valid angle], averaged over the range rows kept."""
        import torch
        rx_antennas = np.array([]) if rx_antennas is None else np.asarray(rx_antennas)
        range_window = np.array([]) if range_window is None else np.asarray(range_window)
        if precise_vel_range is None:
            precise_vel_range = np.array([-0.25, 0.25])
        elif isinstance(precise_vel_range, list):
            precise_vel_range = np.array(precise_vel_range)
        eng, dev, squeeze = stage_cube(adc_cube, self.config_manager)
        ants = [int(a) % eng.V for a in rx_antennas.ravel()] if rx_antennas.size > 0 else None
        n_ant = len(ants) if ants is not None else eng.V
        cm = self.config_manager
        weight = np.hanning(n_ant) if (cm.array_geometry == "standard" and cm.virtual_antennas_enabled) else None
        k0, k1 = self._range_rows(range_window)
        a = int(self.num_angle_bins)
        from .. import _native as nat
        win = nat.RADAR_WIN_RANGE | nat.RADAR_WIN_DOPPLER
        if not use_precise_fft:
            rd, _ = eng.range_doppler(dev, ant_list=ants, window_flags=win, want_rd=True)
            resp = eng.doppler_azimuth(rd, k0, k1, ant_weight=weight, num_angle_bins=a, shift_angle=shift_angle)
        else:
            halves = self._precise_plan(precise_vel_range)
            rc, _ = eng.range_doppler(dev, ant_list=ants, window_flags=win | nat.RADAR_RD_RANGE_ONLY, want_rd=True)
            parts = []
            for m, turns in halves:
                if turns is None:
                    parts.append(torch.zeros((dev.shape[0], m, a), dtype=torch.float32, device=eng.device))
                    continue
                if m != eng.L:
                    raise ValueError(f"zoom transform over {m} of {eng.L} chirps is not supported")
                z = eng.chirp_zoom(rc, turns, k0, k1)
                parts.append(eng.doppler_azimuth(z, 0, k1 - k0, ant_weight=weight, num_angle_bins=a,
                                                 shift_angle=shift_angle))
            resp = torch.cat(parts, dim=1)
        if not squeeze:
            return resp[:, :, torch.from_numpy(self.valid_angle_mask).to(eng.device)]
        return to_host_f64(resp[0])[:, self.valid_angle_mask]

    # ------------------------------------------------------------------ peak helpers (ref :336-417)
    def _floor_db(self, resp_mag, min_threshold_dB):
        db = 20 * np.log10(np.abs(np.asarray(resp_mag, dtype=np.float64)) + 1e-12)
        floor = np.max(db) - min_threshold_dB
        db[db <= floor] = floor
        return db

    def _strongest_peak_per_row(self, rows_db, prominence_db):
        """scipy.signal.find_peaks(row, prominence=...) then the strongest peak, through radar_find_peaks."""
        import torch
        from ..engine import RadarEngine
        eng = RadarEngine.for_config(self.config_manager)
        prof = torch.from_numpy(np.ascontiguousarray(rows_db, dtype=np.float32)).to(eng.device)
        idx, _, cnt = eng.find_peaks(prof, in_db=True, prominence_db=float(prominence_db), threshold_db=float("inf"),
                                     max_peaks=1)
        return idx[:, 0].cpu().numpy(), cnt.cpu().numpy()

    def detect_peaks_rows(self, doppler_azimuth_resp_mag: np.ndarray, vel_bins: np.ndarray,
                          min_threshold_dB: float = 30.0):
        """Per velocity row the strongest peak with >= 4 dB prominence -> Nx2 (angle rad, velocity)."""
        db = self._floor_db(doppler_azimuth_resp_mag, min_threshold_dB)
        best, cnt = self._strongest_peak_per_row(db, 4.0)
        rows = np.nonzero(cnt > 0)[0]
        return np.stack([self.valid_angle_bins[best[rows]], np.asarray(vel_bins)[rows]], axis=1)

    def detect_peak_zero_az(self, doppler_azimuth_resp_mag: np.ndarray, vel_bins: np.ndarray,
                            min_threshold_dB: float = 30.0):
        """Strongest local maximum of the zero-azimuth column -> [0, velocity] or empty (0, 2)."""
        db = self._floor_db(doppler_azimuth_resp_mag, min_threshold_dB)
        zero_az_col = np.argmin(np.abs(self.valid_angle_bins))
        best, cnt = self._strongest_peak_per_row(db[:, zero_az_col][None, :], 0.0)
        if cnt[0] > 0:
            return np.array([0.0, np.asarray(vel_bins)[best[0]]])
        return np.empty(shape=(0, 2))
