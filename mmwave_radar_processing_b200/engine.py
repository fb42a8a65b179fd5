"""Device-level wrapper over the C ABI: one ``RadarEngine`` per (radar shape, CUDA device).

PyTorch is plumbing here (device buffers + streams); every computation is a call into
``libradar_b200.so``.  All methods take and return CUDA tensors and enqueue on the current torch stream.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import _native as nat


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def cfar_params(kind: str, num_train, num_guard, pfa: float = 0.0, rho: float = 0.0, alpha: float = 0.0,
                os_rank: int = 0) -> nat.CfarParams:
    """Build the C struct.  2-D detectors pass (range, doppler) pairs, 1-D detectors pass ints."""
    p = nat.CfarParams()
    p.kind = nat.CFAR_KIND[kind]
    if isinstance(num_train, (tuple, list, np.ndarray)):
        p.train_r, p.train_d = int(num_train[0]), int(num_train[1])
        p.guard_r, p.guard_d = int(num_guard[0]), int(num_guard[1])
    else:
        p.train_r, p.train_d = 0, int(num_train)
        p.guard_r, p.guard_d = 0, int(num_guard)
    p.os_rank = int(os_rank)
    p.pfa, p.rho, p.alpha = float(pfa), float(rho), float(alpha)
    return p


class RadarEngine:
    """Owns a ``radar_ctx``.  Shapes: R rx, S samples, cpl TX slots, L loops; V = R*cpl, C = cpl*L."""

    def __init__(self, num_rx: int, num_samples: int, chirps_per_loop: int, num_loops: int, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("mmwave_radar_processing_b200 needs a CUDA device (no CPU fallback)")
        self.lib = nat.load_library()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.R, self.S, self.cpl, self.L = int(num_rx), int(num_samples), int(chirps_per_loop), int(num_loops)
        self.V, self.C = self.R * self.cpl, self.cpl * self.L
        shape = nat.RadarShape(self.R, self.S, self.cpl, self.L)
        handle = C.c_void_p()
        nat.check(self.lib.radar_ctx_create(C.byref(shape), self.device.index, C.byref(handle)), None)
        self.ctx = handle
        self._int_cache = {}
        self._axes_key = None
        self._angle_key = None

    @classmethod
    def for_config(cls, cfg, device=None) -> "RadarEngine":
        key = (cfg.num_rx_antennas, cfg.get_num_adc_samples(0), cfg.chirp_cfgs_per_loop, cfg.frameCfg_loops,
               str(device) if device is not None else f"cuda:{torch.cuda.current_device()}")
        eng = _ENGINES.get(key)
        if eng is None:
            eng = cls(*key[:4], device=device)
            _ENGINES[key] = eng
        return eng

    def __del__(self):
        try:
            if getattr(self, "ctx", None):
                self.lib.radar_ctx_destroy(self.ctx)
                self.ctx = None
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass

    # ------------------------------------------------------------------ helpers
    def _ints(self, values: Optional[Sequence[int]]):
        """Small int32 device array, cached by content (antenna lists, bin selections)."""
        if values is None:
            return None
        key = tuple(int(v) for v in values)
        t = self._int_cache.get(key)
        if t is None:
            t = torch.tensor(key, dtype=torch.int32, device=self.device) if key else \
                torch.zeros(1, dtype=torch.int32, device=self.device)
            self._int_cache[key] = t
        return t

    def _in_kind(self, adc: torch.Tensor) -> int:
        if adc.dtype == torch.int16:
            exp = (self.C, self.R, self.S, 2)
            kind = nat.RADAR_IN_IQ16
        elif adc.dtype == torch.complex64:
            exp = (self.V, self.S, self.L)
            kind = nat.RADAR_IN_CUBE64
        else:
            raise ValueError(f"ADC tensor must be int16 IQ or complex64 cube, got {adc.dtype}")
        if tuple(adc.shape[-len(exp):]) != exp or adc.dim() != len(exp) + 1:
            raise ValueError(f"ADC tensor shape {tuple(adc.shape)} != (frames,)+{exp}")
        if not adc.is_contiguous() or adc.device != self.device:
            raise ValueError("ADC tensor must be contiguous and on the engine's device")
        return kind

    def set_axes(self, range_bins: np.ndarray, vel_bins: np.ndarray):
        rb = np.ascontiguousarray(range_bins, dtype=np.float64)
        vb = np.ascontiguousarray(vel_bins, dtype=np.float64)
        key = (rb.tobytes(), vb.tobytes())
        if key != self._axes_key:
            dp = C.POINTER(C.c_double)
            nat.check(self.lib.radar_ctx_set_axes(self.ctx, rb.ctypes.data_as(dp), rb.shape[0],
                                                  vb.ctypes.data_as(dp), vb.shape[0]), self.ctx)
            self._axes_key = key

    def set_angle_bins(self, angle_bins: np.ndarray):
        ab = np.ascontiguousarray(angle_bins, dtype=np.float64)
        key = ab.tobytes()
        if key != self._angle_key:
            nat.check(self.lib.radar_ctx_set_angle_bins(self.ctx, ab.ctypes.data_as(C.POINTER(C.c_double)),
                                                        ab.shape[0]), self.ctx)
            self._angle_key = key

    def set_point_mirrors(self, pointers, cap_rows: int = 0):
        """Device pointers (ints) every packed point row is also stored to (peer-mapped gathered buffers); [] switches off."""
        arr = (C.c_uint64 * max(1, len(pointers)))(*[int(p) for p in pointers])
        nat.check(self.lib.radar_ctx_set_point_mirrors(self.ctx, C.cast(arr, C.c_void_p), len(pointers), int(cap_rows)),
                  self.ctx)

    def set_image_mirrors(self, pointers, row_stride: int = 0, col_offset: int = 0):
        """Device pointers (ints) of [S][row_stride] complex64 images every cell ``sar_image`` writes is also stored to, at
        column ``col_offset`` + (direction - g_begin) (peer-mapped full images of all ranks); [] switches off."""
        arr = (C.c_uint64 * max(1, len(pointers)))(*[int(p) for p in pointers])
        nat.check(self.lib.radar_ctx_set_image_mirrors(self.ctx, C.cast(arr, C.c_void_p), len(pointers), int(row_stride),
                                                       int(col_offset)), self.ctx)

    # ------------------------------------------------------------------ kernels
    def last_rd_kernel(self) -> str:
        """Name of the kernel the last range_doppler call launched."""
        return (self.lib.radar_last_rd_kernel(self.ctx) or b"").decode()

    def range_doppler(self, adc: torch.Tensor, ant_list=None, window_flags: int = 3, want_rd: bool = True,
                      mag_antenna: Optional[int] = None, rd_out=None, mag_out=None):
        """-> (rd complex64 [F, n_ant, S, L] or None, mag float32 [F,S,L] / [F,n_ant,S,L] or None)."""
        kind = self._in_kind(adc)
        f = adc.shape[0]
        n_ant = len(ant_list) if ant_list is not None else self.V
        ants = self._ints(ant_list)
        if want_rd and rd_out is None:
            rd_out = torch.empty((f, n_ant, self.S, self.L), dtype=torch.complex64, device=self.device)
        if mag_antenna is not None and mag_out is None:
            shape = (f, self.S, self.L) if mag_antenna >= 0 else (f, n_ant, self.S, self.L)
            mag_out = torch.empty(shape, dtype=torch.float32, device=self.device)
        nat.check(self.lib.radar_range_doppler(
            self.ctx, nat.ptr(adc), kind, f, nat.ptr(ants), n_ant, int(window_flags),
            nat.ptr(rd_out if want_rd else None), nat.ptr(mag_out), int(mag_antenna if mag_antenna is not None else 0),
            _stream_ptr()), self.ctx)
        return (rd_out if want_rd else None), mag_out

    def range_profile(self, adc: torch.Tensor, loop_idx: int = 0, out=None) -> torch.Tensor:
        kind = self._in_kind(adc)
        f = adc.shape[0]
        if out is None:
            out = torch.empty((f, self.S), dtype=torch.float32, device=self.device)
        nat.check(self.lib.radar_range_profile(self.ctx, nat.ptr(adc), kind, f, int(loop_idx), nat.ptr(out),
                                               _stream_ptr()), self.ctx)
        return out

    def range_zoom(self, adc: torch.Tensor, start_m, stop_m, range_max_m: float, loop_idx: int = 0, out=None) -> torch.Tensor:
        """Zoom range profile of every frame (RangeProcessor.zoom_fft): float32 [F, S].  start_m / stop_m: scalars or
        per-frame sequences / device float64 tensors."""
        kind = self._in_kind(adc)
        f = adc.shape[0]

        def per_frame(v):
            if isinstance(v, torch.Tensor):
                t = v.to(device=self.device, dtype=torch.float64).reshape(-1)
            else:
                t = torch.as_tensor(v, dtype=torch.float64).reshape(-1).to(self.device)
            if t.numel() == 1:
                t = t.expand(f)
            if t.numel() != f:
                raise ValueError(f"zoom window must hold 1 or {f} values, got {t.numel()}")
            return t.contiguous()
        lo, hi = per_frame(start_m), per_frame(stop_m)
        if out is None:
            out = torch.empty((f, self.S), dtype=torch.float32, device=self.device)
        nat.check(self.lib.radar_range_zoom(self.ctx, nat.ptr(adc), kind, f, int(loop_idx), nat.ptr(lo), nat.ptr(hi),
                                            float(range_max_m), nat.ptr(out), _stream_ptr()), self.ctx)
        return out

    def find_peaks(self, profiles: torch.Tensor, in_db: bool = False, prominence_db: float = 6.0,
                   threshold_db: float = 20.0, max_peaks: int = 3):
        """Prominence peak picking (RangeProcessor.find_peaks) on float32 profiles [n, len] ->
        (idx int32 [n, max_peaks] (-1 padded), db float64 [n, max_peaks], counts int32 [n])."""
        if profiles.dtype != torch.float32 or profiles.dim() != 2 or not profiles.is_contiguous():
            raise ValueError("profiles must be a contiguous float32 [n, len] tensor")
        n, m = profiles.shape
        idx = torch.empty((n, max_peaks), dtype=torch.int32, device=self.device)
        db = torch.empty((n, max_peaks), dtype=torch.float64, device=self.device)
        cnt = torch.empty((n,), dtype=torch.int32, device=self.device)
        nat.check(self.lib.radar_find_peaks(self.ctx, nat.ptr(profiles), n, m, 1 if in_db else 0, float(prominence_db),
                                            float(threshold_db), int(max_peaks), nat.ptr(idx), nat.ptr(db), nat.ptr(cnt),
                                            _stream_ptr()), self.ctx)
        return idx, db, cnt

    def cfar(self, maps: torch.Tensor, params: nat.CfarParams, one_d: bool = False, row_select=None,
             want_noise: bool = False, want_mask: bool = False, max_dets: Optional[int] = None,
             dets_out=None, count_out=None):
        """CFAR + ordered compaction on float32 maps [n, rows, cols].

        -> dict(noise=f64 or None, mask=u8 or None, dets=int32 [n,max_dets,2], counts=int32 [n])."""
        if maps.dtype != torch.float32 or maps.dim() != 3 or not maps.is_contiguous():
            raise ValueError("maps must be a contiguous float32 [n, rows, cols] tensor")
        n, rows, cols = maps.shape
        if max_dets is None:
            max_dets = rows * cols
        noise = torch.empty((n, rows, cols), dtype=torch.float64, device=self.device) if want_noise else None
        mask = torch.empty((n, rows, cols), dtype=torch.uint8, device=self.device) if want_mask else None
        if dets_out is None:
            dets_out = torch.empty((n, max_dets, 2), dtype=torch.int32, device=self.device)
        if count_out is None:
            count_out = torch.empty((n,), dtype=torch.int32, device=self.device)
        if one_d:
            nat.check(self.lib.radar_cfar_1d_rows(self.ctx, nat.ptr(maps), n, rows, cols, C.byref(params),
                                                  nat.ptr(row_select), nat.ptr(noise), nat.ptr(mask), nat.ptr(dets_out),
                                                  nat.ptr(count_out), int(max_dets), _stream_ptr()), self.ctx)
        else:
            nat.check(self.lib.radar_cfar_2d(self.ctx, nat.ptr(maps), n, rows, cols, C.byref(params), nat.ptr(noise),
                                             nat.ptr(mask), nat.ptr(dets_out), nat.ptr(count_out), int(max_dets),
                                             _stream_ptr()), self.ctx)
        return dict(noise=noise, mask=mask, dets=dets_out, counts=count_out)

    def det_offsets(self, counts: torch.Tensor, max_dets: int, out=None) -> torch.Tensor:
        """Exclusive scan of min(counts, max_dets) -> int32 [F+1]; out[F] is the batch's number of points."""
        f = counts.shape[0]
        if out is None:
            out = torch.empty((f + 1,), dtype=torch.int32, device=self.device)
        nat.check(self.lib.radar_det_offsets(self.ctx, nat.ptr(counts), f, int(max_dets), nat.ptr(out), _stream_ptr()),
                  self.ctx)
        return out

    def point_cloud(self, rd: torch.Tensor, dets: torch.Tensor, counts: torch.Tensor, az_idx, el_idx,
                    shift_az: bool = True, shift_el: bool = False, num_angle_bins: int = 64,
                    points_out=None, want_bins: bool = True, offsets=None):
        """-> (points float64, az_bin int32 | None, el_bin | None).

        offsets None: padded layout, points [F,max_dets,4], bins [F,max_dets].
        offsets int32 [F+1] (``det_offsets``): packed layout, points [rows,4] with rows >= offsets[F]."""
        f, max_dets = dets.shape[0], dets.shape[1]
        if rd.dtype != torch.complex64 or tuple(rd.shape) != (f, self.V, self.S, self.L):
            raise ValueError("rd must be complex64 [F, V, S, L] for all V antennas")
        if points_out is None:
            if offsets is not None:
                raise ValueError("packed output needs a caller-sized points_out")
            points_out = torch.zeros((f, max_dets, 4), dtype=torch.float64, device=self.device)
        rows = points_out.numel() // 4
        az_bin = torch.zeros((rows,) if offsets is not None else (f, max_dets), dtype=torch.int32,
                             device=self.device) if want_bins else None
        el_bin = torch.zeros_like(az_bin) if want_bins else None
        az, el = self._ints(az_idx), self._ints(el_idx)
        nat.check(self.lib.radar_point_cloud(
            self.ctx, nat.ptr(rd), f, nat.ptr(dets), nat.ptr(counts), max_dets,
            nat.ptr(az), len(az_idx), nat.ptr(el), len(el_idx), int(bool(shift_az)), int(bool(shift_el)),
            int(num_angle_bins), nat.ptr(offsets), nat.ptr(points_out), nat.ptr(az_bin), nat.ptr(el_bin),
            _stream_ptr()), self.ctx)
        return points_out, az_bin, el_bin

    def range_angle(self, adc: torch.Tensor, loop_idx: int = 0, ant_list=None, num_angle_bins: int = 64,
                    windowing: bool = True, out=None) -> torch.Tensor:
        kind = self._in_kind(adc)
        f = adc.shape[0]
        n_ant = len(ant_list) if ant_list is not None else self.V
        if out is None:
            out = torch.empty((f, self.S, num_angle_bins), dtype=torch.float32, device=self.device)
        flags = (nat.RADAR_WIN_RANGE | nat.RADAR_WIN_ANTENNA) if windowing else 0
        nat.check(self.lib.radar_range_angle(self.ctx, nat.ptr(adc), kind, f, int(loop_idx), nat.ptr(self._ints(ant_list)),
                                             n_ant, int(num_angle_bins), flags, nat.ptr(out), _stream_ptr()), self.ctx)
        return out

    def dbs(self, adc: torch.Tensor, vel_bin, ang_bin, ant_list=None, num_angle_bins: int = 64, out=None) -> torch.Tensor:
        kind = self._in_kind(adc)
        f = adc.shape[0]
        n_ant = len(ant_list) if ant_list is not None else self.V
        n_out = len(vel_bin)
        if out is None:
            out = torch.empty((f, self.S, n_out), dtype=torch.float32, device=self.device)
        nat.check(self.lib.radar_dbs(self.ctx, nat.ptr(adc), kind, f, nat.ptr(self._ints(ant_list)), n_ant,
                                     int(num_angle_bins), nat.ptr(self._ints(vel_bin)), nat.ptr(self._ints(ang_bin)),
                                     n_out, nat.ptr(out), _stream_ptr()), self.ctx)
        return out

    def _floats(self, values):
        """Small float32 device array cached by content (antenna windows)."""
        if values is None:
            return None
        key = ("f32",) + tuple(float(v) for v in values)
        t = self._int_cache.get(key)
        if t is None:
            t = torch.tensor(key[1:], dtype=torch.float32, device=self.device)
            self._int_cache[key] = t
        return t

    def doppler_azimuth(self, cube: torch.Tensor, k_begin: int = 0, k_end: Optional[int] = None, ant_weight=None,
                        num_angle_bins: int = 64, shift_angle: bool = True, out=None) -> torch.Tensor:
        """cube complex64 [F, n_ant, rows, cols] (range-Doppler or zoom cube) -> float32 [F, cols, A]:
        mean over rows [k_begin, k_end) of |zero-padded angle DFT| (DopplerAzimuthProcessor, coarse / precise tail)."""
        if cube.dtype != torch.complex64 or cube.dim() != 4 or not cube.is_contiguous():
            raise ValueError("cube must be a contiguous complex64 [F, n_ant, rows, cols] tensor")
        f, n_ant, rows, cols = cube.shape
        k_end = rows if k_end is None else k_end
        if out is None:
            out = torch.empty((f, cols, num_angle_bins), dtype=torch.float32, device=self.device)
        nat.check(self.lib.radar_doppler_azimuth(self.ctx, nat.ptr(cube), f, n_ant, rows, cols, int(k_begin), int(k_end),
                                                 nat.ptr(self._floats(ant_weight)), int(num_angle_bins),
                                                 int(bool(shift_angle)), nat.ptr(out), _stream_ptr()), self.ctx)
        return out

    def chirp_zoom(self, range_cube: torch.Tensor, turns, k_begin: int = 0, k_end: Optional[int] = None) -> torch.Tensor:
        """range_cube complex64 [F, n_ant, S, L] (range transform only) -> complex64 [F, n_ant, k_end-k_begin, J]:
        X[j] = sum_l x[l] exp(-2 pi i l turns[j]) (scipy ZoomFFT over the chirp axis)."""
        if range_cube.dtype != torch.complex64 or range_cube.dim() != 4 or tuple(range_cube.shape[2:]) != (self.S, self.L):
            raise ValueError("range_cube must be complex64 [F, n_ant, S, L]")
        f, n_ant = range_cube.shape[:2]
        k_end = self.S if k_end is None else k_end
        tt = torch.as_tensor(np.ascontiguousarray(turns, dtype=np.float64)).to(self.device)
        out = torch.empty((f, n_ant, k_end - k_begin, tt.numel()), dtype=torch.complex64, device=self.device)
        nat.check(self.lib.radar_chirp_zoom(self.ctx, nat.ptr(range_cube.contiguous()), f, n_ant, int(k_begin), int(k_end),
                                            nat.ptr(tt), tt.numel(), nat.ptr(out), _stream_ptr()), self.ctx)
        return out

    def chirp_zoom_static(self, range_cube: torch.Tensor, turns: torch.Tensor, out: torch.Tensor, k_begin: int = 0,
                          k_end: Optional[int] = None) -> torch.Tensor:
        """``chirp_zoom`` through ``radar_chirp_zoom_static``: the twiddle table lives in a static device array of the
        library (L * J <= 65536), so neither the library nor this method allocates - the caller passes ``turns`` (float64
        device tensor [J]) and ``out`` (complex64 [F, n_ant, k_end-k_begin, J]).  No counterpart in the reference; calls
        that overlap on the GPU race on the shared table."""
        if range_cube.dtype != torch.complex64 or range_cube.dim() != 4 or tuple(range_cube.shape[2:]) != (self.S, self.L) \
                or not range_cube.is_contiguous():
            raise ValueError("range_cube must be a contiguous complex64 [F, n_ant, S, L] tensor")
        f, n_ant = range_cube.shape[:2]
        k_end = self.S if k_end is None else k_end
        if turns.dtype != torch.float64 or turns.dim() != 1 or not turns.is_cuda or not turns.is_contiguous():
            raise ValueError("turns must be a contiguous float64 device tensor [J]")
        if out.dtype != torch.complex64 or tuple(out.shape) != (f, n_ant, k_end - k_begin, turns.numel()) or not out.is_contiguous():
            raise ValueError("out must be a contiguous complex64 [F, n_ant, k_end-k_begin, J] tensor")
        nat.check(self.lib.radar_chirp_zoom_static(self.ctx, nat.ptr(range_cube), f, n_ant, int(k_begin), int(k_end),
                                                   nat.ptr(turns), turns.numel(), nat.ptr(out), _stream_ptr()), self.ctx)
        return out

    def row_reduce(self, maps: torch.Tensor, row_select=None, op: str = "max") -> torch.Tensor:
        """maps float32 [n, rows, cols] -> float32 [n, cols]: max / mean over the selected rows."""
        if maps.dtype != torch.float32 or maps.dim() != 3 or not maps.is_contiguous():
            raise ValueError("maps must be a contiguous float32 [n, rows, cols] tensor")
        n, rows, cols = maps.shape
        sel = None
        if row_select is not None:
            sel = torch.as_tensor(np.ascontiguousarray(row_select, dtype=np.uint8)).to(self.device)
            if sel.numel() != rows:
                raise ValueError("row_select must hold one flag per row")
        out = torch.empty((n, cols), dtype=torch.float32, device=self.device)
        nat.check(self.lib.radar_row_reduce(self.ctx, nat.ptr(maps), n, rows, cols, nat.ptr(sel),
                                            nat.RADAR_REDUCE_MAX if op == "max" else nat.RADAR_REDUCE_MEAN,
                                            nat.ptr(out), _stream_ptr()), self.ctx)
        return out

    def ego_velocity(self, points: torch.Tensor, offsets: torch.Tensor, dims: int, table, residual_threshold: float = 0.15,
                     stop_probability: float = 0.99):
        """Packed point clouds float64 [rows, 4] + offsets int32 [F+1] -> (vel float64 [F, 3], r2 [F], inlier
        fraction [F], status int32 [F]); ``table`` is a point_cloud_processing.ransac_subsets.SubsetTable that already
        holds a block for every point count of the batch (status -1 marks frames it does not cover)."""
        if points.dtype != torch.float64 or points.dim() != 2 or points.shape[1] != 4 or not points.is_contiguous():
            raise ValueError("points must be a contiguous float64 [rows, 4] tensor")
        if offsets.dtype != torch.int32 or offsets.dim() != 1:
            raise ValueError("offsets must be int32 [F+1]")
        f = offsets.shape[0] - 1
        blocks, index, max_n = table.device_arrays(self.device)
        vel = torch.empty((f, 3), dtype=torch.float64, device=self.device)
        r2 = torch.empty((f,), dtype=torch.float64, device=self.device)
        frac = torch.empty((f,), dtype=torch.float64, device=self.device)
        status = torch.empty((f,), dtype=torch.int32, device=self.device)
        nat.check(self.lib.radar_ego_velocity(self.ctx, nat.ptr(points), nat.ptr(offsets.contiguous()), f, int(dims),
                                              nat.ptr(blocks), nat.ptr(index), int(max_n), int(table.max_trials),
                                              int(table.min_samples), float(residual_threshold), float(stop_probability),
                                              nat.ptr(vel), nat.ptr(r2), nat.ptr(frac), nat.ptr(status), _stream_ptr()),
                  self.ctx)
        return vel, r2, frac, status

    def sar_image(self, aperture: torch.Tensor, positions: torch.Tensor, dirs: torch.Tensor, lambda_m: float,
                  g_begin: int = 0, g_end: Optional[int] = None, out=None) -> torch.Tensor:
        """aperture complex64 [S,N], positions float64 [3,N], dirs float64 [3,G] -> complex64 [S, g_end-g_begin]."""
        s, n = aperture.shape
        g = dirs.shape[1]
        g_end = g if g_end is None else g_end
        if aperture.dtype != torch.complex64 or positions.dtype != torch.float64 or dirs.dtype != torch.float64:
            raise ValueError("aperture complex64, positions/dirs float64 expected")
        if tuple(positions.shape) != (3, n) or dirs.shape[0] != 3:
            raise ValueError("positions must be [3,N], dirs [3,G]")
        if out is None:
            out = torch.empty((s, g_end - g_begin), dtype=torch.complex64, device=self.device)
        nat.check(self.lib.radar_sar_image(self.ctx, nat.ptr(aperture.contiguous()), s, n, nat.ptr(positions.contiguous()),
                                           nat.ptr(dirs.contiguous()), g, float(lambda_m), int(g_begin), int(g_end),
                                           nat.ptr(out), _stream_ptr()), self.ctx)
        return out


    def sar_calibrate(self, aperture: torch.Tensor, window: torch.Tensor, positions: torch.Tensor, image: torch.Tensor,
                      dirs: torch.Tensor, n_elev: int, lambda_m: float):
        """aperture complex64 [S,N], window float64 [N], positions float64 [3,N], image complex64 [S,G] (all directions),
        dirs float64 [3,G] -> (calibrated positions float64 [3,N], targets int32 [7]: (range, azimuth) x 3 + status)."""
        s, n = aperture.shape
        g = dirs.shape[1]
        if aperture.dtype != torch.complex64 or image.dtype != torch.complex64:
            raise ValueError("aperture / image must be complex64")
        if any(t.dtype != torch.float64 for t in (window, positions, dirs)):
            raise ValueError("window / positions / dirs must be float64")
        if tuple(positions.shape) != (3, n) or tuple(image.shape) != (s, g) or window.numel() != n or dirs.shape[0] != 3:
            raise ValueError("shapes: positions [3,N], window [N], image [S,G], dirs [3,G]")
        out = torch.empty((3, n), dtype=torch.float64, device=self.device)
        targets = torch.empty((7,), dtype=torch.int32, device=self.device)
        nat.check(self.lib.radar_sar_calibrate(self.ctx, nat.ptr(aperture.contiguous()), s, n, nat.ptr(window.contiguous()),
                                               nat.ptr(positions.contiguous()), nat.ptr(image.contiguous()), g, int(n_elev),
                                               nat.ptr(dirs.contiguous()), float(lambda_m), nat.ptr(out), nat.ptr(targets),
                                               _stream_ptr()), self.ctx)
        return out, targets


    def polar_to_cart(self, image: torch.Tensor, n_elev: int, range_bins: torch.Tensor, az_bins: torch.Tensor, x0: float,
                      dx: float, nx: int, y0: float, dy: float, ny: int, diagonal: int = 0, az_order=None,
                      want_inside: bool = False):
        """image complex64 [S,G] -> complex64 [nx,ny] piecewise-linear resampling of image[:, az, el 0] onto the grid
        (x0 + ix dx, y0 + iy dy) (+ uint8 inside mask).  range_bins / az_bins float64, ascending."""
        s, g = image.shape
        if image.dtype != torch.complex64 or range_bins.dtype != torch.float64 or az_bins.dtype != torch.float64:
            raise ValueError("image complex64, range_bins / az_bins float64 expected")
        if range_bins.numel() != s:
            raise ValueError("range_bins must have one entry per image row")
        out = torch.empty((nx, ny), dtype=torch.complex64, device=self.device)
        inside = torch.empty((nx, ny), dtype=torch.uint8, device=self.device) if want_inside else None
        order = None if az_order is None else az_order.to(torch.int32).contiguous()
        nat.check(self.lib.radar_polar_to_cart(self.ctx, nat.ptr(image.contiguous()), s, g, int(n_elev),
                                               nat.ptr(range_bins.contiguous()), nat.ptr(az_bins.contiguous()),
                                               nat.ptr(order) if order is not None else None, az_bins.numel(), float(x0),
                                               float(dx), int(nx), float(y0), float(dy), int(ny), int(diagonal), nat.ptr(out),
                                               nat.ptr(inside) if inside is not None else None, _stream_ptr()), self.ctx)
        return (out, inside) if want_inside else out


_ENGINES = {}


def utility_engine(device=None) -> RadarEngine:
    """Shape-independent engine used by the stand-alone CFAR detector classes."""
    key = ("util", str(device) if device is not None else f"cuda:{torch.cuda.current_device()}")
    eng = _ENGINES.get(key)
    if eng is None:
        eng = RadarEngine(1, 1, 1, 1, device=device)
        _ENGINES[key] = eng
    return eng
