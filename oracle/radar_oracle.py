"""CPU oracle for the ADC-cube -> point-cloud path.   *** TEST INFRASTRUCTURE ONLY ***

A plain-numpy (fp64) restatement of what the reference computes on this path, written as
pure functions so it can be pinned and used as a checker.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference`` legs may
import it; nothing under ``mmwave_radar_processing_b200/`` does (the product path fails
loudly when the CUDA library is missing).

Pinning: the reference's own tests hold no numeric golden vectors for this path
(SURVEY.md section 8c).  The pin is ``tests/golden/*.npz``: outputs of the reference classes
themselves, produced in the build container by ``tests/golden/make_golden.py`` (which
imports ``/root/reference`` read-only) on seeded synthetic cubes; ``tests/test_oracle_golden.py``
checks every function below against them.  The reference's two runnable known-answer checks
(``tests/test_point_cloud_angles.py``, the CA/GO/SO part of ``tests/verify_detectors_manual.py``)
are restated in ``tests/test_reference_known_answers.py``.

Third-party arithmetic the reference leans on (numpy 1.26.4 pinned in its poetry.lock;
pocketfft ``np.fft``, ``np.hanning``/``np.hamming``, ``np.partition``, ``sliding_window_view``)
is re-expressed from the published definitions: forward DFT ``X[k]=sum x[n] e^{-2 pi i kn/N}``,
``hanning(M)[n]=0.5-0.5cos(2 pi n/(M-1))``, ``hamming(M)[n]=0.54-0.46cos(2 pi n/(M-1))``,
k-th order statistic = k-th entry of the sorted training set.

Every function cites the reference lines it follows (paths relative to
``/root/reference/mmwave_radar_processing/``).
"""
from __future__ import annotations

import numpy as np
from numpy.lib.stride_tricks import sliding_window_view

C_M_S = 299792458.0


# --------------------------------------------------------------------------------------
# windows / small helpers
# --------------------------------------------------------------------------------------
def hann(m: int) -> np.ndarray:
    """``np.hanning`` (symmetric, zero end points; hanning(1) == [1.])."""
    if m < 1:
        return np.zeros(0)
    if m == 1:
        return np.ones(1)
    n = np.arange(m)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * n / (m - 1))


def hamm(m: int) -> np.ndarray:
    """``np.hamming``."""
    if m < 1:
        return np.zeros(0)
    if m == 1:
        return np.ones(1)
    n = np.arange(m)
    return 0.54 - 0.46 * np.cos(2.0 * np.pi * n / (m - 1))


def shift_index(n: int) -> np.ndarray:
    """``fftshift`` as a gather: out[i] = in[(i - n//2) mod n]."""
    return (np.arange(n) - n // 2) % n


# --------------------------------------------------------------------------------------
# a1: int16 I/Q -> complex cube, TDM virtual array
# --------------------------------------------------------------------------------------
def iq_to_cube(iq: np.ndarray) -> np.ndarray:
    """Device ADC layout -> the reference's complex cube.

    ``iq``: int16 ``[chirp][rx][sample][2]`` (I then Q), the layout the CUDA path ingests
    (DESIGN.md "ADC layout").  Returns complex128 ``(rx, samples, chirps)`` - the array
    ``CpslDS.get_radar_adc_data`` hands to processors (processors/_processor.py:58).
    The cast int16 -> float64 is exact.
    """
    iq = np.asarray(iq)
    assert iq.dtype == np.int16 and iq.ndim == 4 and iq.shape[-1] == 2
    z = iq[..., 0].astype(np.float64) + 1j * iq[..., 1].astype(np.float64)   # (C, R, S)
    return np.ascontiguousarray(z.transpose(1, 2, 0))                       # (R, S, C)


def virtual_array(cube: np.ndarray, num_rx: int, cpl: int, loops: int) -> np.ndarray:
    """TDM-MIMO regroup (processors/virtual_array_reformater.py:44-65).

    virtual channel ``v = tx_slot*num_rx + rx`` takes chirps ``c = loop*cpl + tx_slot``.
    """
    s = cube.shape[1]
    out = np.zeros((num_rx * cpl, s, loops), dtype=complex)
    for slot in range(cpl):
        out[slot * num_rx:(slot + 1) * num_rx] = cube[:num_rx, :, slot::cpl][:, :, :loops]
    return out


# --------------------------------------------------------------------------------------
# a2-a5: range-Doppler
# --------------------------------------------------------------------------------------
def rd_axes(vel_max, vel_res, range_max, range_res):
    """Bin axes of RangeDopplerProcessor.configure (processors/range_doppler_resp.py:33-47)."""
    vel_bins = np.arange(start=-1 * vel_max, stop=vel_max - vel_res + 1e-3, step=vel_res)
    range_bins = np.arange(start=0, step=range_res, stop=range_max - range_res / 2 + 1e-3)
    return range_bins, vel_bins


def range_doppler(cube_v: np.ndarray) -> np.ndarray:
    """Complex RD cube for all antennas (processors/range_doppler_resp.py:49-103).

    Hann over samples x Hann over chirps, unnormalised forward 2-D DFT over (S, L),
    ``fftshift`` on the Doppler axis only.
    """
    _, s, l = cube_v.shape
    w = hann(s)[None, :, None] * hann(l)[None, None, :]
    spec = np.fft.fft(np.fft.fft(cube_v * w, axis=1), axis=2)
    return spec[:, :, shift_index(l)]


def rd_magnitude0(rd: np.ndarray) -> np.ndarray:
    """CFAR input: |RD| of virtual antenna 0 only (range_doppler_detection/range_doppler_detector.py:72-80)."""
    return np.abs(rd[0])


def coarse_range_profile(cube_v: np.ndarray, chirp_idx: int = 0) -> np.ndarray:
    """RangeProcessor.coarse_fft (processors/range_resp.py:32-57): one chirp, Hann(S), |FFT|, mean over antennas."""
    x = cube_v[:, :, chirp_idx] * hann(cube_v.shape[1])
    return np.mean(np.abs(np.fft.fft(x, axis=1)), axis=0)


# --------------------------------------------------------------------------------------
# a6-a9: CFAR family
# --------------------------------------------------------------------------------------
def ca_alpha(n_train_cells: int, pfa: float) -> float:
    """alpha = N (pfa^(-1/N) - 1)  (detectors/base.py:155-169, :281-293)."""
    return n_train_cells * (pfa ** (-1.0 / n_train_cells) - 1.0)


def os_rank(rho: float, n_train_cells: int) -> int:
    """k = clip(int(rho*N), 1, N)  (detectors/os_cfar.py:25-27, :131-132)."""
    return max(1, min(int(rho * n_train_cells), n_train_cells))


def _train_mask_1d(t: int, g: int) -> np.ndarray:
    m = np.ones(2 * (t + g) + 1, dtype=bool)
    m[t:t + 2 * g + 1] = False
    return m


def cfar_1d(x: np.ndarray, kind: str, num_train: int, num_guard: int, pfa: float = 0.0,
            rho: float = 0.0, alpha: float = 0.0):
    """1-D CFAR thresholds (detectors/ca_cfar.py:11-77, os_cfar.py:29-86, go_so_cfar.py:11-123).

    Returns ``(thresholds, noise, mask)``; outside the valid interior thresholds = +inf, noise = 0.
    ``kind`` in {"ca","go","so","os"}.  GO/SO use alpha computed with N = num_train (one side).
    """
    x = np.asarray(x, dtype=np.float64)
    n = x.shape[0]
    thr = np.full(n, np.inf)
    noise = np.zeros(n)
    t, g = int(num_train), int(num_guard)
    win = 2 * (t + g) + 1
    if n >= win:
        w = sliding_window_view(x, win)                      # (n-win+1, win)
        left, right = w[:, :t], w[:, t + 2 * g + 1:]
        if kind == "ca":
            est = np.mean(w[:, _train_mask_1d(t, g)], axis=1)
            a = ca_alpha(2 * t, pfa)
        elif kind == "go":
            est = np.maximum(np.mean(left, axis=1), np.mean(right, axis=1))
            a = ca_alpha(t, pfa)
        elif kind == "so":
            est = np.minimum(np.mean(left, axis=1), np.mean(right, axis=1))
            a = ca_alpha(t, pfa)
        elif kind == "os":
            k = os_rank(rho, 2 * t)
            est = np.partition(np.concatenate((left, right), axis=1), k - 1, axis=1)[:, k - 1]
            a = alpha
        else:
            raise ValueError(kind)
        c = t + g
        thr[c:c + est.shape[0]] = a * est
        noise[c:c + est.shape[0]] = est
    return thr, noise, x > thr


def cfar_2d(x: np.ndarray, kind: str, num_train, num_guard, pfa: float = 0.0,
            rho: float = 0.0, alpha: float = 0.0):
    """2-D CFAR thresholds (detectors/ca_cfar.py:85-155, os_cfar.py:134-195; decision base.py:208-230).

    Training set = outer (2(Tr+Gr)+1)x(2(Td+Gd)+1) rectangle minus the centre (2Gr+1)x(2Gd+1).
    Returns ``(thresholds, noise, mask)``.
    """
    x = np.asarray(x, dtype=np.float64)
    r, d = x.shape
    tr, td = int(num_train[0]), int(num_train[1])
    gr, gd = int(num_guard[0]), int(num_guard[1])
    wr, wd = 2 * (tr + gr) + 1, 2 * (td + gd) + 1
    thr = np.full((r, d), np.inf)
    noise = np.zeros((r, d))
    if r >= wr and d >= wd:
        w = sliding_window_view(x, (wr, wd))                 # (r', d', wr, wd)
        m = np.ones((wr, wd), dtype=bool)
        m[tr:tr + 2 * gr + 1, td:td + 2 * gd + 1] = False
        n_train = int(m.sum())
        if kind == "ca":
            est = (w * m).sum(axis=(2, 3)) / n_train
            a = ca_alpha(n_train, pfa)
        elif kind == "os":
            k = os_rank(rho, n_train)
            est = np.partition(w[..., m], k - 1, axis=-1)[..., k - 1]
            a = alpha
        else:
            raise ValueError(kind)
        cr, cd = tr + gr, td + gd
        thr[cr:cr + est.shape[0], cd:cd + est.shape[1]] = a * est
        noise[cr:cr + est.shape[0], cd:cd + est.shape[1]] = est
    return thr, noise, x > thr


def mask_to_dets(mask: np.ndarray) -> np.ndarray:
    """Row-major (range, Doppler) detection list, ``(0,2)`` int when empty
    (detectors/base.py:226-230; range_doppler_detection/range_doppler_detector_2d.py:61-66)."""
    rows, cols = np.nonzero(mask)
    if rows.size == 0:
        return np.empty((0, 2), dtype=int)
    return np.stack((rows, cols), axis=1).astype(int)


def detect_sequential(cube_v, rd_mag0, rng_kind, rng_params, vel_kind, vel_params) -> np.ndarray:
    """RangeDopplerDetectorSequential.process (range_doppler_detection/range_doppler_detector_sequential.py:72-108)."""
    prof = coarse_range_profile(cube_v, 0)
    _, _, rmask = cfar_1d(prof, rng_kind, **rng_params)
    out = []
    for ri in np.nonzero(rmask)[0]:
        _, _, vmask = cfar_1d(rd_mag0[ri], vel_kind, **vel_params)
        out.extend((int(ri), int(di)) for di in np.nonzero(vmask)[0])
    return np.array(out, dtype=int) if out else np.empty((0, 2), dtype=int)


def detect_range_band(rd_mag0, range_bins, altitude_m, vel_kind, vel_params) -> np.ndarray:
    """CFAR part of RangeDopplerGroundDetector.process
    (range_doppler_detection/range_doppler_ground_detector.py:84-119) for a given altitude."""
    lo = int(np.argmin(np.abs(range_bins - altitude_m)))
    max_rng = min(np.max(range_bins), altitude_m / np.cos(np.deg2rad(60)))
    hi = int(np.argmin(np.abs(range_bins - max_rng)))
    rows = [lo] if hi == lo else list(range(lo, hi + 1))
    out = []
    for ri in rows:
        _, _, vmask = cfar_1d(rd_mag0[ri], vel_kind, **vel_params)
        out.extend((int(ri), int(di)) for di in np.nonzero(vmask)[0])
    return np.array(out, dtype=int) if out else np.empty((0, 2), dtype=int)


# --------------------------------------------------------------------------------------
# a13-a15: angle estimation and point cloud
# --------------------------------------------------------------------------------------
def angle_axis(num_angle_bins: int):
    """phase_shifts / angle_bins of PointCloudGenerator.configure and RangeAngleProcessor.configure
    (processors/point_cloud_generator.py:89-105, range_angle_resp.py:38-48).  Length is A or A+1
    depending on float rounding of ``np.arange`` - keep the identical expression."""
    a = num_angle_bins
    phase = np.arange(start=np.pi, stop=-np.pi - 2 * np.pi / (a - 1), step=-2 * np.pi / (a - 1))
    phase[-1] = -1 * np.pi
    return phase, np.arcsin(phase / np.pi)


def angle_indices(rd: np.ndarray, r_idx, d_idx, ant_idxs, num_angle_bins: int, shift: bool) -> np.ndarray:
    """Per-detection zero-padded angle FFT + first-max argmax
    (processors/point_cloud_generator.py:143-214).  Returns the argmax indices (N,)."""
    ant = np.asarray(ant_idxs, dtype=int)
    n = len(r_idx)
    if ant.size == 0:
        return np.zeros(n, dtype=int)
    snap = rd[ant][:, r_idx, d_idx].T                                       # (N, n_ant)
    pad = np.zeros((n, num_angle_bins), dtype=complex)
    pad[:, :ant.size] = snap
    spec = np.fft.fft(pad, axis=1)
    if shift:
        spec = spec[:, shift_index(num_angle_bins)]
    return np.argmax(np.abs(spec), axis=1)


def point_cloud(rd, dets, range_bins, vel_bins, az_idxs, el_idxs, num_angle_bins=64,
                shift_az=True, shift_el=False):
    """PointCloudGenerator.process after detection (processors/point_cloud_generator.py:108-140,
    :216-248).  Returns ``(points (N,4) f64, az_index (N,), el_index (N,))``."""
    if dets.shape[0] == 0:
        return np.empty((0, 4)), np.zeros(0, dtype=int), np.zeros(0, dtype=int)
    r_idx = dets[:, 0].astype(int)
    d_idx = dets[:, 1].astype(int)
    _, bins = angle_axis(num_angle_bins)
    az_i = angle_indices(rd, r_idx, d_idx, az_idxs, num_angle_bins, shift_az)
    el_i = angle_indices(rd, r_idx, d_idx, el_idxs, num_angle_bins, shift_el)
    az = bins[az_i] if len(az_idxs) else np.zeros(len(r_idx))
    el = bins[el_i] if len(el_idxs) else np.zeros(len(r_idx))
    rng = range_bins[r_idx]
    vel = vel_bins[d_idx]
    x = rng * np.cos(el) * np.cos(az)
    y = rng * np.cos(el) * np.sin(az)
    z = rng * np.sin(el)
    return np.column_stack((x, y, z, vel)), az_i, el_i


# --------------------------------------------------------------------------------------
# a16: range-angle heat map
# --------------------------------------------------------------------------------------
def range_angle(cube_v, num_angle_bins=64, chirp_idx=0, rx_antennas=(), perform_windowing=True):
    """RangeAngleProcessor.process (processors/range_angle_resp.py:68-122).

    Hann over samples and over ALL antennas of the cube is applied BEFORE the antenna subset;
    the subset's rank is the antenna's slot in the zero-padded angle FFT.
    """
    v, s, _ = cube_v.shape
    x = cube_v[:, :, chirp_idx]
    if perform_windowing:
        x = x * hann(s)[None, :] * hann(v)[:, None]
    ants = np.asarray(rx_antennas, dtype=int)
    if ants.size > 0:
        x = x[ants]
    data = np.zeros((s, num_angle_bins), dtype=complex)
    data[:, :x.shape[0]] = x.T
    spec = np.fft.fft(np.fft.fft(data, axis=0), axis=1)
    return np.abs(spec[:, shift_index(num_angle_bins)])


# --------------------------------------------------------------------------------------
# a17: Doppler beam sharpening
# --------------------------------------------------------------------------------------
def dbs_bin_selection(angle_bins_coarse, angle_bins_dbs, vel_bins, velocity_ned):
    """Host-side (Doppler bin, coarse angle bin) choice per output angle
    (processors/range_angle_resp_dbs_enhanced.py:200-258)."""
    vel_bin = np.empty(len(angle_bins_dbs), dtype=int)
    ang_bin = np.empty(len(angle_bins_dbs), dtype=int)
    v = np.asarray(velocity_ned, dtype=float)
    for i, th in enumerate(angle_bins_dbs):
        rhat = np.array([np.cos(th), np.sin(th), 0])
        dop = -1 * np.dot(rhat / np.linalg.norm(rhat), v)
        vel_bin[i] = np.argmin(np.abs(vel_bins - dop))
        ang_bin[i] = np.argmin(np.abs(angle_bins_coarse - th))
    return vel_bin, ang_bin


def dbs_enhanced(cube_v, velocity_ned, vel_bins, num_angle_bins=64, num_angle_bins_dbs=64, rx_antennas=()):
    """RangeAngleProcessorDBSEnhanced.process_dbs_enhanced
    (processors/range_angle_resp_dbs_enhanced.py:137-198, :216-304): antennas are subset FIRST,
    then Hann over range, Doppler and the subset's antennas; 3-D FFT; pick one
    (angle, Doppler) cell per output angle."""
    ants = np.asarray(rx_antennas, dtype=int)
    x = cube_v[ants] if ants.size > 0 else cube_v
    n, s, l = x.shape
    a = num_angle_bins
    xr = np.fft.fft(x * hann(s)[None, :, None], axis=1)
    xd = np.fft.fft(xr * hann(l)[None, None, :], axis=2)[:, :, shift_index(l)]
    pad = np.zeros((a, s, l), dtype=complex)
    pad[:n] = xd * hann(n)[:, None, None]
    mag = np.abs(np.fft.fft(pad, axis=0)[shift_index(a)])
    _, coarse = angle_axis(a)
    fine = np.linspace(coarse[0], coarse[-1], num_angle_bins_dbs)
    vb, ab = dbs_bin_selection(coarse, fine, vel_bins, velocity_ned)
    return mag[ab, :, vb].T                                              # (S, A_dbs)


# --------------------------------------------------------------------------------------
# a18: multi-frame synthetic array (SAR)
# --------------------------------------------------------------------------------------
def sar_chirp_times_us(cpl, loops, chirp_period_us, chirp_cfg_idx=0, stride=1, start_index=0):
    """Valid-chirp mask and (negative) chirp start times, last chirp of the frame at 0
    (processors/simple_synthetic_array_beamformer_processor_multiFrame.py:207-234)."""
    n = cpl * loops
    cfg_of_chirp = np.tile(np.arange(start_index, start_index + cpl), loops)
    valid = np.nonzero(cfg_of_chirp == chirp_cfg_idx)[0][::stride]
    mask = np.zeros(n, dtype=bool)
    mask[valid] = True
    times = np.arange(n - 1, -1, -1) * -chirp_period_us
    return mask, times[mask]


def sar_geometry(history_vel, chirp_times_us, frame_period_ms):
    """Element positions chained backwards over frames (…multiFrame.py:377-471).
    ``history_vel`` (F,3) oldest->newest.  Returns (F,3,Lv)."""
    f = history_vel.shape[0]
    geo = np.zeros((f, 3, chirp_times_us.shape[0]))
    p0 = np.zeros(3)
    for i in range(f - 1, -1, -1):
        vel = history_vel[i]
        geo[i] = 2 * chirp_times_us[None, :] * 1e-6 * vel[:, None] + p0[:, None]
        p0 = p0 + 2 * vel * (-1 * frame_period_ms * 1e-3)
    return geo


def sar_velocity_gate(history_vel, min_vel, max_vel, max_vel_stdev) -> bool:
    """Validity gate of _update_vel_history (…multiFrame.py:352-375)."""
    h = np.asarray(history_vel, dtype=float)
    ok = np.all((min_vel <= np.abs(h)) & (np.abs(h) <= max_vel))
    ok = ok and np.all(np.std(h, axis=0) <= max_vel_stdev)
    d = h / (np.linalg.norm(h, axis=1, keepdims=True) + 1e-6)
    return bool(ok and np.all(d @ d.T > 0.95))


def sar_steering(az_rad, el_rad):
    """Unit pointing vectors d[(x,y,z), az, el] (…multiFrame.py:474-488)."""
    th, ph = np.meshgrid(np.asarray(az_rad), np.asarray(el_rad), indexing="ij")
    return np.array([np.cos(th) * np.cos(ph), np.sin(th) * np.cos(ph), np.sin(ph)])


def sar_response(history, geometry, d, lambda_m):
    """compute_synthetic_response (…multiFrame.py:499-585).

    ``history`` (F,S,Lv) complex oldest->newest, ``geometry`` (F,3,Lv), ``d`` (3,A,E).
    out[:,a,e] = FFT_S( hann(S) * sum_n hamming(N)[n] x[s,n] e^{+2 pi i d.p_n / lambda} ).
    """
    f, s, lv = history.shape
    x = history.transpose(1, 0, 2).reshape(s, f * lv) * hamm(f * lv)[None, :]
    p = geometry.transpose(1, 0, 2).reshape(3, f * lv)
    _, na, ne = d.shape
    steer = np.exp(1j * 2 * np.pi * np.einsum("cae,cn->nae", d, p) / lambda_m)      # (N, A, E)
    beams = np.einsum("sn,nae->sae", x, steer)
    return np.fft.fft(beams * hann(s)[:, None, None], axis=0)


# ------------------------------------------------------------------------------------------------
# zoom range profile, prominence peak picking, altimeter recurrence
# ------------------------------------------------------------------------------------------------
def zoom_range_profile(cube_v: np.ndarray, range_start_m: float, range_stop_m: float, range_max_m: float,
                       chirp_idx: int = 0):
    """RangeProcessor.zoom_fft (reference processors/range_resp.py:59-103).  scipy.signal.ZoomFFT(n, [f0, f1], fs=fs)
    with m = n output points and endpoint=False evaluates X[k] = sum_s x[s] exp(-2 pi i s f_k / fs),
    f_k = f0 + k (f1 - f0)/m (scipy 1.14 _czt.py: a = exp(2 pi i f0/fs), w = exp(-2 pi i (f1-f0)/(m fs))); with the
    reference's f = range * fs / range_max that is a phase of s * r_k / range_max turns, r_k = start + k (stop-start)/n.
    Restated as the direct sum (no chirp-Z convolution).  Returns (mean_v |X|, linspace(start, stop, n))."""
    x = np.asarray(cube_v)[:, :, chirp_idx].astype(np.complex128)
    n = x.shape[1]
    xw = x * hann(n)
    k = np.arange(n)
    r_k = range_start_m + k * (range_stop_m - range_start_m) / n
    s = np.arange(n)
    phase = np.outer(r_k / range_max_m, s)                      # turns, [k, s]
    kern = np.exp(-2j * np.pi * phase)
    zoom = xw @ kern.T                                           # [v, k]
    return np.mean(np.abs(zoom), axis=0), np.linspace(range_start_m, range_stop_m, n)


def local_maxima_1d(x: np.ndarray) -> np.ndarray:
    """scipy.signal._peak_finding_utils._local_maxima_1d (scipy 1.14): strict rise, plateau, strict fall -> the
    plateau's midpoint (left + right)//2; the first and last samples are never peaks."""
    x = np.asarray(x, dtype=np.float64)
    peaks = []
    i, i_max = 1, x.shape[0] - 1
    while i < i_max:
        if x[i - 1] < x[i]:
            i_ahead = i + 1
            while i_ahead < i_max and x[i_ahead] == x[i]:
                i_ahead += 1
            if x[i_ahead] < x[i]:
                peaks.append((i + i_ahead - 1) // 2)
                i = i_ahead
        i += 1
    return np.asarray(peaks, dtype=np.intp)


def peak_prominences(x: np.ndarray, peaks: np.ndarray) -> np.ndarray:
    """scipy.signal.peak_prominences with wlen=None (scipy 1.14 _peak_prominences): walk left / right from the peak
    until a sample is HIGHER than it (or the signal ends), keeping the minimum; prominence = peak - max(left min, right min)."""
    x = np.asarray(x, dtype=np.float64)
    out = np.empty(len(peaks), dtype=np.float64)
    for j, p in enumerate(peaks):
        i, left_min = p, x[p]
        while 0 <= i and x[i] <= x[p]:
            left_min = min(left_min, x[i])
            i -= 1
        i, right_min = p, x[p]
        while i <= x.shape[0] - 1 and x[i] <= x[p]:
            right_min = min(right_min, x[i])
            i += 1
        out[j] = x[p] - max(left_min, right_min)
    return out


def find_range_peaks(resp_db: np.ndarray, bins: np.ndarray, max_peaks: int = 3, threshold_db: float = 20.0,
                     prominence: float = 6.0):
    """RangeProcessor.find_peaks (reference processors/range_resp.py:105-150): scipy find_peaks(prominence=6), keep peaks
    within threshold_db of the strongest, strongest first, at most max_peaks.  Returns (bins[idx], resp_db[idx], idx)."""
    resp_db = np.asarray(resp_db, dtype=np.float64)
    peaks = local_maxima_1d(resp_db)
    if len(peaks):
        prom = peak_prominences(resp_db, peaks)
        peaks = peaks[prom >= prominence]
    if len(peaks) == 0:
        return np.array([]), np.array([]), np.array([], dtype=np.intp)
    vals = resp_db[peaks]
    keep = vals >= (np.max(vals) - threshold_db)
    peaks, vals = peaks[keep], vals[keep]
    top = peaks[np.argsort(vals)[::-1]][:max_peaks]
    return np.asarray(bins)[top], resp_db[top], top


def altimeter_step(state: dict, profile: np.ndarray, range_bins: np.ndarray, cube_v=None, range_max_m: float = 0.0,
                   precise: bool = False) -> float:
    """One Altimeter.process call (reference processors/altimeter.py:56-139).  state: min_altitude_m, zoom_search_region_m,
    altitude_search_limit_m, range_bias, measured, corrected (updated in place).  Returns the corrected altitude."""
    def ground_peak(peaks_m):
        if peaks_m.size > 0:
            ok = peaks_m[(peaks_m >= state["min_altitude_m"]) &
                         (np.abs(peaks_m - state["measured"]) <= state["altitude_search_limit_m"])]
            if ok.size > 0:
                return float(np.min(ok))
        return -1.0
    with np.errstate(divide="ignore"):
        db = 20 * np.log10(np.asarray(profile, dtype=np.float64))
    peaks_m, _, _ = find_range_peaks(db, range_bins, max_peaks=3)
    if peaks_m.size == 0:
        return state["corrected"]
    g = ground_peak(peaks_m)
    if g < 0:
        return state["corrected"]
    if precise:
        lo = max(1e-6, g - state["zoom_search_region_m"])
        hi = min(np.max(range_bins) - 1e-6, g + state["zoom_search_region_m"])
        zoom, zbins = zoom_range_profile(cube_v, lo, hi, range_max_m, 0)
        zp, _, _ = find_range_peaks(20 * np.log10(zoom), zbins, max_peaks=2)
        g = ground_peak(zp) if zp.size > 0 else -1.0
        if g < 0:
            return state["corrected"]
    state["measured"] = g
    state["corrected"] = g + state["range_bias"]
    return state["corrected"]


# ------------------------------------------------------------------------------------------------
# Doppler-azimuth (coarse + zoom), micro-Doppler, strip-map SAR
# ------------------------------------------------------------------------------------------------
def doppler_azimuth_windowed(cube_v: np.ndarray, rx_antennas=(), antenna_window: bool = False) -> np.ndarray:
    """DopplerAzimuthProcessor.process head (reference processors/doppler_azimuth_resp.py:460-466) +
    apply_hanning_windows (:84-100): the antenna subset is taken FIRST, then Hann over samples, Hann over chirps and -
    for standard arrays with TDM-MIMO (``antenna_window``) - Hann over the subset's antennas."""
    x = np.asarray(cube_v).astype(np.complex128)
    ants = np.asarray(rx_antennas, dtype=int).ravel()
    if ants.size > 0:
        x = x[ants]
    x = x * hann(x.shape[1])[None, :, None] * hann(x.shape[2])[None, None, :]
    if antenna_window:
        x = x * hann(x.shape[0])[:, None, None]
    return x


def doppler_azimuth_coarse(cube_v, range_bins, num_angle_bins=64, rx_antennas=(), range_window=(), shift_angle=True,
                           antenna_window=False, range_max_m=None):
    """Coarse Doppler-azimuth response BEFORE the valid-angle column selection (reference :102-128, :296-334, :491):
    range FFT, rows inside ``range_window`` (default [0, range_max]), zero-padded fft2 over (chirp, antenna), fftshift
    over Doppler (always) and angle (``shift_angle``), abs, mean over the kept range rows.  -> [L, A]."""
    x = doppler_azimuth_windowed(cube_v, rx_antennas, antenna_window)
    rf = np.fft.fft(x, axis=1)
    rw = np.asarray(range_window, dtype=float).ravel()
    if rw.size == 0:
        rw = np.array([0.0, range_max_m])
    keep = (range_bins >= rw[0]) & (range_bins <= rw[1])
    rf = rf[:, keep, :]
    n, k, l = rf.shape
    data = np.zeros((k, l, num_angle_bins), dtype=complex)
    data[:, :, :n] = np.transpose(rf, (1, 2, 0))
    spec = np.fft.fft(np.fft.fft(data, axis=1), axis=2)
    idx_d = shift_index(l)
    spec = spec[:, idx_d, :]
    if shift_angle:
        spec = spec[:, :, shift_index(num_angle_bins)]
    return np.mean(np.abs(spec), axis=0)


def zoom_vel_bins(vel_range, num_vel_bins: int):
    """DopplerAzimuthProcessor.set_zoomed_fft_vel_bins (reference :164-205)."""
    neg = np.linspace(vel_range[0], min(-1e-4, vel_range[1]), num_vel_bins if vel_range[0] <= 0 else 0, endpoint=False)
    pos = np.linspace(max(1e-4, vel_range[0]), vel_range[1], num_vel_bins if vel_range[1] > 0 else 0, endpoint=False)
    if neg.size > 0 and pos.size > 0:
        return np.concatenate((neg, pos))
    if neg.size > 0:
        return neg
    if pos.size > 0:
        return pos
    return np.array([])


def zoom_turns(vel_lo: float, vel_hi: float, m: int, vel_res: float, vel_max: float) -> np.ndarray:
    """Turns per chirp of the m output bins of DopplerAzimuthProcessor.zoom_fft (reference :130-162):
    fs = 1/vel_res, f = v*fs/vel_max, ZoomFFT(m, [f0, f1], fs=2*fs) -> nu_j = (f0 + j (f1-f0)/m)/(2 fs)."""
    fs = 1 / vel_res
    f0 = vel_lo * fs / vel_max
    f1 = vel_hi * fs / vel_max
    return (f0 + np.arange(m) * (f1 - f0) / m) / (fs * 2)


def doppler_azimuth_precise_plan(vel_range, vel_bins, vel_res, vel_max, min_zoom_span=0.1):
    """Host logic of DopplerAzimuthProcessor.precise_doppler_azimuth_fft (reference :207-294): clamp the velocity range,
    enforce the minimum spread, build the zoomed velocity bins, and decide for the negative and the positive half
    whether a zoom transform is computed (span > min_zoom_span) or the half is zero.
    -> (zoomed_vel_bins, [(n_bins, turns or None), (n_bins, turns or None)])."""
    vr = np.array(vel_range, dtype=float)
    vr[0] = max(vr[0], -1 * vel_max)
    vr[1] = min(vr[1], vel_max)
    spread = 2 * min_zoom_span
    if (vr[1] - vr[0]) < spread:
        to_max = abs(vr[1] - vel_max)
        to_min = abs(vr[0] - (-1 * vel_max))
        if to_max > to_min:
            vr[1] = vr[0] + spread
        elif to_min > to_max:
            vr[0] = vr[1] - spread
    zb = zoom_vel_bins(vr, vel_bins.size)
    halves = []
    neg = zb[zb <= 0]
    if len(neg) > 0 and np.abs(max(neg) - min(neg)) > min_zoom_span:
        halves.append((len(neg), zoom_turns(np.min(neg) + 2 * vel_max, np.max(neg) + 2 * vel_max, len(neg), vel_res, vel_max)))
    else:
        halves.append((len(neg), None))
    pos = zb[zb > 0]
    if len(pos) > 0 and np.abs(max(pos) - min(pos)) > min_zoom_span:
        halves.append((len(pos), zoom_turns(np.min(pos), np.max(pos), len(pos), vel_res, vel_max)))
    else:
        halves.append((len(pos), None))
    return zb, halves


def doppler_azimuth_precise(cube_v, range_bins, vel_bins, vel_res, vel_max, num_angle_bins=64, rx_antennas=(),
                            range_window=(), shift_angle=True, antenna_window=False, range_max_m=None,
                            vel_range=(-0.25, 0.25), min_zoom_span=0.1):
    """Precise (zoom) Doppler-azimuth response before the valid-angle selection (reference :130-162, :207-294, :491).
    Each computed half is a direct DFT over the first n_bins chirps at the zoom frequencies, then the zero-padded
    angle FFT, abs; halves are concatenated along Doppler; only the angle axis is shifted.  -> (resp [J, A], zoomed bins)."""
    x = doppler_azimuth_windowed(cube_v, rx_antennas, antenna_window)
    rf = np.fft.fft(x, axis=1)
    rw = np.asarray(range_window, dtype=float).ravel()
    if rw.size == 0:
        rw = np.array([0.0, range_max_m])
    keep = (range_bins >= rw[0]) & (range_bins <= rw[1])
    rf = rf[:, keep, :]
    n, k, l = rf.shape
    zb, halves = doppler_azimuth_precise_plan(vel_range, vel_bins, vel_res, vel_max, min_zoom_span)
    parts = []
    for m, turns in halves:
        if turns is None:
            parts.append(np.zeros((k, m, num_angle_bins)))
            continue
        y = rf[:, :, :m]                                                  # [n, k, m] first m chirps
        kern = np.exp(-2j * np.pi * np.outer(np.arange(m), turns))        # [chirp, j]
        z = np.einsum("nkl,lj->kjn", y, kern)
        data = np.zeros((k, m, num_angle_bins), dtype=complex)
        data[:, :, :n] = z
        parts.append(np.abs(np.fft.fft(data, axis=2)))
    resp = np.concatenate(parts, axis=1)
    if shift_angle:
        resp = resp[:, :, shift_index(num_angle_bins)]
    return np.mean(resp, axis=0), zb


def micro_doppler_column(cube_v, range_bins, target_ranges=(0.0, 1.0), rx_idx=0) -> np.ndarray:
    """One new column of MicroDopplerProcessor.process (reference processors/micro_doppler_resp.py:91-114):
    UN-windowed fft2 of one antenna, fftshift over Doppler, abs, max over the range rows inside target_ranges."""
    x = np.asarray(cube_v)[rx_idx].astype(np.complex128)
    resp = np.abs(np.fft.fft2(x, axes=(-2, -1))[:, shift_index(x.shape[1])])
    keep = (range_bins >= target_ranges[0]) & (range_bins <= target_ranges[1])
    return np.max(resp[keep, :], axis=0)


def strip_map_response(cube_v, rx_index=0) -> np.ndarray:
    """StripMapSARProcessor.process before the range / angle slices (reference processors/strip_map_SAR_processor.py:
    176-188): un-windowed fft2 of one antenna, fftshift over the chirp axis, complex."""
    x = np.asarray(cube_v)[rx_index].astype(np.complex128)
    return np.fft.fft2(x, axes=(-2, -1))[:, shift_index(x.shape[1])]


# ------------------------------------------------------------------------------------------------
# point cloud -> ego velocity (least squares inside RANSAC)
# ------------------------------------------------------------------------------------------------
_RANSAC_EPS = np.spacing(1)


def ransac_subsets(n_points: int, min_samples: int = 10, max_trials: int = 20, seed: int = 42,
                   policy: str = "permutation") -> np.ndarray:
    """Random subsets sklearn's RANSACRegressor.fit draws with random_state=seed: ``check_random_state(seed)`` builds a
    fresh ``np.random.RandomState(seed)`` on every fit, and every trial calls
    ``sample_without_replacement(n, min_samples, random_state=rng)`` exactly once, so the subsets depend only on
    (n_points, trial).  sklearn 1.9.0 (the build the goldens ran on; utils/_random.pyx) takes
    ``rng.permutation(n)[:k]`` for 0.01 < k/n < 0.99 ("permutation"); outside that band, and in the older releases the
    reference pins (1.5.2: "reservoir" for every ratio below 0.99 - not verifiable here), it is reservoir sampling
    (ratio >= 0.2) or tracking selection.  -> int [max_trials, min_samples]."""
    rng = np.random.RandomState(seed)
    out = np.empty((max_trials, min_samples), dtype=np.int64)
    ratio = min_samples / n_points
    for t in range(max_trials):
        if policy == "permutation" and 0.01 < ratio < 0.99:
            out[t] = rng.permutation(n_points)[:min_samples]
        elif ratio < 0.2 and not (policy == "reservoir" and ratio >= 0.01):
            sel, chosen = [], set()
            for _ in range(min_samples):
                j = rng.randint(n_points)
                while j in chosen:
                    j = rng.randint(n_points)
                chosen.add(j)
                sel.append(j)
            out[t] = sel
        else:
            res = np.arange(min_samples)
            for i in range(min_samples, n_points):
                j = rng.randint(0, i + 1)
                if j < min_samples:
                    res[j] = i
            out[t] = res
    return out


def _r2(y, pred):
    """sklearn.metrics.r2_score with force_finite=True on 1-D targets; nan below two samples."""
    if y.shape[0] < 2:
        return float("nan")
    num = float(np.sum((y - pred) ** 2))
    den = float(np.sum((y - np.mean(y)) ** 2))
    if den == 0.0:
        return 1.0 if num == 0.0 else 0.0
    return 1.0 - num / den


def _dynamic_max_trials(n_inliers, n_samples, min_samples, probability):
    ratio = n_inliers / float(n_samples)
    nom = max(_RANSAC_EPS, 1 - probability)
    denom = max(_RANSAC_EPS, 1 - ratio ** min_samples)
    if nom == 1:
        return 0
    if denom == 1:
        return float("inf")
    return abs(float(np.ceil(np.log(nom) / np.log(denom))))


def ego_velocity_ransac(points: np.ndarray, dims: int = 2, residual_threshold: float = 0.15, min_samples: int = 10,
                        max_trials: int = 20, seed: int = 42, policy: str = "permutation"):
    """VelocityEstimator.lsq_fit_ego_vel_ransac_points_2D / _3D (reference point_cloud_processing/vel_estimator.py:
    76-170): y = -v, H = xyz[:dims]/|xyz[:dims]|, RANSACRegressor(LinearRegression(fit_intercept=False),
    residual_threshold=0.15, random_state=42, max_trials=20, min_samples=10) restated from sklearn's fit loop
    (linear_model/_ransac.py): per trial a least-squares fit on the subset, inliers = |y - H c| <= threshold, keep the
    trial with more inliers (ties: the better R^2 on its inliers), shrink max_trials with the inlier ratio
    (stop_probability 0.99), final fit on the best inlier set.  Any ValueError inside sklearn (fewer points than
    min_samples, non-finite input, no consensus set) is swallowed by the reference -> zeros.
    -> (coef [dims] or zeros(2) on failure, r2 of the final model on its inliers (0 if <= 3 inliers), inlier fraction,
        inlier mask or None)."""
    pts = np.asarray(points, dtype=np.float64)
    fail = (np.array([0.0, 0.0]), 0.0, 0.0, None)
    if pts.shape[0] < min_samples:
        return fail
    y = -1 * pts[:, 3]
    with np.errstate(invalid="ignore", divide="ignore"):
        h = pts[:, 0:dims] / np.linalg.norm(pts[:, 0:dims], axis=1, keepdims=True)
    fit = ransac_fit(h, y, residual_threshold, min_samples, max_trials, seed, policy)
    return fail if fit is None else fit


def ransac_fit(h, y, residual_threshold, min_samples: int = 10, max_trials: int = 20, seed: int = 42,
               policy: str = "permutation"):
    """RANSACRegressor(LinearRegression(fit_intercept=False), random_state=seed).fit(h, y) restated from sklearn's
    linear_model/_ransac.py (see ego_velocity_ransac).  h [N, D], y [N].
    -> (coef [D], r2 of the final model on its inliers (0 if <= 3 inliers), inlier fraction, inlier mask), or None where
    sklearn raises ValueError."""
    h = np.asarray(h, dtype=np.float64).reshape(len(y), -1)
    y = np.asarray(y, dtype=np.float64)
    n = y.shape[0]
    if n < min_samples or not (np.all(np.isfinite(h)) and np.all(np.isfinite(y))):
        return None
    subsets = ransac_subsets(n, min_samples, max_trials, seed, policy)
    n_best, score_best, mask_best = 1, -np.inf, None
    trials, limit = 0, max_trials
    while trials < limit:
        idx = subsets[trials]
        trials += 1
        coef = np.linalg.lstsq(h[idx], y[idx], rcond=None)[0]
        mask = np.abs(y - h @ coef) <= residual_threshold
        n_in = int(mask.sum())
        if n_in < n_best:
            continue
        score = _r2(y[mask], h[mask] @ coef)
        if n_in == n_best and score < score_best:
            continue
        n_best, score_best, mask_best = n_in, score, mask
        limit = min(limit, _dynamic_max_trials(n_best, n, min_samples, 0.99))
    if mask_best is None:
        return None
    coef = np.linalg.lstsq(h[mask_best], y[mask_best], rcond=None)[0]
    r2 = _r2(y[mask_best], h[mask_best] @ coef) if mask_best.sum() > 3 else 0.0
    return coef, r2, mask_best.sum() / n, mask_best


# ---------------------------------------------------------------------------------------------------------------------
# ADC-domain ego velocity (reference processors/velocity_estimator.py) - test infrastructure like the rest of this file
# ---------------------------------------------------------------------------------------------------------------------
def da_valid_angles(num_angle_bins: int, valid_angle_range):
    """DopplerAzimuthProcessor.configure (reference processors/doppler_azimuth_resp.py:66-76) -> (mask, valid bins)."""
    a = num_angle_bins
    ph = np.arange(start=np.pi, stop=-np.pi - 2 * np.pi / (a - 1), step=-2 * np.pi / (a - 1))
    ph[-1] = -np.pi
    with np.errstate(invalid="ignore"):
        ang = np.arcsin(ph / np.pi)
    mask = (ang >= valid_angle_range[0]) & (ang <= valid_angle_range[1])
    return mask, ang[mask]


def _da_floor_db(resp, thr_db):
    db = 20 * np.log10(np.abs(resp) + 1e-12)
    floor = np.max(db) - thr_db
    db[db <= floor] = floor
    return db


def da_peaks_rows(resp, vel_bins, valid_angle_bins, thr_db=30.0):
    """detect_peaks_rows (reference doppler_azimuth_resp.py:336-377): per Doppler row, scipy find_peaks(prominence=4) on the
    floored dB row, strongest survivor -> rows of (angle, velocity)."""
    db = _da_floor_db(resp, thr_db)
    out = []
    for r, row in enumerate(db):
        pk = local_maxima_1d(row)
        if len(pk):
            pk = pk[peak_prominences(row, pk) >= 4.0]
        if len(pk):
            out.append((valid_angle_bins[pk[np.argmax(row[pk])]], vel_bins[r]))
    return np.array(out, dtype=np.float64).reshape(-1, 2)


def da_peak_zero_az(resp, vel_bins, valid_angle_bins, thr_db=30.0):
    """detect_peak_zero_az (reference doppler_azimuth_resp.py:379-417): strongest local maximum down the zero-azimuth column."""
    col = _da_floor_db(resp, thr_db)[:, np.argmin(np.abs(valid_angle_bins))]
    pk = local_maxima_1d(col)
    if len(pk):
        return np.array([0.0, vel_bins[pk[np.argmax(col[pk])]]])
    return np.empty((0, 2))


def adc_velocity_sets(geometry: str, virtual: bool):
    """Antenna sets of compute_azimuth_response / compute_elevation_response (reference velocity_estimator.py:129-264)
    -> (azimuth sets, elevation sets)."""
    if geometry == "standard":
        return [list(range(8 if virtual else 4))], []
    if virtual:
        return [[0, 3, 4, 7], [1, 2, 5, 6]], [[10, 11, 6, 7], [9, 8, 5, 4]]
    return [[0, 3], [1, 2]], [[1, 0], [3, 4]]


def adc_fit_vy(peaks, vx):
    """lsq_fit_ego_vy_ransac (reference velocity_estimator.py:344-507) -> (vy, r2, inlier share); the standard model for
    vx >= 0.1, the small-vx model (threshold 0.20, vy = -1/a) below."""
    if vx >= 0.1:
        fit = ransac_fit(np.sin(peaks[:, 0]), -1 * peaks[:, 1] - vx * np.cos(peaks[:, 0]), 0.15)
        return (0.0, 0.0, 0.0) if fit is None else (fit[0][0], fit[1], fit[2])
    fit = ransac_fit(peaks[:, 1] - vx, peaks[:, 0], 0.20)
    if fit is None:
        return 0.0, 0.0, 0.0
    a = fit[0][0]
    return (-1 / a if a != 0 else 0.0), fit[1], fit[2]


def adc_velocity_estimate(cube_v, geometry, virtual, range_bins, vel_bins, vel_res, vel_max, range_max_m, range_window,
                          valid_angle_range, current, precise=False, precise_vel_bound=0.25, thr_db=30.0,
                          min_r2=0.6, min_inlier=0.75, x_only=False):
    """VelocityEstimator.process with an ADC cube (reference velocity_estimator.py:783-865).  ``current`` is the running
    [vx, vy, vz] estimate (updated copy returned).  -> dict(current, proposed, vx, az/el (vy, r2, inlier), peaks)."""
    mask, vbins = da_valid_angles(64, valid_angle_range)
    az_sets, el_sets = adc_velocity_sets(geometry, virtual)
    ant_win = geometry == "standard" and virtual
    state = dict(az=(0.0, 0.0, 0.0), el=(0.0, 0.0, 0.0))

    def responses(use_precise, center):
        out, bins = [], vel_bins
        for sets, shift in ((az_sets, True), (el_sets, False)):
            acc = []
            for s in sets:
                if use_precise:
                    r, bins = doppler_azimuth_precise(cube_v, range_bins, vel_bins, vel_res, vel_max, 64, s, range_window,
                                                      shift, ant_win, range_max_m,
                                                      vel_range=[center - precise_vel_bound, center + precise_vel_bound])
                else:
                    r = doppler_azimuth_coarse(cube_v, range_bins, 64, s, range_window, shift, ant_win, range_max_m)
                acc.append(r[:, mask])
            out.append(sum(acc) / len(acc) if acc else None)
        return out, bins

    def vx_from(resps, bins):
        z = [da_peak_zero_az(r, bins, vbins, thr_db) for r in resps if r is not None]
        z = [p[1] for p in z if p.shape[0] > 0]
        return -1 * sum(z) / len(z) if z else 0.0

    resps, bins = responses(False, 0.0)
    vx = vx_from(resps, bins)
    if precise:
        resps, bins = responses(True, -1 * vx)
        vx = vx_from(resps, bins)
    peaks = [None, None]
    if not x_only:
        for i, key in enumerate(("az", "el")):
            if resps[i] is not None:
                peaks[i] = da_peaks_rows(resps[i], bins, vbins, thr_db)
                if peaks[i].shape[0] > 0:
                    state[key] = adc_fit_vy(peaks[i], vx)
    cur = np.array(current, dtype=np.float64)
    if x_only:
        proposed = np.array([vx])
        cur[0] = vx
    elif geometry == "ods":
        proposed = np.array([state["az"][0], state["el"][0], vx])
        for i, key in enumerate(("az", "el")):
            ok = state[key][1] >= min_r2 and state[key][2] >= min_inlier
            cur[i] = proposed[i] if ok else 0.0
        cur[2] = proposed[2]
    else:
        proposed = np.array([vx, state["az"][0], 0.0])
        if vx < 0.0:
            cur = proposed.copy() if state["az"][1] >= min_r2 else np.zeros(3)
        else:
            cur = np.array([proposed[0], proposed[1] if state["az"][1] >= min_r2 else 0.0, 0.0])
    return dict(current=cur, proposed=proposed, vx=vx, az=state["az"], el=state["el"], az_peaks=peaks[0],
                el_peaks=peaks[1])


# ---------------------------------------------------------------------------------------------------------------------
# SAR self-calibration (reference processors/simple_synthetic_array_beamformer_processor_multiFrame.py:673-816)
# ---------------------------------------------------------------------------------------------------------------------
def _peaks_height0(x):
    """scipy.signal.find_peaks(x, height=0): local maxima (plateau midpoints) whose value is >= 0."""
    pk = local_maxima_1d(x)
    return pk[np.asarray(x)[pk] >= 0] if len(pk) else pk


def sar_calibration_targets(freq_resps, image_el0, num_targets=3):
    """_find_calibration_targets (ref :698-729): the num_targets strongest peaks of the mean range profile in dB, then per
    range the strongest peak over azimuth of |10 log10 |image||.  Raises IndexError like the reference when the profile has
    fewer peaks than targets.  -> list of (range idx, azimuth idx)."""
    with np.errstate(divide="ignore"):
        avg = np.mean(20 * np.log10(np.abs(freq_resps)), axis=1)
    pk = _peaks_height0(avg)
    order = pk[np.argsort(avg[pk], kind="stable")[-num_targets:][::-1]]
    targets = []
    for i in range(num_targets):
        rng = order[i]
        with np.errstate(divide="ignore"):
            row = np.abs(10 * np.log10(np.abs(image_el0[rng])))
        p = _peaks_height0(row)
        if len(p) == 0:
            continue
        targets.append((int(rng), int(p[np.argmax(row[p])])))
    return targets


def sar_calibrate(aperture, geometry, image_el0, d_el0, lambda_m):
    """perform_array_calibration (ref :673-816).  aperture [S, N] (history reshaped), geometry (3, N), image_el0 [S, A]
    (beamformed response at elevation index 0), d_el0 (3, A) steering vectors.  -> (new geometry (3, N) or None when fewer
    than three targets qualify, targets)."""
    n = aperture.shape[1]
    freq = np.fft.fft(np.hamming(n).reshape(1, -1) * aperture, axis=0)
    targets = sar_calibration_targets(freq, image_el0)
    if len(targets) < 3:
        return None, targets
    diffs, vecs = [], []
    for rng, az in targets:
        sv = d_el0[:, az]
        shifted = freq[rng, :] * np.exp(1j * 2 * np.pi * (sv @ geometry) / lambda_m)
        diffs.append(np.diff(np.unwrap(np.angle(shifted))))
        vecs.append(sv)
    diffs, vecs = np.array(diffs), np.array(vecs)
    dj = 2 * np.pi / lambda_m * vecs[:, 0:2]
    corr = np.stack([np.linalg.lstsq(dj, diffs[:, i], rcond=None)[0] for i in range(diffs.shape[1])], axis=1)
    out = np.array(geometry, dtype=np.float64)
    out[0:2, 1:] -= np.cumsum(corr, axis=1)
    return out, targets


# ---------------------------------------------------------------------------------------------------------------------
# Polar -> cartesian resampling of the SAR image (reference ...multiFrame.py:587-613: scipy griddata(method="linear"))
# ---------------------------------------------------------------------------------------------------------------------
def polar_to_cart(values, range_bins, az_bins, xq, yq, diagonal=0):
    """Piecewise-linear interpolation of ``values`` [S, A], given at (r cos az, r sin az), at the points (xq, yq) over the
    triangulation that splits every polar cell (a cyclic trapezoid: BOTH of its diagonals give a Delaunay triangulation,
    so Qhull's pick inside scipy.griddata is arbitrary per cell) along diagonal 0 = (i, j)-(i+1, j+1) or 1 =
    (i, j+1)-(i+1, j); 0 outside the convex hull.  range_bins ascending from 0, az_bins ascending (column 0 of ``values``
    is the first sample of the flattened grid: the duplicate Qhull keeps at the origin).
    -> (interpolated values, inside mask)."""
    values = np.asarray(values)
    r = np.asarray(range_bins, dtype=np.float64)
    az = np.asarray(az_bins, dtype=np.float64)
    xq = np.asarray(xq, dtype=np.float64).ravel()
    yq = np.asarray(yq, dtype=np.float64).ravel()
    out = np.zeros(xq.shape, dtype=values.dtype)
    inside = np.zeros(xq.shape, dtype=bool)
    th = np.arctan2(yq, xq)
    j = np.clip(np.searchsorted(az, th, side="right") - 1, 0, len(az) - 2)
    half = 0.5 * (az[j + 1] - az[j])
    mid = 0.5 * (az[j + 1] + az[j])
    u = (xq * np.cos(mid) + yq * np.sin(mid)) / np.cos(half)          # chord index coordinate along the sector's bisector
    i = np.clip(np.searchsorted(r, u, side="right") - 1, 0, len(r) - 2)
    ok = (th >= az[0]) & (th <= az[-1]) & (u >= r[0]) & (u <= r[-1])

    def corner(ii, jj):
        # at r == 0 all azimuth samples coincide; Qhull keeps the first duplicate (flat index 0 = column 0)
        return r[ii] * np.cos(az[jj]), r[ii] * np.sin(az[jj]), np.where(r[ii] == 0.0, values[ii, 0], values[ii, jj])

    p00, p01, p10, p11 = corner(i, j), corner(i, j + 1), corner(i + 1, j), corner(i + 1, j + 1)
    tris = ((p00, p11, p10), (p00, p01, p11)) if diagonal == 0 else ((p00, p01, p10), (p01, p11, p10))
    for a, b, c in tris:
        det = (b[1] - c[1]) * (a[0] - c[0]) + (c[0] - b[0]) * (a[1] - c[1])
        with np.errstate(divide="ignore", invalid="ignore"):
            l0 = ((b[1] - c[1]) * (xq - c[0]) + (c[0] - b[0]) * (yq - c[1])) / det
            l1 = ((c[1] - a[1]) * (xq - c[0]) + (a[0] - c[0]) * (yq - c[1])) / det
            l2 = 1.0 - l0 - l1
            eps = -1e-12
            hit = ok & ~inside & (det != 0) & (l0 >= eps) & (l1 >= eps) & (l2 >= eps)
            out[hit] = (l0[hit] * a[2][hit] + l1[hit] * b[2][hit] + l2[hit] * c[2][hit])
        inside |= hit
    return out, inside
