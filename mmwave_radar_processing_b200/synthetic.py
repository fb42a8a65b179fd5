"""Seeded synthetic ADC frames (SURVEY.md section 8d recipe) in the device ingest layout.

The device layout is int16 ``[frame][chirp][rx][sample][2]`` (I, Q): chirp-major capture order,
each chirp holding all RX rows back to back (DESIGN.md "ADC layout").  For R=4 every chirp block
is ``16*S`` bytes, i.e. 16-byte aligned for any S.

Per frame: T point targets, each a complex exponential over (sample, loop, virtual antenna), plus
white Gaussian noise; rounded to nearest and clipped to int16.
"""
from __future__ import annotations

import numpy as np


def synth_frame_iq(num_rx: int, num_samples: int, cpl: int, loops: int, seed: int,
                   num_targets: int = 6, noise_sigma: float = 30.0) -> np.ndarray:
    """One frame, int16 ``[chirp][rx][sample][2]``; ``np.random.default_rng(seed)``."""
    rng = np.random.default_rng(seed)
    r, s, l = num_rx, num_samples, loops
    c = cpl * l
    amp = rng.uniform(100.0, 900.0, num_targets)
    rho = rng.uniform(3.0, s - 3.0, num_targets)                 # range bin
    lo, hi = -l / 2 + 2, l / 2 - 2
    if hi <= lo:                                                # very short Doppler axes (L=8)
        lo, hi = -l / 4, l / 4
    dlt = rng.uniform(lo, hi, num_targets)                       # Doppler bin
    sin_t = rng.uniform(-0.8, 0.8, num_targets)

    chirp = np.arange(c)
    tx = chirp % cpl
    loop = chirp // cpl
    rx = np.arange(r)
    n = np.arange(s)
    v = tx[:, None] * r + rx[None, :]                            # (C, R) virtual index
    x = np.zeros((c, r, s), dtype=np.complex128)
    for t in range(num_targets):
        ph = (rho[t] * n[None, None, :] / s
              + dlt[t] * loop[:, None, None] / l
              + 0.5 * sin_t[t] * v[:, :, None])
        x += amp[t] * np.exp(2j * np.pi * ph)
    x += noise_sigma * (rng.standard_normal(x.shape) + 1j * rng.standard_normal(x.shape))
    iq = np.empty((c, r, s, 2), dtype=np.int16)
    iq[..., 0] = np.clip(np.rint(x.real), -32768, 32767).astype(np.int16)
    iq[..., 1] = np.clip(np.rint(x.imag), -32768, 32767).astype(np.int16)
    return iq


def synth_scene_iq(num_rx: int, num_samples: int, cpl: int, loops: int, seed: int, targets,
                   noise_sigma: float = 30.0) -> np.ndarray:
    """One frame with EXPLICIT point targets ``[(amplitude, range_bin, doppler_bin, sin_theta), ...]`` (fractional bins
    allowed) plus seeded white noise; int16 ``[chirp][rx][sample][2]``.  Used for scripted sequences (a ground return whose
    range drifts from frame to frame drives the altimeter recurrence of the ground detector)."""
    rng = np.random.default_rng(seed)
    r, s, l = num_rx, num_samples, loops
    c = cpl * l
    chirp = np.arange(c)
    loop = chirp // cpl
    v = (chirp % cpl)[:, None] * r + np.arange(r)[None, :]      # (C, R) virtual index
    n = np.arange(s)
    x = np.zeros((c, r, s), dtype=np.complex128)
    for amp, rho, dlt, sin_t in targets:
        ph = rho * n[None, None, :] / s + dlt * loop[:, None, None] / l + 0.5 * sin_t * v[:, :, None]
        x += amp * np.exp(2j * np.pi * ph)
    if noise_sigma > 0:
        x += noise_sigma * (rng.standard_normal(x.shape) + 1j * rng.standard_normal(x.shape))
    iq = np.empty((c, r, s, 2), dtype=np.int16)
    iq[..., 0] = np.clip(np.rint(x.real), -32768, 32767).astype(np.int16)
    iq[..., 1] = np.clip(np.rint(x.imag), -32768, 32767).astype(np.int16)
    return iq


def synth_batch_iq(num_rx, num_samples, cpl, loops, seeds, **kw) -> np.ndarray:
    """Stack of frames ``[frame][chirp][rx][sample][2]`` for the given seeds."""
    return np.stack([synth_frame_iq(num_rx, num_samples, cpl, loops, int(sd), **kw) for sd in seeds])


def synth_batch_for_cfg(cfg, seeds, **kw) -> np.ndarray:
    """Frames shaped for a loaded ``ConfigManager``."""
    return synth_batch_iq(cfg.num_rx_antennas, cfg.get_num_adc_samples(0), cfg.chirp_cfgs_per_loop,
                          cfg.frameCfg_loops, seeds, **kw)


def iq_to_complex_cube(iq: np.ndarray) -> np.ndarray:
    """int16 ``[chirp][rx][sample][2]`` -> complex128 ``(rx, samples, chirps)``: the cube the
    reference's dataset loader hands to processors (exact cast)."""
    z = iq[..., 0].astype(np.float64) + 1j * iq[..., 1].astype(np.float64)
    return np.ascontiguousarray(np.moveaxis(z, -3, -1))         # (..., R, S, C)


# (row, column) of each virtual antenna, in half wavelengths.  "ods": the 4x4-minus-corners ODS array, consistent with
# the antenna sets the reference's ADC-domain velocity estimator uses (azimuth rows [0,3,4,7] / [1,2,5,6], elevation
# columns [10,11,6,7] / [9,8,5,4]); "standard": a uniform line.
ODS_ROW_COL = {0: (3, 0), 3: (3, 1), 4: (3, 2), 7: (3, 3), 1: (2, 0), 2: (2, 1), 5: (2, 2), 6: (2, 3),
               10: (0, 3), 11: (1, 3), 9: (0, 2), 8: (1, 2)}


def synth_ego_motion_iq(num_rx: int, num_samples: int, cpl: int, loops: int, seed: int, ego, vel_res: float,
                        geometry: str = "standard", range_bins=(8.0, 20.0), num_scatterers: int = 60,
                        mover_frac: float = 0.15, noise_sigma: float = 20.0) -> np.ndarray:
    """One frame of a static scene seen from a platform moving at ``ego`` = (boresight, azimuth-lateral,
    elevation-lateral) m/s: every scatterer at (theta, phi) shows the Doppler -(e0 cos(theta) cos(phi) + e1 sin(theta) +
    e2 sin(phi)); a fraction are movers with unrelated Doppler.  int16 ``[chirp][rx][sample][2]``."""
    rng = np.random.default_rng(seed)
    r, s, l = num_rx, num_samples, loops
    c = cpl * l
    k = num_scatterers
    theta = np.deg2rad(rng.uniform(-55.0, 55.0, k))
    phi = np.deg2rad(rng.uniform(-55.0, 55.0, k)) if geometry == "ods" else np.zeros(k)
    amp = rng.uniform(200.0, 600.0, k)
    rho = rng.uniform(range_bins[0], range_bins[1], k)
    dop = -(ego[0] * np.cos(theta) * np.cos(phi) + ego[1] * np.sin(theta) + ego[2] * np.sin(phi))
    movers = rng.random(k) < mover_frac
    dop[movers] = rng.uniform(-0.4, 0.4, int(movers.sum())) * vel_res * l
    chirp = np.arange(c)
    v = (chirp % cpl)[:, None] * r + np.arange(r)[None, :]       # (C, R) virtual index
    if geometry == "ods":
        row = np.vectorize(lambda i: ODS_ROW_COL.get(int(i), (0, 0))[0])(v)
        col = np.vectorize(lambda i: ODS_ROW_COL.get(int(i), (0, 0))[1])(v)
    else:
        row, col = np.zeros_like(v), v
    loop = (chirp // cpl)[:, None, None]
    n = np.arange(s)[None, None, :]
    x = np.zeros((c, r, s), dtype=np.complex128)
    for t in range(k):
        ph = rho[t] * n / s + dop[t] / (vel_res * l) * loop + \
            0.5 * (np.sin(theta[t]) * col + np.sin(phi[t]) * row)[:, :, None]
        x += amp[t] * np.exp(2j * np.pi * ph)
    x += noise_sigma * (rng.standard_normal(x.shape) + 1j * rng.standard_normal(x.shape))
    iq = np.empty((c, r, s, 2), dtype=np.int16)
    iq[..., 0] = np.clip(np.rint(x.real), -32768, 32767).astype(np.int16)
    iq[..., 1] = np.clip(np.rint(x.imag), -32768, 32767).astype(np.int16)
    return iq
