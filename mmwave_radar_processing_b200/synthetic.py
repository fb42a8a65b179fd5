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
