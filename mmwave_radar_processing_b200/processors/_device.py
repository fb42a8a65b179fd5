"""Host <-> device staging shared by the processor classes.

The reference hands processors a host ``np.ndarray`` complex cube ``(antennas, samples, chirps)``; the
compatibility path uploads it as complex64 in the ``RADAR_IN_CUBE64`` layout and downloads results as
fp64 / complex128.  CUDA tensors (int16 IQ frames ``[F][chirp][rx][sample][2]`` or complex64 cubes
``[F][V][S][L]``) are accepted as an extension and keep results on the device.
"""
from __future__ import annotations

import numpy as np


def is_device_tensor(x) -> bool:
    try:
        import torch
    except ImportError:  # pragma: no cover
        return False
    return isinstance(x, torch.Tensor)


def engine_for_cube_shape(v: int, s: int, l: int):
    """Engine for a host cube: treated as V 'receivers', one TX slot."""
    from ..engine import RadarEngine, _ENGINES
    import torch
    key = (int(v), int(s), 1, int(l), f"cuda:{torch.cuda.current_device()}")
    eng = _ENGINES.get(key)
    if eng is None:
        eng = RadarEngine(int(v), int(s), 1, int(l))
        _ENGINES[key] = eng
    return eng


def stage_cube(adc_cube, config_manager=None):
    """-> (engine, device tensor with a leading frame axis, squeeze flag).

    numpy complex (V,S,L) -> complex64 [1,V,S,L] on the device (values of an ADC cube are integers, the
    cast is exact).  torch tensors must already be in one of the two device layouts."""
    import torch
    if is_device_tensor(adc_cube):
        from ..engine import RadarEngine
        if config_manager is None:
            raise ValueError("device tensors need a config manager to fix the radar shape")
        t = adc_cube
        if t.dtype == torch.int16:
            squeeze = t.dim() == 4
            t = t[None] if squeeze else t
            eng = RadarEngine.for_config(config_manager)
        elif t.dtype == torch.complex64:
            squeeze = t.dim() == 3
            t = t[None] if squeeze else t
            eng = engine_for_cube_shape(t.shape[1], t.shape[2], t.shape[3])
        else:
            raise ValueError(f"unsupported device tensor dtype {t.dtype}")
        return eng, t.contiguous(), squeeze
    cube = np.asarray(adc_cube)
    if cube.ndim != 3:
        raise ValueError("adc_cube must be (antennas, samples, chirps)")
    eng = engine_for_cube_shape(*cube.shape)
    host = np.ascontiguousarray(cube, dtype=np.complex64)[None]
    return eng, torch.from_numpy(host).to(eng.device), True


def to_host_f64(t):
    return t.detach().cpu().numpy().astype(np.float64)


def to_host_c128(t):
    return t.detach().cpu().numpy().astype(np.complex128)
