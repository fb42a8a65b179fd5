"""ctypes binding of ``libradar_b200.so`` (the C ABI in ``include/radar_b200.h``).

There is no CPU fallback: if the shared library has not been built (``python __graft_entry__.py``) the
import of any compute entry point raises ``RuntimeError``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libradar_b200.so")

RADAR_IN_IQ16 = 0
RADAR_IN_CUBE64 = 1
RADAR_WIN_RANGE = 1
RADAR_WIN_DOPPLER = 2
RADAR_WIN_ANTENNA = 4
RADAR_RD_RANGE_ONLY = 8
RADAR_REDUCE_MAX, RADAR_REDUCE_MEAN = 0, 1
RADAR_CFAR_CA, RADAR_CFAR_OS, RADAR_CFAR_GO, RADAR_CFAR_SO = 0, 1, 2, 3
CFAR_KIND = {"ca": RADAR_CFAR_CA, "os": RADAR_CFAR_OS, "go": RADAR_CFAR_GO, "so": RADAR_CFAR_SO}
ABI_VERSION = 1


class RadarShape(C.Structure):
    _fields_ = [("num_rx", C.c_int32), ("num_samples", C.c_int32),
                ("chirps_per_loop", C.c_int32), ("num_loops", C.c_int32)]


class CfarParams(C.Structure):
    _fields_ = [("kind", C.c_int32), ("train_r", C.c_int32), ("train_d", C.c_int32),
                ("guard_r", C.c_int32), ("guard_d", C.c_int32), ("os_rank", C.c_int32),
                ("pfa", C.c_double), ("rho", C.c_double), ("alpha", C.c_double)]


_P = C.c_void_p
_I = C.c_int
# name -> (restype, argtypes); must list every symbol include/radar_b200.h declares
SIGNATURES = {
    "radar_abi_version": (_I, []),
    "radar_ctx_create": (_I, [C.POINTER(RadarShape), _I, C.POINTER(_P)]),
    "radar_ctx_destroy": (_I, [_P]),
    "radar_last_error": (C.c_char_p, [_P]),
    "radar_last_rd_kernel": (C.c_char_p, [_P]),
    "radar_ctx_set_axes": (_I, [_P, C.POINTER(C.c_double), _I, C.POINTER(C.c_double), _I]),
    "radar_ctx_set_angle_bins": (_I, [_P, C.POINTER(C.c_double), _I]),
    "radar_range_doppler": (_I, [_P, _P, _I, _I, _P, _I, _I, _P, _P, _I, _P]),
    "radar_range_profile": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "radar_range_zoom": (_I, [_P, _P, _I, _I, _I, _P, _P, C.c_double, _P, _P]),
    "radar_find_peaks": (_I, [_P, _P, _I, _I, _I, C.c_double, C.c_double, _I, _P, _P, _P, _P]),
    "radar_cfar_2d": (_I, [_P, _P, _I, _I, _I, C.POINTER(CfarParams), _P, _P, _P, _P, _I, _P]),
    "radar_cfar_1d_rows": (_I, [_P, _P, _I, _I, _I, C.POINTER(CfarParams), _P, _P, _P, _P, _P, _I, _P]),
    "radar_cfar_ca_alpha": (C.c_double, [_I, C.c_double]),
    "radar_point_cloud": (_I, [_P, _P, _I, _P, _P, _I, _P, _I, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P]),
    "radar_det_offsets": (_I, [_P, _P, _I, _I, _P, _P]),
    "radar_range_angle": (_I, [_P, _P, _I, _I, _I, _P, _I, _I, _I, _P, _P]),
    "radar_dbs": (_I, [_P, _P, _I, _I, _P, _I, _I, _P, _P, _I, _P, _P]),
    "radar_sar_image": (_I, [_P, _P, _I, _I, _P, _P, _I, C.c_double, _I, _I, _P, _P]),
    "radar_sar_calibrate": (_I, [_P, _P, _I, _I, _P, _P, _P, _I, _I, _P, C.c_double, _P, _P, _P]),
    "radar_polar_to_cart": (_I, [_P, _P, _I, _I, _I, _P, _P, _P, _I, C.c_double, C.c_double, _I, C.c_double, C.c_double, _I, _I,
                                 _P, _P, _P]),
    "radar_ctx_set_point_mirrors": (_I, [_P, _P, _I, _I]),
    "radar_ctx_set_image_mirrors": (_I, [_P, _P, _I, C.c_int64, C.c_int64]),
    "radar_doppler_azimuth": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P, _I, _I, _P, _P]),
    "radar_chirp_zoom": (_I, [_P, _P, _I, _I, _I, _I, _P, _I, _P, _P]),
    "radar_chirp_zoom_static": (_I, [_P, _P, _I, _I, _I, _I, _P, _I, _P, _P]),
    "radar_chirp_zoom_static_capacity": (_I, []),
    "radar_row_reduce": (_I, [_P, _P, _I, _I, _I, _P, _I, _P, _P]),
    "radar_ego_velocity": (_I, [_P, _P, _P, _I, _I, _P, _P, _I, _I, _I, C.c_double, C.c_double, _P, _P, _P, _P, _P]),
}

_lib = None


def load_library():
    """dlopen the in-tree shared library and bind every declared symbol (no GPU needed for this)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build the CUDA library first (python __graft_entry__.py). "
            "mmwave_radar_processing_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.radar_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libradar_b200 ABI {lib.radar_abi_version()} != binding {ABI_VERSION}")
    _lib = lib
    return lib


class RadarError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libradar_b200 error {code}: {message}")
        self.code = code


def check(rc, ctx=None):
    if rc != 0:
        msg = load_library().radar_last_error(ctx)
        raise RadarError(rc, msg.decode() if msg else "unknown")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())
