"""CFAR detector base classes with the reference's interface, evaluated by the CUDA CFAR kernel.

Host-side mirror of ``BaseCFAR1D`` / ``BaseCFAR2D`` (reference: detectors/base.py:7-327): same constructor
arguments, ``detect`` return types (``list[int]`` / ``list[(row, col)]`` in ``np.where`` order) and cached
``thresholds`` / ``detections`` / ``noise_estimates`` arrays.  The arithmetic runs in
``libradar_b200.so`` (``radar_cfar_2d`` / ``radar_cfar_1d_rows``): the input is cast to fp32 on upload,
training statistics and the decision ``x > alpha*noise`` are evaluated in fp64 on the device, and
``thresholds`` is formed here with that same fp64 product (``+inf`` outside the valid interior, like the
reference).
"""
from __future__ import annotations

from abc import ABC
from typing import List, Optional, Tuple

import numpy as np


def _engine():
    from ..engine import utility_engine
    return utility_engine()


class _DeviceCfar(ABC):
    """Shared upload / launch / download logic."""
    _kind: str = ""

    thresholds: Optional[np.ndarray]
    detections: Optional[np.ndarray]
    noise_estimates: Optional[np.ndarray]

    def _alpha(self, n_all: int, one_side: int) -> float:
        raise NotImplementedError

    def _os_rank(self) -> int:
        return 0

    def _window(self):
        """-> (tr, td, gr, gd)"""
        raise NotImplementedError

    def _c_params(self):
        from ..engine import cfar_params
        tr, td, gr, gd = self._window()
        n_all = (2 * (tr + gr) + 1) * (2 * (td + gd) + 1) - (2 * gr + 1) * (2 * gd + 1)
        alpha = self._alpha(n_all, td)
        p = cfar_params(self._kind, (tr, td), (gr, gd), pfa=float(getattr(self, "pfa", 0.0) or 0.0),
                        alpha=alpha, os_rank=self._os_rank())
        return p, alpha

    def _run(self, x2d: np.ndarray, one_d: bool):
        """x2d: float array [rows, cols]; fills the cached arrays and returns int64 (N,2) detections."""
        import torch
        from .._native import RadarError
        eng = _engine()
        rows, cols = x2d.shape
        params, alpha = self._c_params()
        maps = torch.from_numpy(np.ascontiguousarray(x2d, dtype=np.float32)[None]).to(eng.device)
        try:
            res = eng.cfar(maps, params, one_d=one_d, want_noise=True, want_mask=True)
        except RadarError as exc:
            if exc.code == -1:
                raise ValueError(str(exc)) from None
            raise
        n = int(res["counts"][0].item())
        dets = res["dets"][0, :n].cpu().numpy().astype(np.int64)
        noise = res["noise"][0].cpu().numpy()
        mask = res["mask"][0].cpu().numpy().astype(bool)
        tr, td, gr, gd = self._window()
        cr, cd = tr + gr, td + gd
        thr = np.full((rows, cols), np.inf)
        if rows - 2 * cr > 0 and cols - 2 * cd > 0:
            thr[cr:rows - cr, cd:cols - cd] = alpha * noise[cr:rows - cr, cd:cols - cd]
        return dets, thr, noise, mask


class BaseCFAR1D(_DeviceCfar):
    """1-D CFAR: ``num_train`` / ``num_guard`` cells on each side of the cell under test."""

    def __init__(self, num_train: int, num_guard: int, pfa: float, **kwargs):
        self.num_train = num_train
        self.num_guard = num_guard
        self.pfa = pfa
        self.thresholds = None
        self.detections = None
        self.noise_estimates = None

    def _window(self):
        return 0, int(self.num_train), 0, int(self.num_guard)

    def detect(self, x: np.ndarray) -> List[int]:
        x = np.asarray(x)
        if x.ndim != 1:
            raise ValueError("Input x must be a 1D array.")
        dets, thr, noise, mask = self._run(x[None, :], one_d=True)
        self.thresholds, self.noise_estimates, self.detections = thr[0], noise[0], mask[0]
        return dets[:, 1].tolist()

    @staticmethod
    def compute_alpha_ca(num_train_cells: int, pfa: float) -> float:
        return num_train_cells * (pfa ** (-1.0 / num_train_cells) - 1.0)


class BaseCFAR2D(_DeviceCfar):
    """2-D CFAR: ``num_train`` / ``num_guard`` are (range, doppler) half-widths."""

    def __init__(self, num_train: Tuple[int, int], num_guard: Tuple[int, int], pfa: float, **kwargs):
        self.num_train = num_train
        self.num_guard = num_guard
        self.pfa = pfa
        self.thresholds = None
        self.detections = None
        self.noise_estimates = None

    def _window(self):
        return int(self.num_train[0]), int(self.num_train[1]), int(self.num_guard[0]), int(self.num_guard[1])

    def detect(self, X: np.ndarray) -> List[Tuple[int, int]]:
        X = np.asarray(X)
        if X.ndim != 2:
            raise ValueError("Input X must be a 2D array.")
        dets, thr, noise, mask = self._run(X, one_d=False)
        self.thresholds, self.noise_estimates, self.detections = thr, noise, mask
        return list(zip(dets[:, 0], dets[:, 1]))

    @staticmethod
    def compute_alpha_ca(num_train_cells: int, pfa: float) -> float:
        return num_train_cells * (pfa ** (-1.0 / num_train_cells) - 1.0)
