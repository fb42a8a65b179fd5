"""Ordered-statistic CFAR (reference: detectors/os_cfar.py:6-195).

``k_rank = clip(int(rho*N), 1, N)`` with N = 2*num_train (1-D) or window cells minus the guard block
(2-D); the noise estimate is the k-th smallest training cell (exact selection on the device) and the
user supplies ``alpha`` directly."""
from typing import Tuple

from .base import BaseCFAR1D, BaseCFAR2D


class OsCFAR1D(BaseCFAR1D):
    _kind = "os"

    def __init__(self, num_train: int, num_guard: int, rho: float, alpha: float, **kwargs):
        super().__init__(num_train, num_guard, pfa=0.0, **kwargs)
        self.alpha = alpha
        n_cells = 2 * num_train
        self.k_rank = max(1, min(int(rho * n_cells), n_cells))

    def _alpha(self, n_all, one_side):
        return float(self.alpha)

    def _os_rank(self):
        return int(self.k_rank)


class OsCFAR2D(BaseCFAR2D):
    _kind = "os"

    def __init__(self, num_train: Tuple[int, int], num_guard: Tuple[int, int], rho: float, alpha: float, **kwargs):
        super().__init__(num_train, num_guard, pfa=0.0, **kwargs)
        self.alpha = alpha
        win_r = 2 * (num_train[0] + num_guard[0]) + 1
        win_d = 2 * (num_train[1] + num_guard[1]) + 1
        n_cells = win_r * win_d - (2 * num_guard[0] + 1) * (2 * num_guard[1] + 1)
        self.k_rank = max(1, min(int(rho * n_cells), n_cells))

    def _alpha(self, n_all, one_side):
        return float(self.alpha)

    def _os_rank(self):
        return int(self.k_rank)
