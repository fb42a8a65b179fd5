"""Greatest-of / smallest-of CFAR, 1-D only (reference: detectors/go_so_cfar.py:6-123).
alpha is computed from ONE side's cell count (N = num_train), as the reference does (:58, :112)."""
from .base import BaseCFAR1D


class GoCFAR1D(BaseCFAR1D):
    _kind = "go"

    def _alpha(self, n_all, one_side):
        return self.compute_alpha_ca(self.num_train, self.pfa)


class SoCFAR1D(BaseCFAR1D):
    _kind = "so"

    def _alpha(self, n_all, one_side):
        return self.compute_alpha_ca(self.num_train, self.pfa)
