"""Cell-averaging CFAR (reference: detectors/ca_cfar.py:6-155)."""
from .base import BaseCFAR1D, BaseCFAR2D


class CaCFAR1D(BaseCFAR1D):
    _kind = "ca"

    def _alpha(self, n_all, one_side):
        return self.compute_alpha_ca(n_all, self.pfa)          # N = 2*num_train


class CaCFAR2D(BaseCFAR2D):
    _kind = "ca"

    def _alpha(self, n_all, one_side):
        return self.compute_alpha_ca(n_all, self.pfa)          # N = window cells - guard block
