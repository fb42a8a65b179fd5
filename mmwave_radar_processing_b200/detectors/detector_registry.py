"""String-keyed detector registry (reference: detectors/detector_registry.py:15-27)."""
from __future__ import annotations

from typing import Dict, Type, Union

from .base import BaseCFAR1D, BaseCFAR2D
from .ca_cfar import CaCFAR1D, CaCFAR2D
from .go_so_cfar import GoCFAR1D, SoCFAR1D
from .os_cfar import OsCFAR1D, OsCFAR2D


def get_detector_registry() -> Dict[str, Type[Union[BaseCFAR2D, BaseCFAR1D]]]:
    return {
        "ca_cfar_1d": CaCFAR1D,
        "ca_cfar_2d": CaCFAR2D,
        "os_cfar_1d": OsCFAR1D,
        "os_cfar_2d": OsCFAR2D,
        "go_cfar_1d": GoCFAR1D,
        "so_cfar_1d": SoCFAR1D,
    }
