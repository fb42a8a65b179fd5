"""Registry of Range-Doppler detectors (reference: processors/range_doppler_detection/registry.py:10-23)."""
from typing import Dict, Type

from .range_doppler_detector import RangeDopplerDetector
from .range_doppler_detector_2d import RangeDopplerDetector2D
from .range_doppler_detector_sequential import RangeDopplerDetectorSequential
from .range_doppler_ground_detector import RangeDopplerGroundDetector

_RANGE_DOPPLER_DETECTOR_REGISTRY: Dict[str, Type[RangeDopplerDetector]] = {
    "range_doppler_detector_2d": RangeDopplerDetector2D,
    "range_doppler_detector_sequential": RangeDopplerDetectorSequential,
    "range_doppler_ground_detector": RangeDopplerGroundDetector,
}


def get_range_doppler_detector_registry() -> Dict[str, Type[RangeDopplerDetector]]:
    return _RANGE_DOPPLER_DETECTOR_REGISTRY
