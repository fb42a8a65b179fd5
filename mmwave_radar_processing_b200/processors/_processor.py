"""Processor base class (reference: processors/_processor.py:6-64): stores the config manager, the
estimate / ground-truth histories and a logger; the constructor calls ``configure()`` so subclasses set
their attributes BEFORE ``super().__init__``."""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ..logging.logger import get_logger


class _Processor:

    def __init__(self, config_manager: ConfigManager, **kwargs) -> None:
        self.config_manager = config_manager
        self.history_estimated = []
        self.history_gt = []
        self.logger = get_logger(__name__)
        self.configure()

    def configure(self):
        pass

    def reset(self):
        self.history_estimated = []
        self.history_gt = []

    def update_history(self, estimated: np.ndarray = np.empty(0), ground_truth: np.ndarray = np.empty(0)) -> None:
        if estimated.size > 0:
            self.history_estimated.append(estimated.copy())
        if ground_truth.size > 0:
            self.history_gt.append(ground_truth.copy())

    def process(self, adc_cube: np.ndarray) -> np.ndarray:
        pass
