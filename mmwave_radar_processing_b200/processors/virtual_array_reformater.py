"""TDM-MIMO virtual-array regrouping (reference: processors/virtual_array_reformater.py:6-65).

Pure re-indexing: virtual channel ``v = tx_slot*num_rx + rx`` takes chirps ``tx_slot, tx_slot+cpl, ...``.
For host cubes this is a numpy gather (no arithmetic).  On the device path it does not exist as a step at
all: the range-Doppler kernel reads int16 frames in capture order and folds this mapping into its load
addresses, so CUDA int16 frames pass through unchanged.
"""
import numpy as np

from ..config_managers.cfgManager import ConfigManager
from ._device import is_device_tensor
from ._processor import _Processor


class VirtualArrayReformatter(_Processor):

    def __init__(self, config_manager: ConfigManager, **kwargs) -> None:
        self.chirp_cfg_idxs: np.ndarray = None
        self.chirp_cfg_idxs_for_frame: np.ndarray = None
        self.chirp_cfgs_per_loop: int = 0
        self.adc_samples_per_chirp: int = 0
        super().__init__(config_manager)

    def configure(self):
        cm = self.config_manager
        self.chirp_cfg_idxs = np.arange(start=cm.frameCfg_start_index, stop=cm.frameCfg_end_index + 1)
        self.chirp_cfg_idxs_for_frame = np.tile(A=self.chirp_cfg_idxs, reps=cm.frameCfg_loops)
        self.chirp_cfgs_per_loop = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
        self.adc_samples_per_chirp = cm.get_num_adc_samples(profile_idx=0)

    def process(self, adc_cube, **kwargs):
        if is_device_tensor(adc_cube):
            return adc_cube
        cm = self.config_manager
        r, cpl, loops = cm.num_rx_antennas, self.chirp_cfgs_per_loop, cm.frameCfg_loops
        out = np.zeros((r * cpl, self.adc_samples_per_chirp, loops), dtype=complex)
        for slot in range(cpl):
            out[slot * r:(slot + 1) * r, :, :] = adc_cube[0:r, :, slot::cpl][:, :, :loops]
        return out
