"""TI mmWave SDK ``.cfg`` parsing for the B200 radar front-end.

Host-side mirror of the reference ``ConfigManager``
(reference: mmwave_radar_processing/config_managers/cfgManager.py:16-363).
The attribute names, the derived-quantity arithmetic and the parsing quirks are
kept so that every processor written against the reference sees the same
numbers:

* only ``channelCfg, adcCfg, adcbufCfg, profileCfg, chirpCfg, frameCfg`` are read
  (ref :246-264); lines containing ``%`` anywhere are skipped (ref :234);
* ``adcbufCfg`` fields are taken from the END of the line so both the 4- and
  5-field forms parse (ref :294-302);
* a second profile / chirp whose id is already taken prints a warning and is
  dropped (ref :323-328, :345-350);
* range / velocity resolution use the same floating-point expression order
  (ref :113-158) so ``np.arange`` bin axes built from them have identical
  lengths and values.

The implementation is table driven (one field spec per command) instead of the
reference's per-command methods; this stays pure Python - it is configuration,
not the hot path.
"""
from __future__ import annotations

import numpy as np

# scipy.constants.c (ref cfgManager.py:5,119) - exact SI value, no scipy import needed.
SPEED_OF_LIGHT_M_S = 299792458.0


class InvalidConfiguration(Exception):
    pass


class ConfigNotLoaded(Exception):
    pass


_PROFILE_FIELDS = (
    ("profileId", int), ("startFreq_GHz", float), ("idleTime_us", float),
    ("adcStartTime_us", float), ("rampEndTime_us", float), ("txOutPower", float),
    ("txPhaseShifter", float), ("freqSlope_MHz_us", float), ("txStartTime_us", float),
    ("adcSamples", int), ("sampleRate_kSps", int), ("hpfCornerFreq1", int),
    ("hpfCornerFreq2", int), ("rxGain_dB", float),
)

_CHIRP_FIELDS = (
    ("startIndex", int), ("endIndex", int), ("profile", int),
    ("startFreqVariation_Hz", float), ("freqSlopVariation_MHz_us", float),
    ("idleTimeVariation_us", float), ("ADCStartTimeVariation_us", float), ("txMask", int),
)

_ADC_BITS = {0: 12, 1: 14, 2: 16}


def _popcount(v: int) -> int:
    return bin(v).count("1")


class ConfigManager:
    """Radar configuration: parsed TI commands + derived performance numbers."""

    def __init__(self):
        # channelCfg
        self.channelCfg_tx_chan_enabled: int = 0
        self.channelCfg_rx_chan_enabled: int = 0
        self.channelCfg_cascading: bool = 0
        # adcCfg
        self.adcCfg_num_adc_bits: int = 0
        self.adcCfg_adcOutputFmt: int = 0
        # adcbufCfg
        self.adcbufCfg_adc_output_fmt: int = 0
        self.adcbufCfg_sample_swap: bool = False
        self.adcbufCfg_channel_interleave: bool = False
        self.adcbufCfg_chirp_threshold: int = 1
        # profileCfg / chirpCfg: one placeholder entry until the first real one arrives
        placeholder_profile = {name: kind(0) for name, kind in _PROFILE_FIELDS}
        placeholder_profile["profileId"] = -1
        placeholder_profile["startFreq_GHz"] = 77.0
        self.profile_cfgs: list = [placeholder_profile]
        placeholder_chirp = {name: kind(0) for name, kind in _CHIRP_FIELDS}
        placeholder_chirp["startIndex"] = -1
        placeholder_chirp["endIndex"] = -1
        self.chirp_cfgs: list = [placeholder_chirp]
        # frameCfg
        self.frameCfg_start_index: int = 0
        self.frameCfg_end_index: int = 0
        self.frameCfg_loops: int = 0
        self.frameCfg_frames: int = 0
        self.frameCfg_periodicity_ms: float = 0
        self.frameCfg_hardware_trigger_enabled: bool = False
        self.frameCfg_trigger_delay_ms: float = 0.0
        # derived
        self.range_res_m: float = 0.0
        self.range_bin_size_m: float = 0.0
        self.range_max_m: float = 0.0
        self.range_bins_m: np.ndarray = np.empty(shape=0, dtype=float)
        self.vel_res_m_s: float = 0.0
        self.vel_max_m_s: float = 0.0
        self.num_tx_antennas: int = 3
        self.num_rx_antennas: int = 4
        self.virtual_antennas_enabled: bool = False
        self.config_loaded = False
        self.array_geometry = "standard"   # "standard" or "ods"
        self.array_direction = "down"      # "down" or "out"

    # ------------------------------------------------------------------ derived numbers
    def compute_radar_perforance(self, profile_idx: int = 0):
        # (sic) the reference spells it this way and callers use the name
        self._compute_range_performance(profile_idx)
        self._compute_vel_performance(profile_idx)
        self._compute_angular_performance()

    def _compute_range_performance(self, profile_idx: int = 0):
        # ref :113-132. Keep the expression order: c*fs*1e3 / (2*slope*(1e6/1e-6)*S)
        n_samp = self.get_num_adc_samples(profile_idx)
        fs_ksps = self.get_adc_sample_rate_kSps(profile_idx)
        slope = self.get_chirp_slope_MHz_us(profile_idx)
        pow2_bins = np.power(2, np.ceil(np.log2(n_samp)))
        numer = SPEED_OF_LIGHT_M_S * fs_ksps * 1e3
        self.range_res_m = numer / (2 * slope * (1e6 / 1e-6) * n_samp)
        self.range_bin_size_m = numer / (2 * slope * (1e6 / 1e-6) * pow2_bins)
        self.range_max_m = numer / (2 * slope * (1e6 / 1e-6))

    def _compute_vel_performance(self, profile_idx: int = 0):
        # ref :134-158
        prof = self.profile_cfgs[profile_idx]
        lambda_m = SPEED_OF_LIGHT_M_S / (float(prof["startFreq_GHz"]) * 1e9)
        chirps_per_loop = self.frameCfg_end_index - self.frameCfg_start_index + 1
        loops = float(self.frameCfg_loops)
        chirp_period_us = float(prof["rampEndTime_us"]) + float(prof["idleTime_us"])
        self.vel_res_m_s = lambda_m / (2 * chirp_period_us * chirps_per_loop * 1e-6 * loops)
        self.vel_max_m_s = lambda_m / (4 * chirp_period_us * chirps_per_loop * 1e-6)

    def _compute_angular_performance(self):
        # ref :160-169 - TDM-MIMO is "on" whenever more than one chirp cfg is looped
        self.virtual_antennas_enabled = (self.frameCfg_end_index - self.frameCfg_start_index + 1) > 1

    # ------------------------------------------------------------------ helpers (ref :175-185)
    def get_adc_sample_rate_kSps(self, profile_idx: int = 0) -> int:
        return self.profile_cfgs[profile_idx]["sampleRate_kSps"]

    def get_num_adc_samples(self, profile_idx: int = 0) -> int:
        return self.profile_cfgs[profile_idx]["adcSamples"]

    def get_chirp_slope_MHz_us(self, profile_idx: int = 0) -> float:
        return self.profile_cfgs[profile_idx]["freqSlope_MHz_us"]

    # shape helpers used by the B200 engine (not in the reference)
    @property
    def chirp_cfgs_per_loop(self) -> int:
        return self.frameCfg_end_index - self.frameCfg_start_index + 1

    @property
    def num_virtual_antennas(self) -> int:
        return self.num_rx_antennas * self.chirp_cfgs_per_loop

    @property
    def chirps_per_frame(self) -> int:
        return self.chirp_cfgs_per_loop * self.frameCfg_loops

    def print_cfg_overview(self):
        prof = self.profile_cfgs[0]
        chirp_period_us = prof["idleTime_us"] + prof["rampEndTime_us"]
        n_slots = self.chirp_cfgs_per_loop
        print("---- Radar Configuration Overview ----")
        print("Radar Performance")
        print(f"\trange res: {self.range_res_m:.2f}m")
        print(f"\trange max: {self.range_max_m:.2f}m")
        print(f"\tvel res: {self.vel_res_m_s:.2f}m/s")
        print(f"\tvel max: {self.vel_max_m_s:.2f}m/s")
        print("Profile Configuration:")
        print(f"\tprofile config: \n\t{prof}")
        print(f"\tFrame loops: {self.frameCfg_loops}")
        print(f"\tChirp period: {chirp_period_us} us")
        print(f"\tActive frame time: {n_slots * self.frameCfg_loops * chirp_period_us * 1e-3} ms")
        print(f"\tFrame period: {self.frameCfg_periodicity_ms}")
        print(f"\tAntenna Geometry: {self.array_geometry}")
        print(f"\tOperating Frequency: {prof['startFreq_GHz']} GHz")

    # ------------------------------------------------------------------ loading
    def load_cfg(self, cfg_file_path: str, array_geometry: str = "standard", array_direction: str = "down"):
        """Parse a TI ``.cfg`` file (ref :210-239)."""
        self.array_geometry = array_geometry
        self.array_direction = array_direction
        with open(cfg_file_path) as fh:
            for line in fh:
                if "%" in line:
                    continue
                self._load_cfg_command_from_line(line)
        self.config_loaded = True
        self.compute_radar_perforance(profile_idx=0)

    def _load_cfg_command_from_line(self, line: str):
        tokens = line.strip("\n").split(" ")
        handler = self._HANDLERS.get(tokens[0])
        if handler is not None:
            handler(self, tokens)

    def _on_channel(self, tok):
        self.channelCfg_rx_chan_enabled = int(tok[1])
        self.num_rx_antennas = _popcount(self.channelCfg_rx_chan_enabled)
        self.channelCfg_tx_chan_enabled = int(tok[2])
        self.num_tx_antennas = _popcount(self.channelCfg_tx_chan_enabled)
        self.channelCfg_cascading = int(tok[3])

    def _on_adc(self, tok):
        code = int(tok[1])
        if code in _ADC_BITS:
            self.adcCfg_num_adc_bits = _ADC_BITS[code]
        self.adcCfg_adcOutputFmt = int(tok[2])

    def _on_adcbuf(self, tok):
        self.adcbufCfg_adc_output_fmt = int(tok[-4])          # 0 complex, 1 real
        self.adcbufCfg_sample_swap = int(tok[-3]) != 0
        self.adcbufCfg_channel_interleave = int(tok[-2]) == 0
        self.adcbufCfg_chirp_threshold = int(tok[-1])

    @staticmethod
    def _record(fields, tok):
        return {name: kind(tok[i + 1]) for i, (name, kind) in enumerate(fields)}

    def _on_profile(self, tok):
        rec = self._record(_PROFILE_FIELDS, tok)
        if self.profile_cfgs[0]["profileId"] == -1:
            self.profile_cfgs[0] = rec
        elif rec["profileId"] < len(self.profile_cfgs):
            print("cfgManager: attempted to load multiple profiles with the same ID")
        else:
            self.profile_cfgs.append(rec)

    def _on_chirp(self, tok):
        rec = self._record(_CHIRP_FIELDS, tok)
        if self.chirp_cfgs[0]["startIndex"] == -1:
            self.chirp_cfgs[0] = rec
        elif rec["startIndex"] < len(self.chirp_cfgs):
            print("cfgManager: attempted to load multiple chirps with the same ID")
        else:
            self.chirp_cfgs.append(rec)

    def _on_frame(self, tok):
        self.frameCfg_start_index = int(tok[1])
        self.frameCfg_end_index = int(tok[2])
        self.frameCfg_loops = int(tok[3])
        self.frameCfg_frames = int(tok[4])
        self.frameCfg_periodicity_ms = float(tok[5])
        self.frameCfg_hardware_trigger_enabled = int(tok[6]) != 1
        self.frameCfg_trigger_delay_ms = float(tok[7])

    _HANDLERS = {
        "channelCfg": _on_channel,
        "adcCfg": _on_adc,
        "adcbufCfg": _on_adcbuf,
        "profileCfg": _on_profile,
        "chirpCfg": _on_chirp,
        "frameCfg": _on_frame,
    }
