import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

GOLDEN_DIR = os.path.join(REPO, "tests", "golden")
CONFIG_DIR = os.path.join(REPO, "configs")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_cfg(name, geometry="standard"):
    from mmwave_radar_processing_b200.config_managers.cfgManager import ConfigManager
    cm = ConfigManager()
    cm.load_cfg(os.path.join(CONFIG_DIR, name + ".cfg"), array_geometry=geometry)
    return cm


def golden(name, seed):
    return np.load(os.path.join(GOLDEN_DIR, f"{name}_seed{seed}.npz"))


def frame_iq(cm, seed):
    from mmwave_radar_processing_b200.synthetic import synth_frame_iq
    return synth_frame_iq(cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop,
                          cm.frameCfg_loops, seed)


GOLDEN_CASES = [
    ("6843_RadVel_ods_20Hz", 0), ("6843_RadVel_ods_20Hz", 1), ("6843_RadVel_ods_20Hz", 2),
    ("1843_RadVel", 0), ("1843_RadVel", 1), ("RadarHD", 0), ("RadCloud", 0), ("RadSAR_1843_official", 0),
]


def ant_sets(v):
    if v >= 12:
        return dict(az_a=[0, 1, 2, 3], el_a=[], az_b=[0, 3, 4, 7], el_b=[9, 8, 5, 4], ra=[0, 3, 4, 7])
    if v >= 8:
        return dict(az_a=[0, 1, 2, 3], el_a=[], az_b=[0, 1, 2, 3, 4, 5, 6, 7], el_b=[1, 5], ra=[0, 3, 4, 7])
    return dict(az_a=[0, 1, 2, 3], el_a=[], az_b=[0, 1, 2], el_b=[3, 2], ra=[0, 1, 3])


CFAR2D_SETS = {
    "os_test": ("os", dict(num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5)),
    "ca_test": ("ca", dict(num_train=(4, 4), num_guard=(2, 2), pfa=1e-5)),
    "os_gui": ("os", dict(num_train=[5, 5], num_guard=[3, 2], rho=0.7, alpha=2)),
}
CFAR1D_SETS = {
    "ca": ("ca", dict(num_train=10, num_guard=2, pfa=1e-3)),
    "go": ("go", dict(num_train=10, num_guard=2, pfa=1e-3)),
    "so": ("so", dict(num_train=10, num_guard=2, pfa=1e-3)),
    "os": ("os", dict(num_train=5, num_guard=3, rho=0.5, alpha=2)),
}
SEQ_PARAMS = dict(rng_kind="os", rng_params=dict(num_train=5, num_guard=3, rho=0.6, alpha=2),
                  vel_kind="os", vel_params=dict(num_train=5, num_guard=2, rho=0.7, alpha=3))
GROUND_VEL = ("os", dict(num_train=12, num_guard=4, rho=0.5, alpha=15))


def rel_l2(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    den = np.linalg.norm(b.ravel())
    return float(np.linalg.norm((a - b).ravel()) / (den if den > 0 else 1.0))
