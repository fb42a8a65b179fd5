import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_cfg
from mmwave_radar_processing_b200.engine import cfar_params, RadarEngine
from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
def eng_env(cm, v):
    os.environ["RADAR_B200_CFAR_LADDER"] = str(v)
    e = RadarEngine(cm.num_rx_antennas, cm.get_num_adc_samples(0), cm.chirp_cfgs_per_loop, cm.frameCfg_loops)
    del os.environ["RADAR_B200_CFAR_LADDER"]
    return e
rng = np.random.default_rng(11)
cm = load_cfg("6843_RadVel_ods_20Hz", "ods")
lad, ref = eng_env(cm, 1), eng_env(cm, 0)
iq = torch.from_numpy(synth_batch_for_cfg(cm, range(12))).cuda()
_, mag = lad.range_doppler(iq, want_rd=False, mag_antenna=0)
s, l = mag.shape[-2:]
extra = [np.abs(rng.normal(size=(s, l)) + 1j * rng.normal(size=(s, l))) * 1e3, rng.normal(size=(s, l)) - 0.5, np.full((s, l), 3.0), np.zeros((s, l)),
         rng.exponential(1.0, (s, l)) * np.exp(rng.normal(size=(s, l)) * 3.0)]
m = rng.exponential(1.0, (s, l)); m[s // 2, l // 2] = np.nan; m[s // 3, l // 3] = np.inf; extra.append(m)
maps = torch.cat([mag, torch.from_numpy(np.stack(extra).astype(np.float32)).cuda()])
p = cfar_params("os", (8, 4), (2, 1), rho=0.75, alpha=3.5)
a = lad.cfar(maps, p, want_mask=True); b = ref.cfar(maps, p, want_mask=True)
for i in range(maps.shape[0]):
    d = (a["mask"][i] != b["mask"][i]).nonzero()
    print(i, "med", float(maps[i].flatten().nan_to_num().median()), "dets", int(a["counts"][i]), int(b["counts"][i]), "diff cells", d.shape[0], d[:4].tolist())
