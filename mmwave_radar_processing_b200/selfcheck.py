"""Smoke check used by ``__graft_entry__.smoke()``: one small pass of the hot path on cuda:0, compared with
the oracle (the oracle is imported here ONLY as the checker)."""
from __future__ import annotations

import os
import sys

import numpy as np


def run_smoke(verbose: bool = False, cfg_name: str = "6843_RadVel_ods_20Hz", n_frames: int = 4):
    import torch
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if repo not in sys.path:
        sys.path.insert(0, repo)
    from oracle import radar_oracle as orc
    from .config_managers.cfgManager import ConfigManager
    from .pipeline import DetectorSpec, FramePipeline
    from .synthetic import synth_batch_for_cfg

    cfg = ConfigManager()
    cfg.load_cfg(os.path.join(repo, "configs", cfg_name + ".cfg"), array_geometry="ods")
    iq = synth_batch_for_cfg(cfg, range(n_frames))
    pipe = FramePipeline(cfg, detector=DetectorSpec("ca", (4, 4), (2, 2), pfa=1e-5), az_antenna_idxs=(0, 3, 4, 7),
                         el_antenna_idxs=(9, 8, 5, 4), max_dets=512, device="cuda:0")
    out = pipe.run(torch.from_numpy(iq).to("cuda:0"))
    torch.cuda.synchronize()
    clouds = pipe.gather_host(out)
    rd_dev = out["rd"].cpu().numpy()
    worst = 0.0
    for i in range(n_frames):
        cube = orc.virtual_array(orc.iq_to_cube(iq[i]), cfg.num_rx_antennas, cfg.chirp_cfgs_per_loop, cfg.frameCfg_loops)
        rd = orc.range_doppler(cube)
        err = np.linalg.norm(rd_dev[i] - rd) / np.linalg.norm(rd)
        worst = max(worst, err)
        _, _, mask = orc.cfar_2d(np.abs(rd[0]), "ca", (4, 4), (2, 2), pfa=1e-5)
        dets = orc.mask_to_dets(mask)
        pts, _, _ = orc.point_cloud(rd, dets, pipe.range_bins, pipe.vel_bins, [0, 3, 4, 7], [9, 8, 5, 4])
        if clouds[i].shape != pts.shape or not np.allclose(clouds[i], pts, rtol=1e-9, atol=1e-9):
            raise AssertionError(f"smoke: point cloud of frame {i} differs from the oracle "
                                 f"({clouds[i].shape} vs {pts.shape})")
    if worst > 1e-4:
        raise AssertionError(f"smoke: RD relative L2 error {worst:.3e} > 1e-4")
    if verbose:
        print(f"smoke ok: {n_frames} frames of {cfg_name}, RD rel-L2 {worst:.2e}, "
              f"{sum(c.shape[0] for c in clouds)} points identical to the oracle")
    return worst
