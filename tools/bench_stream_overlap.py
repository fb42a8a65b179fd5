#!/usr/bin/env python
"""One GPU, the 6843 ODS stream: one batch of F frames per step on one stream against two half batches on two streams
(own buffers each), to see whether CFAR / point cloud of one half hides under the range-Doppler kernel of the other.
One JSON line per mode."""
import json
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    from bench import DEFAULT_WORKLOAD, DETECTOR, MAX_DETS, N_UNIQUE, WORKLOADS, load_cfg
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    F = int(os.environ.get("FRAMES", "32768"))
    cfg_name, geom, az, el, ra_ants = WORKLOADS[DEFAULT_WORKLOAD]
    cm = load_cfg(cfg_name, geom)
    uniq = synth_batch_for_cfg(cm, range(N_UNIQUE))
    iq = torch.from_numpy(np.concatenate([uniq] * (F // N_UNIQUE))).cuda()

    def make():
        return FramePipeline(cm, detector=DetectorSpec(**DETECTOR), az_antenna_idxs=az, el_antenna_idxs=el, max_dets=MAX_DETS,
                             with_range_angle=True, ra_antennas=ra_ants)
    steps = 10
    for parts in (1, 2, 4):
        pipes = [make() for _ in range(parts)]
        streams = [torch.cuda.Stream() for _ in range(parts)]
        chunks = list(iq.chunk(parts))
        for p, c in zip(pipes, chunks):
            for _ in range(2):
                p.run(c)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for st in streams:
            st.wait_stream(torch.cuda.current_stream())
        for _ in range(steps):
            for p, c, st in zip(pipes, chunks, streams):
                with torch.cuda.stream(st):
                    p.run(c)
        for st in streams:
            torch.cuda.current_stream().wait_stream(st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        print(json.dumps({"streams": parts, "frames_per_step": F, "ms_per_step": ms, "frames_per_s": F / ms * 1e3}), flush=True)
        del pipes
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
