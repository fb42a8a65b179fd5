"""Run under torchrun on >= 2 GPUs (tests/test_gpu_multi.py launches it): the peer-memory exchange of the point clouds
(PeerPointExchange + radar_ctx_set_point_mirrors) must deliver exactly what the NCCL all-gather delivers."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))


def main():
    from conftest import load_cfg
    from mmwave_radar_processing_b200 import distributed as rdist
    from mmwave_radar_processing_b200.pipeline import DetectorSpec, FramePipeline
    from mmwave_radar_processing_b200.synthetic import synth_batch_for_cfg
    rank, local, world = rdist.init_from_env("nccl")
    dev = torch.device("cuda", local)
    cm = load_cfg("1843_RadVel")
    f = 48
    pipe = FramePipeline(cm, detector=DetectorSpec(kind="os", num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5),
                         az_antenna_idxs=(0, 1, 2, 3), el_antenna_idxs=(), max_dets=256, device=dev)
    iq = torch.from_numpy(synth_batch_for_cfg(cm, range(rank * f, rank * f + f + rank))[:f]).to(dev)   # different frames per rank
    out = pipe.run(iq)
    total = out["offsets"][f:f + 1]
    t = total.to(torch.int64).clone()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    cap = int((int(t.item()) * 5 // 4 + 63) // 64 * 64)
    want = rdist.allgather_point_clouds(out["points"], out["counts"], total, cap)
    want = [w.clone() for w in want]
    ex = rdist.PeerPointExchange(pipe.engine, f, cap)
    for step in range(4):                                   # both slots, twice
        k = step & 1
        ex.arm(k)
        out = pipe.run(iq)
        ex.publish(k, out["counts"], out["offsets"][f:f + 1])
        torch.cuda.synchronize()
        pts, cnt, tot = ex.gathered(k)
        assert torch.equal(tot, want[2]), (rank, step, tot, want[2])
        assert torch.equal(cnt, want[1]), (rank, step)
        for r in range(world):
            n = int(tot[r])
            assert n > 0 and torch.equal(pts[r, :n], want[0][r, :n]), (rank, step, r)
        dist.barrier()
    for step in range(3):                                   # kernel stores stay local, the peers get copy-engine pushes
        k = step % 2
        ex.buf[k].zero_()
        torch.cuda.synchronize()
        dist.barrier()
        ex.arm(k, remote=False)
        out = pipe.run(iq)
        ex.publish(k, out["counts"], out["offsets"][f:f + 1], push_points=True)
        torch.cuda.synchronize()
        pts, cnt, tot = ex.gathered(k)
        assert torch.equal(tot, want[2]) and torch.equal(cnt, want[1]), (rank, step)
        for r in range(world):
            n = int(tot[r])
            assert torch.equal(pts[r, :n], want[0][r, :n]), (rank, step, r)
        dist.barrier()
    ex.disarm()
    # capacity smaller than the rank's rows: rows beyond it are dropped on the mirrors, the local output is complete
    small = rdist.PeerPointExchange(pipe.engine, f, 64)
    small.arm(0)
    out = pipe.run(iq)
    small.publish(0, out["counts"], out["offsets"][f:f + 1])
    torch.cuda.synchronize()
    pts, _, tot = small.gathered(0)
    for r in range(world):
        assert torch.equal(pts[r, :64], want[0][r, :64]) and int(tot[r]) == int(want[2][r])
    small.disarm()
    dist.barrier()
    if rank == 0:
        print("PEER_EXCHANGE_OK world", world, "totals", want[2].tolist(), "cap", cap)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
