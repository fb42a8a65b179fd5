"""Run under torchrun on >= 2 GPUs (tests/test_gpu_multi.py launches it): the voxel-sharded synthetic-array image
(NCCL broadcast of the aperture cube + geometry from rank 0, radar_sar_image(g_begin, g_end) per rank, NCCL all-gather of
the tiles) must equal the image one GPU computes alone, on every rank - and so must the collective-free variant
(distributed.PeerImageExchange: the range-FFT kernel stores its tile into every rank's image over NVLink peer memory).
Reference arithmetic: processors/simple_synthetic_array_beamformer_processor_multiFrame.py:543-585."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    from mmwave_radar_processing_b200 import distributed as rdist
    from mmwave_radar_processing_b200.engine import RadarEngine
    rank, local, world = rdist.init_from_env("nccl")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    for s, n, g in ((90, 500, 90), (90, 2000, 3001), (64, 333, 777)):        # ragged shards included
        rng = np.random.default_rng(7)
        x = (rng.integers(-2000, 2000, (s, n)) + 1j * rng.integers(-2000, 2000, (s, n))).astype(np.complex64)
        t = np.arange(n) * 1e-4
        pos = np.stack([0.01 * t, 0.5 * t, 0.002 * np.sin(t * 40.0)])
        az = np.linspace(-1.0, 1.0, g)
        el = 0.2 * np.cos(np.arange(g) * 0.37)
        dirs = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)])
        lam = 0.0039
        eng = RadarEngine(1, s, 1, 1, device=dev)
        ds = torch.from_numpy(dirs).to(dev)
        if rank == 0:
            xs, ps = torch.from_numpy(x).to(dev), torch.from_numpy(pos).to(dev)
        else:                                               # the other ranks start with garbage: the broadcast must fill it
            xs = torch.full((s, n), 7.0 + 1.0j, dtype=torch.complex64, device=dev)
            ps = torch.full((3, n), -3.0, dtype=torch.float64, device=dev)
        img, (lo, hi) = rdist.sar_image_sharded(lambda ap, po, di, a, b: eng.sar_image(ap, po, di, lam, a, b), xs, ps, ds,
                                                src=0, gather=True)
        assert (lo, hi) == rdist.shard_range(g, rank, world)
        full = eng.sar_image(torch.from_numpy(x).to(dev), torch.from_numpy(pos).to(dev), ds, lam, 0, g)
        torch.cuda.synchronize()
        assert img.shape == full.shape == (s, g), (img.shape, full.shape)
        err = float((img - full).abs().max() / full.abs().max())
        # slices and the full grid tile the directions differently; the kernel's tilings agree to rounding of the segment sums
        assert err < 2e-6, (rank, s, n, g, err)
        assert torch.equal(img[:, lo:hi], eng.sar_image(xs, ps, ds, lam, lo, hi)), (rank, "own tile differs")
        dist.barrier()
        # the same image without a collective: the range-FFT kernel stores every cell into the full image of every rank
        # (radar_ctx_set_image_mirrors over symmetric memory), the kernels read rank 0's aperture over NVLink
        ex = rdist.PeerImageExchange(eng, s, n, g, depth=2, src=0)
        for k in range(3):                                  # k = 2 reuses slot 0
            scale = float(k + 1)
            fused = ex.run(k, ds, lam, xs * scale if rank == 0 else None, ps if rank == 0 else None)
            torch.cuda.synchronize()
            assert fused.shape == (s, g)
            want = img * scale                              # power-of-two / small-integer scale of the input: exact for k = 0, 1
            if k < 2:
                assert torch.equal(fused, want), (rank, s, n, g, k, "fused exchange differs from the NCCL all-gather")
            else:
                e3 = float((fused - want).abs().max() / want.abs().max())
                assert e3 < 2e-6, (rank, k, e3)
            dist.barrier()
        ex.disarm()
        # two images in flight: one stream and one context per image slot, consecutive images overlap
        exp = rdist.PeerImageExchange(eng, s, n, g, depth=2, src=0, pipelined=True)
        for pair in range(2):
            keep = []                                           # inputs stay alive until their images are done
            for k in (2 * pair, 2 * pair + 1):
                scale = float(2 ** (k % 3))
                xin = xs * scale if rank == 0 else None
                keep.append(xin)
                exp.run(k, ds, lam, xin, ps if rank == 0 else None)
            for k in (2 * pair, 2 * pair + 1):
                exp.wait(k)
                torch.cuda.synchronize()
                assert torch.equal(exp.image(k), img * float(2 ** (k % 3))), (rank, s, n, g, k, "pipelined exchange differs")
            dist.barrier()
        exp.disarm()
        after = eng.sar_image(xs, ps, ds, lam, lo, hi)      # mirrors off again: a plain call
        assert torch.equal(after, img[:, lo:hi])
        dist.barrier()
    if rank == 0:
        print("SAR_SHARDED_OK world", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
