"""Multi-GPU plumbing: one process per GPU, ``torch.distributed`` (NCCL over NVLink on the GPU box, gloo in
the CPU tests).  Only the two exchange steps the path really has:

  * streaming frames are sharded by contiguous frame blocks; the variable-length per-rank point clouds are
    exchanged with a counts all-gather followed by a fixed-capacity all-gather of the packed rows;
  * SAR / DBS image grids are sharded by voxel (direction) after a broadcast of the aperture cube and the
    element geometry; the image tiles are all-gathered when every rank wants the full image.

There is no data-path collective inside the kernels: frames and voxels are independent.
"""
from __future__ import annotations

import os
from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


def init_from_env(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """Join the process group described by RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* (torchrun).
    Returns (rank, local_rank, world).  A single process (no WORLD_SIZE or 1) does not create a group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            kw["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    elif torch.cuda.is_available():
        torch.cuda.set_device(local_rank)
    return rank, local_rank, world


def world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of ``n`` items owned by ``rank``; the first ``n % world`` ranks get one extra.
    Contiguous so a cross-frame recurrence (the ground detector's altimeter) stays inside a block."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


# -------------------------------------------------------------------------------------------- streaming frames
def allgather_point_clouds(points_packed: torch.Tensor, counts: torch.Tensor, total: torch.Tensor, cap_rows: int,
                           out_points: Optional[torch.Tensor] = None, out_counts: Optional[torch.Tensor] = None,
                           out_totals: Optional[torch.Tensor] = None):
    """Exchange the per-rank point clouds of one batch.

    points_packed [>=cap_rows, 4] f64: this rank's packed rows (frame after frame); counts int32 [F_local];
    total int32 [1] = number of valid rows.  Every rank contributes exactly ``cap_rows`` rows (NCCL has no
    all-gather-v): rows beyond ``total`` are padding.  Returns (points [W, cap_rows, 4], counts [W, F_local],
    totals [W]) - all device tensors, no host synchronisation.  The caller checks ``totals.max() <= cap_rows``
    when it next touches the host (``check_capacity``) and re-gathers with a larger capacity if not.
    Equal F_local on every rank is required (pad the last block with empty frames otherwise)."""
    w = world_size()
    f = counts.shape[0]
    dev = points_packed.device
    if out_points is None:
        out_points = torch.empty((w, cap_rows, 4), dtype=points_packed.dtype, device=dev)
    if out_counts is None:
        out_counts = torch.empty((w, f), dtype=counts.dtype, device=dev)
    if out_totals is None:
        out_totals = torch.empty((w,), dtype=total.dtype, device=dev)
    if w == 1:
        out_points[0].copy_(points_packed[:cap_rows])
        out_counts[0].copy_(counts)
        out_totals.copy_(total)
        return out_points, out_counts, out_totals
    dist.all_gather_into_tensor(out_counts.view(-1), counts.contiguous())
    dist.all_gather_into_tensor(out_totals, total.contiguous())
    dist.all_gather_into_tensor(out_points.view(-1), points_packed[:cap_rows].contiguous().view(-1))
    return out_points, out_counts, out_totals


def check_capacity(totals: torch.Tensor, cap_rows: int) -> int:
    """Host-side check after a batch: largest per-rank total; raises if the gather truncated a rank."""
    m = int(totals.max().item())
    if m > cap_rows:
        raise RuntimeError(f"a rank produced {m} points but the all-gather capacity is {cap_rows} rows")
    return m


def split_gathered(points: torch.Tensor, counts: torch.Tensor, max_dets: int) -> List:
    """[W, cap, 4], [W, F] -> list over the global frame order of (N_i, 4) numpy arrays."""
    pts = points.cpu().numpy()
    cnt = counts.cpu().numpy().clip(max=max_dets)
    clouds = []
    for r in range(cnt.shape[0]):
        row = 0
        for c in cnt[r]:
            clouds.append(pts[r, row:row + int(c)])
            row += int(c)
    return clouds


# -------------------------------------------------------------------------------------------- SAR / DBS grids
def sar_image_sharded(compute_slice: Callable[[torch.Tensor, torch.Tensor, torch.Tensor, int, int], torch.Tensor],
                      aperture: torch.Tensor, positions: torch.Tensor, dirs: torch.Tensor, src: int = 0,
                      gather: bool = True, broadcast: bool = True):
    """Voxel-sharded synthetic-array image.

    Rank ``src`` holds the aperture cube [S, N] c64 and the geometry [3, N] f64; they are broadcast, every rank
    computes directions [g_lo, g_hi) with ``compute_slice(aperture, positions, dirs, g_lo, g_hi)`` -> [S, g_hi-g_lo]
    and, if ``gather``, the tiles are all-gathered into the full [S, G] image on every rank (tiles are padded to
    the widest shard for the collective)."""
    w = world_size()
    rank = dist.get_rank() if dist.is_initialized() else 0
    if w > 1 and broadcast:
        # complex tensors travel as their real view (NCCL / gloo have no complex type)
        dist.broadcast(torch.view_as_real(aperture) if aperture.is_complex() else aperture, src=src)
        dist.broadcast(positions, src=src)
    g = dirs.shape[1]
    lo, hi = shard_range(g, rank, w)
    tile = compute_slice(aperture, positions, dirs, lo, hi)
    if not gather or w == 1:
        return tile, (lo, hi)
    # Tiles travel as they are, [S][width] row-major into slot `rank` of a [world][S][widest] buffer (no transposes; ragged
    # shards are padded on the right), then ONE strided copy lays the slots side by side as the [S][G] image.
    s = tile.shape[0]
    widest = shard_range(g, 0, w)[1]
    if hi - lo == widest and tile.is_contiguous():
        mine = tile
    else:
        mine = torch.zeros((s, widest), dtype=tile.dtype, device=tile.device)
        mine[:, :hi - lo] = tile
    allt = torch.empty((w, s, widest), dtype=tile.dtype, device=tile.device)
    # complex tensors travel as their real view (NCCL / gloo have no complex all-gather)
    if tile.is_complex():
        dist.all_gather_into_tensor(torch.view_as_real(allt).view(-1), torch.view_as_real(mine).reshape(-1))
    else:
        dist.all_gather_into_tensor(allt.view(-1), mine.reshape(-1))
    if g == w * widest:
        image = allt.permute(1, 0, 2).reshape(s, g)            # one copy: [S][world][width] -> [S][G]
    else:
        image = torch.empty((s, g), dtype=tile.dtype, device=tile.device)
        for r in range(w):
            rlo, rhi = shard_range(g, r, w)
            image[:, rlo:rhi] = allt[r, :, :rhi - rlo]
    return image, (lo, hi)


class PeerPointExchange:
    """The all-gather of the streaming path's point clouds over NVLink peer memory, with the point rows stored by the
    point-cloud kernel itself (``radar_ctx_set_point_mirrors``): no staging copy and no collective kernel competing with
    the next step for SMs.  One symmetric allocation (``torch.distributed._symmetric_memory``) holds ``depth`` gathered
    slots on every rank; slot layout per (k, source rank): points float64 [cap_rows, 4] | counts int32 [F_local] |
    total int32 [1].  Per step:

        ex.arm(k)                      # the next radar_point_cloud call also stores its packed rows into slot k of every rank
        out = pipe.run(...)            # kernels (main stream)
        ex.publish(k, out["counts"], out["offsets"][F:F+1])   # counts / total by peer copies, then the cross-rank barrier

    ``gathered(k)`` are views of this rank's slot k: valid once ``publish(k)`` has completed on the stream it ran on; the
    reader must be done with slot k before this rank's ``publish`` of the step after next (depth 2)."""

    def __init__(self, engine, n_frames_local: int, cap_rows: int, depth: int = 2):
        import torch.distributed._symmetric_memory as symm
        self.engine, self.f, self.cap, self.depth = engine, int(n_frames_local), int(cap_rows), int(depth)
        self.world, self.rank = world_size(), dist.get_rank()
        self.pts_bytes = self.cap * 32
        self.cnt_off = self.pts_bytes
        self.tot_off = self.cnt_off + 4 * self.f
        self.slot = (self.tot_off + 4 + 255) // 256 * 256
        dev = engine.device
        self.buf = symm.empty((self.depth, self.world, self.slot), dtype=torch.uint8, device=dev)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, dist.group.WORLD)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.peers = [self.hdl.get_buffer(r, self.buf.shape, torch.uint8) for r in range(self.world)]
        torch.cuda.synchronize(dev)
        self.hdl.barrier(channel=0)

    def arm(self, k: int, remote: bool = True):
        """remote False: only this rank's own gathered slot is a mirror; the peers get the rows from ``publish(points=...)``."""
        off = (k * self.world + self.rank) * self.slot
        ptrs = self.ptrs if remote else [self.ptrs[self.rank]]
        self.engine.set_point_mirrors([p + off for p in ptrs], self.cap)

    def disarm(self):
        self.engine.set_point_mirrors([], 0)

    def publish(self, k: int, counts: torch.Tensor, total: torch.Tensor, push_points: bool = False,
                rows: Optional[int] = None):
        """On the current stream, after the point-cloud kernel of this step (stream order or an event wait).
        ``push_points``: this rank's rows (its own slot k, filled by the kernel's local mirror) are pushed to the OTHER
        ranks by peer-to-peer copies (copy engines) - the variant for ``arm(k, remote=False)``, where the kernel's own
        stores stay on the local GPU.  ``rows``: push only the first ``rows`` rows of the slot instead of its whole
        capacity (the caller's estimate from earlier steps); the totals travel with the barrier, so a reader that finds
        ``totals.max() > rows`` (``check_capacity(totals, rows)``) must ask for that step again with ``rows=None``."""
        c8 = counts.contiguous().view(torch.uint8)
        t8 = total.contiguous().view(torch.uint8)
        nbytes = self.pts_bytes if rows is None else min(self.pts_bytes, int(rows) * 32)
        p8 = self.buf[k, self.rank, :nbytes] if push_points else None
        for i in range(self.world):
            dst = self.peers[(self.rank + i) % self.world][k, self.rank]
            if p8 is not None and i > 0:
                dst[:nbytes].copy_(p8, non_blocking=True)
            dst[self.cnt_off:self.cnt_off + c8.numel()].copy_(c8, non_blocking=True)
            dst[self.tot_off:self.tot_off + 4].copy_(t8, non_blocking=True)
        self.hdl.barrier(channel=k % 8)

    def gathered(self, k: int):
        """-> (points [W, cap_rows, 4] f64, counts [W, F_local] i32, totals [W] i32) views of this rank's slot k."""
        s = self.buf[k]
        pts = s[:, :self.pts_bytes].view(torch.float64).view(self.world, self.cap, 4)
        cnt = s[:, self.cnt_off:self.tot_off].view(torch.int32)
        tot = s[:, self.tot_off:self.tot_off + 4].view(torch.int32).view(self.world)
        return pts, cnt, tot


class PeerImageExchange:
    """The voxel-sharded synthetic-array image over NVLink peer memory, without a collective: one symmetric allocation
    holds, per image slot (``depth`` of them), the aperture cube + geometry of the feeding rank and a full [S][G] image on
    every rank.

        ex.put_inputs(k, aperture, positions)   # rank ``src`` only: into slot k of its symmetric region; then every rank:
        ex.inputs_ready(k)                      # cross-rank barrier on the slot's stream
        ap, po = ex.inputs(k)                   # this rank's copy, pulled from rank src's region by one peer copy each (1.5 MB)
        ex.arm(k)                               # radar_sar_image also stores its cells into image k of EVERY rank, at this rank's columns
        engine_k.sar_image(ap, po, dirs, lam, lo, hi, out=tile_k)
        ex.finish(k)                            # cross-rank barrier: all tiles of image k have landed everywhere
        img = ex.image(k)                       # [S][G] complex64 view of this rank's image k

    (``run(k, ...)`` is that sequence.)  The all-gather of the tiles is the store stream of the range-FFT kernel
    (``radar_ctx_set_image_mirrors``), the broadcast is a pull by the copy engines (kernels that load the cube straight
    from the peer run 1.5x longer: the small preparation kernels are latency chains).  ``finish(k)`` also fences slot k's
    inputs; the reader must be done with image k before ``run`` of step k + depth.

    ``pipelined=True`` gives every slot its own stream and its own ``RadarEngine`` (the scratch of a context serves one
    image at a time), so consecutive images overlap: the peer stores and barriers of image k run under the GEMM of image
    k + 1.  ``image(k)`` is then valid once ``streams[k % depth]`` has reached the end of ``run(k)``."""

    def __init__(self, engine, num_samples: int, n_elems: int, n_dirs: int, depth: int = 2, src: int = 0,
                 pipelined: bool = False):
        import torch.distributed._symmetric_memory as symm
        self.s, self.n, self.g, self.depth, self.src = int(num_samples), int(n_elems), int(n_dirs), int(depth), int(src)
        self.world, self.rank = world_size(), dist.get_rank()
        up = lambda v: (v + 255) // 256 * 256                                  # noqa: E731
        self.ap_bytes, self.pos_bytes, self.img_bytes = self.s * self.n * 8, 3 * self.n * 8, self.s * self.g * 8
        self.pos_off = up(self.ap_bytes)
        self.img_off = self.pos_off + up(self.pos_bytes)
        self.slot = self.img_off + up(self.img_bytes)                          # bytes of one slot: inputs | image
        dev = engine.device
        self.engines = [engine]
        self.streams = [None] * self.depth
        if pipelined:
            from .engine import RadarEngine
            self.engines += [RadarEngine(engine.R, engine.S, engine.cpl, engine.L, device=dev) for _ in range(self.depth - 1)]
            self.streams = [torch.cuda.Stream(device=dev) for _ in range(self.depth)]
        self.pipelined = bool(pipelined)
        self.buf = symm.empty((self.depth * self.slot,), dtype=torch.uint8, device=dev)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, dist.group.WORLD)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.src_buf = self.hdl.get_buffer(self.src, self.buf.shape, torch.uint8)
        self.lo, self.hi = shard_range(self.g, self.rank, self.world)
        self.tiles = [torch.empty((self.s, self.hi - self.lo), dtype=torch.complex64, device=dev) for _ in range(self.depth)]
        self.ap_local = [torch.zeros((self.s, self.n), dtype=torch.complex64, device=dev) for _ in range(self.depth)]
        self.pos_local = [torch.zeros((3, self.n), dtype=torch.float64, device=dev) for _ in range(self.depth)]
        torch.cuda.synchronize(dev)
        self.hdl.barrier(channel=0)
        torch.cuda.synchronize(dev)

    def engine_of(self, k: int):
        return self.engines[(k % self.depth) % len(self.engines)]

    def stream_of(self, k: int):
        st = self.streams[k % self.depth]
        return torch.cuda.current_stream() if st is None else st

    def _views(self, base: torch.Tensor, k: int):
        off = (k % self.depth) * self.slot
        return (base[off:off + self.ap_bytes].view(torch.complex64).view(self.s, self.n),
                base[off + self.pos_off:off + self.pos_off + self.pos_bytes].view(torch.float64).view(3, self.n))

    def put_inputs(self, k: int, aperture: torch.Tensor, positions: torch.Tensor):
        if self.rank != self.src:
            return
        ap, po = self._views(self.buf, k)
        ap.copy_(aperture, non_blocking=True)
        po.copy_(positions, non_blocking=True)

    def inputs_ready(self, k: int):
        self.hdl.barrier(channel=1 + (k % self.depth))

    def inputs(self, k: int):
        """After ``inputs_ready(k)``: pull rank src's aperture cube + geometry of slot k into this rank's memory (rank src
        uses its own region in place) -> (aperture [S, N] c64, positions [3, N] f64)."""
        if self.rank == self.src:
            return self._views(self.buf, k)
        ap, po = self._views(self.src_buf, k)
        self.ap_local[k % self.depth].copy_(ap, non_blocking=True)
        self.pos_local[k % self.depth].copy_(po, non_blocking=True)
        return self.ap_local[k % self.depth], self.pos_local[k % self.depth]

    def arm(self, k: int):
        off = (k % self.depth) * self.slot + self.img_off
        self.engine_of(k).set_image_mirrors([p + off for p in self.ptrs], self.g, self.lo)

    def disarm(self):
        for e in self.engines:
            e.set_image_mirrors([], 0, 0)

    def finish(self, k: int):
        self.hdl.barrier(channel=1 + self.depth + (k % self.depth))

    def image(self, k: int) -> torch.Tensor:
        off = (k % self.depth) * self.slot + self.img_off
        return self.buf[off:off + self.img_bytes].view(torch.complex64).view(self.s, self.g)

    def _run_on_current_stream(self, k: int, dirs, lambda_m, aperture, positions):
        if aperture is not None:
            self.put_inputs(k, aperture, positions)
        self.inputs_ready(k)
        ap, po = self.inputs(k)
        self.arm(k)
        self.engine_of(k).sar_image(ap, po, dirs, lambda_m, self.lo, self.hi, out=self.tiles[k % self.depth])
        self.finish(k)

    def run(self, k: int, dirs: torch.Tensor, lambda_m: float, aperture: Optional[torch.Tensor] = None,
            positions: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One image: rank ``src`` passes the aperture cube + geometry, the others None.  Pipelined: the work is queued on
        the slot's stream after whatever the current stream has queued so far (the inputs)."""
        st = self.streams[k % self.depth]
        if st is None:
            self._run_on_current_stream(k, dirs, lambda_m, aperture, positions)
        else:
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                self._run_on_current_stream(k, dirs, lambda_m, aperture, positions)
        return self.image(k)

    def wait(self, k: int):
        """Make the current stream wait for image k (pipelined mode; a no-op otherwise)."""
        st = self.streams[k % self.depth]
        if st is not None:
            torch.cuda.current_stream().wait_stream(st)

    def capture(self, dirs: torch.Tensor, lambda_m: float, aperture: Optional[torch.Tensor] = None,
                positions: Optional[torch.Tensor] = None):
        """CUDA graphs of ``run`` for the ``depth`` image slots (one image is ten short launches: at these sizes the host
        would otherwise set the pace).  Call after a few eager ``run`` calls (scratch sized, symmetric memory warm) with the
        tensors that will hold the inputs of every later image; returns ``replay(k) -> image k``.  Every rank must capture
        and replay in step.  Pipelined: slot k's graph is replayed on its stream (after what the current stream has
        queued), so consecutive replays overlap; ``wait(k)`` joins."""
        torch.cuda.synchronize()
        graphs = []
        for k in range(self.depth):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                self._run_on_current_stream(k, dirs, lambda_m, aperture, positions)
            graphs.append(gr)

        def replay(k: int) -> torch.Tensor:
            st = self.streams[k % self.depth]
            if st is None:
                graphs[k % self.depth].replay()
            else:
                st.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(st):
                    graphs[k % self.depth].replay()
            return self.image(k)
        return replay


class LocalImagePipeline:
    """One GPU, ``depth`` synthetic-array images in flight: every slot has its own ``RadarEngine`` (the scratch of a context
    serves one image at a time), stream, output image and CUDA graph, so the max / prep / range-FFT kernels of one image
    run under the GEMM of the next (measured 0.182 -> 0.151 ms per image with two slots, tools/bench_sar_overlap.py).
    ``run(k)`` queues image k on its slot's stream after what the current stream has queued (the inputs); ``wait(k)`` makes
    the current stream wait for it; the reader must be done with image k before ``run(k + depth)``."""

    def __init__(self, engine, n_dirs: int, depth: int = 2):
        from .engine import RadarEngine
        dev = engine.device
        self.depth = int(depth)
        self.engines = [engine] + [RadarEngine(engine.R, engine.S, engine.cpl, engine.L, device=dev) for _ in range(self.depth - 1)]
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(self.depth)]
        self.images = [torch.empty((engine.S, int(n_dirs)), dtype=torch.complex64, device=dev) for _ in range(self.depth)]
        self.graphs = None

    def capture(self, aperture: torch.Tensor, positions: torch.Tensor, dirs: torch.Tensor, lambda_m: float):
        """After a few eager ``run`` calls: one CUDA graph per slot over the tensors that will hold every later input."""
        torch.cuda.synchronize()
        self.graphs = []
        for e, img in zip(self.engines, self.images):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                e.sar_image(aperture, positions, dirs, lambda_m, out=img)
            self.graphs.append(gr)

    def run(self, k: int, aperture: torch.Tensor, positions: torch.Tensor, dirs: torch.Tensor, lambda_m: float) -> torch.Tensor:
        i = k % self.depth
        st = self.streams[i]
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            if self.graphs is not None:
                self.graphs[i].replay()
            else:
                self.engines[i].sar_image(aperture, positions, dirs, lambda_m, out=self.images[i])
        return self.images[i]

    def wait(self, k: int):
        torch.cuda.current_stream().wait_stream(self.streams[k % self.depth])
