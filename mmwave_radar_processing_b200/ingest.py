"""On-disk int16 ADC frame batches -> pinned host chunks -> FramePipeline.run_host (first part).

The reference reads recordings through ``cpsl_datasets`` (mmwave_radar_processing/visualization/models/dataset_model.py:11,
67, 91, 106), a submodule absent from the reference tree, so its on-disk format is unknown and is NOT guessed here.  This
module fixes the one layout the device path already consumes (``RADAR_IN_IQ16``):

    header  64 bytes: magic b"RB2IQ16\\0", version u32 = 1, num_rx u32, num_samples u32, chirps_per_frame u32,
                      n_frames u64, reserved (zeros)
    body    int16 [frame][chirp][rx][sample][2]   (I, Q interleaved; capture order of a DCA1000-style LVDS dump after
            de-interleaving the lanes), little endian

``FrameFile.write`` produces such a file from an array, ``FrameFile(path)`` memory-maps one, ``FrameFile.chunks`` yields
pinned host tensors of whole frames (double buffered, so the copy of chunk i+1 from the page cache overlaps the device
work on chunk i), and ``process_file`` drives ``FramePipeline.run_host`` over the file and returns the per-frame clouds.
A converter from the reference's ``(rx, sample, chirp)`` complex cubes is ``cube_to_iq16``.
"""
from __future__ import annotations

import struct
from typing import Iterator, List, Tuple

import numpy as np
import torch

MAGIC = b"RB2IQ16\0"
HEADER_BYTES = 64
_HEADER = struct.Struct("<8sIIIIQ")


def cube_to_iq16(cube: np.ndarray) -> np.ndarray:
    """Reference ADC cube ``(rx, sample, chirp)`` complex (integer valued, as the ADC delivers it) -> int16
    ``[chirp][rx][sample][2]``.  Raises if a value is not an int16 (the device ingest is exact only for ADC integers)."""
    c = np.asarray(cube)
    if c.ndim != 3 or not np.iscomplexobj(c):
        raise ValueError("cube must be a complex (rx, sample, chirp) array")
    re, im = np.rint(c.real), np.rint(c.imag)
    if np.abs(re - c.real).max(initial=0.0) > 0 or np.abs(im - c.imag).max(initial=0.0) > 0:
        raise ValueError("cube holds non-integer samples; the int16 ingest would round them")
    if max(np.abs(re).max(initial=0.0), np.abs(im).max(initial=0.0)) > 32767:
        raise ValueError("cube exceeds the int16 range")
    out = np.empty((c.shape[2], c.shape[0], c.shape[1], 2), dtype=np.int16)
    out[..., 0] = re.transpose(2, 0, 1)
    out[..., 1] = im.transpose(2, 0, 1)
    return out


class FrameFile:
    """Memory-mapped frame batch in the layout above."""

    def __init__(self, path: str):
        self.path = path
        with open(path, "rb") as f:
            head = f.read(HEADER_BYTES)
        if len(head) < HEADER_BYTES:
            raise ValueError(f"{path}: shorter than the {HEADER_BYTES}-byte header")
        magic, version, rx, samples, chirps, frames = _HEADER.unpack(head[:_HEADER.size])
        if magic != MAGIC:
            raise ValueError(f"{path}: not an RB2IQ16 frame file")
        if version != 1:
            raise ValueError(f"{path}: unsupported version {version}")
        self.num_rx, self.num_samples, self.chirps_per_frame, self.n_frames = int(rx), int(samples), int(chirps), int(frames)
        self.frame_shape = (self.chirps_per_frame, self.num_rx, self.num_samples, 2)
        want = HEADER_BYTES + 2 * self.n_frames * int(np.prod(self.frame_shape))
        import os
        have = os.path.getsize(path)
        if have < want:
            raise ValueError(f"{path}: truncated ({have} bytes, header promises {want})")
        self.data = np.memmap(path, dtype="<i2", mode="r", offset=HEADER_BYTES, shape=(self.n_frames,) + self.frame_shape)

    @staticmethod
    def write(path: str, frames: np.ndarray) -> None:
        """frames: int16 [frame][chirp][rx][sample][2]."""
        a = np.ascontiguousarray(frames)
        if a.dtype != np.int16 or a.ndim != 5 or a.shape[-1] != 2:
            raise ValueError("frames must be int16 [frame][chirp][rx][sample][2]")
        head = _HEADER.pack(MAGIC, 1, a.shape[2], a.shape[3], a.shape[1], a.shape[0])
        with open(path, "wb") as f:
            f.write(head + b"\0" * (HEADER_BYTES - len(head)))
            f.write(a.astype("<i2", copy=False).tobytes())

    def check_engine(self, engine) -> None:
        if (engine.C, engine.R, engine.S) != (self.chirps_per_frame, self.num_rx, self.num_samples):
            raise ValueError(f"{self.path}: frames are (chirps, rx, samples) = {(self.chirps_per_frame, self.num_rx, self.num_samples)}, "
                             f"the radar configuration expects {(engine.C, engine.R, engine.S)}")

    def chunks(self, chunk_frames: int, pinned: bool = True) -> Iterator[Tuple[int, torch.Tensor]]:
        """Yields ``(first_frame, int16 tensor [n, chirp, rx, sample, 2])``; two alternating pinned buffers, so a yielded
        tensor stays valid until the one after next is requested."""
        bufs = [torch.empty((chunk_frames,) + self.frame_shape, dtype=torch.int16) for _ in range(2)]
        if pinned and torch.cuda.is_available():
            bufs = [b.pin_memory() for b in bufs]
        for k, lo in enumerate(range(0, self.n_frames, chunk_frames)):
            hi = min(self.n_frames, lo + chunk_frames)
            buf = bufs[k & 1][:hi - lo]
            buf.numpy()[...] = self.data[lo:hi]          # page cache -> pinned memory, one pass
            yield lo, buf


def process_file(pipe, path: str, batch_frames: int = 4096, chunk_frames: int = 512, points_per_frame: int = 64) -> List[np.ndarray]:
    """Point clouds of every frame of ``path`` through ``FramePipeline.run_host`` (the end-to-end API of bench.py's
    ``e2e`` leg): list of ``(n_i, 4)`` float64 arrays ``[x, y, z, v]`` in frame order."""
    ff = FrameFile(path)
    ff.check_engine(pipe.engine)
    clouds: List[np.ndarray] = []
    counts = torch.zeros((batch_frames,), dtype=torch.int32)
    points = torch.zeros((batch_frames * points_per_frame, 4), dtype=torch.float64)
    if torch.cuda.is_available():
        counts, points = counts.pin_memory(), points.pin_memory()
    for lo, iq in ff.chunks(batch_frames):
        n = iq.shape[0]
        totals, regions, _, _ = pipe.run_host(iq, counts[:n], points, chunk_frames=chunk_frames)
        # the staging buffers are reused by the next batch: the clouds are copied out
        clouds.extend(c.copy() for c in pipe.split_host(points, counts[:n], totals, regions, chunk_frames, pipe.max_dets))
    return clouds
