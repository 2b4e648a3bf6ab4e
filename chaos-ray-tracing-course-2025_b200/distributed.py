"""Multi-GPU stills and animations: one process per GPU (torch.distributed), scene replicated, work partitioned.

Stills   rows are cut into bands of ``band_rows`` rows; band b belongs to rank b % world (interleaving balances the
         load: hit density varies strongly over the image).  Every rank renders its bands into a COMPACT buffer
         (crt_render_params.rank/world/band_rows); the only exchange step of the path is the gather of those buffers
         to rank 0 (NCCL over NVLink on GPUs, gloo in the CPU tests), followed by the band un-interleave.
Frames   frame k of an animation belongs to rank k % world; no exchange at all (each rank keeps its frames).

torch is plumbing here (process group, device tensors wrapping OUR framebuffers); no pixel is computed by torch."""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def owned_rows(height: int, rank: int, world: int, band_rows: int) -> np.ndarray:
    if world <= 1:
        return np.arange(height)
    bands = (height + band_rows - 1) // band_rows
    rows = [np.arange(b * band_rows, min(height, (b + 1) * band_rows)) for b in range(rank, bands, world)]
    return np.concatenate(rows) if rows else np.zeros(0, dtype=np.int64)


def owned_frames(n_frames: int, rank: int, world: int) -> list[int]:
    return list(range(rank, n_frames, max(1, world)))


class _DeviceBuffer:
    """exposes a raw device pointer through __cuda_array_interface__ so torch can wrap it without a copy"""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def wrap_device_bytes(ptr: int, nbytes: int, device) -> torch.Tensor:
    return torch.as_tensor(_DeviceBuffer(ptr, nbytes), device=device)


class BandGather:
    """Gathers every rank's compact band buffer to rank 0 and un-interleaves the bands into the full frame.

    ``channels_bytes`` = bytes per pixel (3 for the quantised u8 frame, 12 for float rgb).  Buffers are padded to the
    largest per-rank row count so the collective moves equal-sized messages."""

    def __init__(self, width: int, height: int, band_rows: int, bytes_per_pixel: int, device, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.width, self.height, self.band_rows, self.bpp = width, height, band_rows, bytes_per_pixel
        self.device = torch.device(device)
        self.rows = [owned_rows(height, r, self.world, band_rows) for r in range(self.world)]
        self.max_rows = max(len(r) for r in self.rows) if self.rows else 0
        self.row_bytes = width * bytes_per_pixel
        self.send = torch.zeros(self.max_rows * self.row_bytes, dtype=torch.uint8, device=self.device)
        if self.rank == 0:
            self.recv = [torch.empty_like(self.send) for _ in range(self.world)]
            self.frame = torch.empty(height * self.row_bytes, dtype=torch.uint8, device=self.device)
            self.index = [torch.as_tensor(r, dtype=torch.long, device=self.device) for r in self.rows]
        else:
            self.recv, self.frame, self.index = None, None, None

    def gather(self, local: torch.Tensor) -> torch.Tensor | None:
        """local: uint8 tensor holding this rank's compact rows (len(rows[rank]) * row_bytes bytes).
        Returns the assembled frame (uint8, height * row_bytes) on rank 0, None elsewhere."""
        n = len(self.rows[self.rank]) * self.row_bytes
        self.send[:n].copy_(local.reshape(-1)[:n])
        if self.world > 1:
            dist.gather(self.send, self.recv, dst=0, group=self.group)
        elif self.rank == 0:
            self.recv[0].copy_(self.send)
        if self.rank != 0:
            return None
        frame = self.frame.view(self.height, self.row_bytes)
        for r in range(self.world):
            k = len(self.rows[r])
            if k:
                frame.index_copy_(0, self.index[r], self.recv[r][: k * self.row_bytes].view(k, self.row_bytes))
        return self.frame


class PeerFrame:
    """The exchange step without a gather: rank 0 owns the full quantised frame, every rank's push kernel stores its
    row bands into it through a CUDA-IPC peer mapping (NVLink), un-interleaving on the way (csrc/crt_push.cuh).
    torch.distributed only ships the 64-byte handle once and provides the completion barrier (a 1-element all-reduce
    on the render stream, so nothing blocks the host)."""

    def __init__(self, renderer, width: int, height: int, device, group=None):
        self.renderer, self.group = renderer, group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.width, self.height = width, height
        self.device = torch.device(device)
        handle = torch.zeros(64, dtype=torch.uint8, device=self.device)
        if self.rank == 0:
            raw = renderer.frame_create(width, height)
            handle.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
        if self.world > 1:
            dist.broadcast(handle, src=0, group=group)
            if self.rank != 0:
                renderer.frame_open(bytes(handle.cpu().numpy().tobytes()), width, height)
        self.flag = torch.zeros(1, dtype=torch.int32, device=self.device)

    def push(self):
        """enqueue this rank's bands -> rank 0's frame, then the completion barrier (asynchronous on the stream)"""
        self.renderer.push_bands()
        if self.world > 1:
            dist.all_reduce(self.flag, group=self.group)

    def frame(self) -> torch.Tensor | None:
        """the assembled frame (uint8 [H, W, 3]) on rank 0 once the stream has passed the barrier"""
        if self.rank != 0:
            return None
        ptr, w, h = self.renderer.frame_pointer()
        return wrap_device_bytes(ptr, w * h * 3, self.device).view(h, w, 3)
