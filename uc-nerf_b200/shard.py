"""Multi-GPU ray-tile sharding: one process per GPU, contiguous row tiles of the target image per rank,
one image gather at the end.  The reference has no multi-GPU code (train.py:433); rays are
independent (network/renderer.py:215-255 has no cross-ray term), so no data-path collective is needed
until the final gather of [R/P, 4] floats (rgb + depth, 16 B per ray).

Host logic only -- works with any torch.distributed backend (NCCL on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def row_tile(H: int, W: int, rank: int, world: int, row_align: int = 1) -> Tuple[int, int]:
    """Rows [r0, r1) of rank `rank`: as even as possible, tile heights multiples of `row_align`
    (except possibly the last).  Returns (ray0, n_rays) in the row-major pixel grid."""
    units = (H + row_align - 1) // row_align
    base, rem = divmod(units, world)
    u0 = rank * base + min(rank, rem)
    u1 = u0 + base + (1 if rank < rem else 0)
    r0, r1 = min(H, u0 * row_align), min(H, u1 * row_align)
    return r0 * W, (r1 - r0) * W


def gather_image(local: torch.Tensor, H: int, W: int, row_align: int = 1, group=None) -> torch.Tensor:
    """All-gather per-rank [n_rays_local, C] tiles into the full [H*W, C] image on every rank."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = [row_tile(H, W, p, world, row_align)[1] for p in range(world)]
    C_ = local.shape[1]
    if len(set(sizes)) == 1:
        out = torch.empty((H * W, C_), device=local.device, dtype=local.dtype)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    mx = max(sizes)
    pad = torch.zeros((mx, C_), device=local.device, dtype=local.dtype)
    pad[:local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:n] for p, n in zip(parts, sizes)], 0)


def render_sharded(render_rays: Callable[[int, int], Tuple[torch.Tensor, torch.Tensor]], H: int, W: int,
                   row_align: int = 1, group=None, gather: bool = True):
    """`render_rays(ray0, n)` -> (rgb [n,3], depth [n]) for this rank's tile; returns the full-frame
    (rgb [H*W,3], depth [H*W]) on every rank (or the local tile when gather=False)."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    ray0, n = row_tile(H, W, rank, world, row_align)
    rgb, depth = render_rays(ray0, n)
    if not gather or world == 1:
        return rgb, depth
    full = gather_image(torch.cat([rgb, depth[:, None]], 1), H, W, row_align, group)
    return full[:, :3].contiguous(), full[:, 3].contiguous()


class _DeviceBlob:
    """Raw device memory handed to torch through __cuda_array_interface__ (keeps `owner` alive)."""

    def __init__(self, addr: int, shape, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(addr), False),
                                         "version": 2, "strides": None}


class PeerFrameBuffer:
    """Frame buffer owned by the root rank and mapped into every other rank over NVLink (CUDA IPC peer memory).

    SURVEY.md 8e: the only exchange of a sharded render is 16 B per ray at the end; instead of an all-gather, every
    rank's compositing kernel (K4) stores its row tile straight into the root's buffer -- the "collective" is the
    epilogue's own stores over NVSwitch, plus one barrier.  One process per GPU, all GPUs visible to every process
    (torchrun's default).  Collective constructor: call it on every rank of `group`, with this rank's GPU current
    (ucn_peer_alloc / ucn_peer_open, include/ucnerf.h: the mapping is opened with the local device current, which is
    what lets that device's kernels address it)."""

    def __init__(self, H: int, W: int, device, root: int = 0, group=None):
        import ctypes as C
        from . import lib as L
        self.lib = L.load()
        self.root, self.group, self.H, self.W = root, group, H, W
        self.is_root = dist.get_rank(group) == root
        n = H * W
        self._base = C.c_void_p()
        handle = C.create_string_buffer(64)
        with torch.cuda.device(device):
            if self.is_root:
                L.check(self.lib.ucn_peer_alloc(16 * n, C.byref(self._base), handle), "ucn_peer_alloc")
            payload = [handle.raw if self.is_root else None]
            dist.broadcast_object_list(payload, src=root, group=group)
            if not self.is_root:
                L.check(self.lib.ucn_peer_open(payload[0], C.byref(self._base)), "ucn_peer_open")
            L._PEER_RANGES.append((self._base.value, self._base.value + 16 * n))
            # [n][3] rgb followed by [n] depth inside the one allocation
            # (no device= here: torch must alias the mapping, never copy it to the local GPU)
            self.rgb = torch.as_tensor(_DeviceBlob(self._base.value, (n, 3), self))
            self.depth = torch.as_tensor(_DeviceBlob(self._base.value + 12 * n, (n,), self))
            self.raw = torch.as_tensor(_DeviceBlob(self._base.value, (4 * n,), self))      # [3n | n], one copy for the read-out
            if self.rgb.data_ptr() != self._base.value or self.depth.data_ptr() != self._base.value + 12 * n:
                raise RuntimeError("PeerFrameBuffer: torch copied the peer mapping instead of aliasing it")
        dist.barrier(group=group)

    def close(self):
        """Collective: unmap on the peers, then free on the root."""
        if self._base:
            torch.cuda.synchronize()
            dist.barrier(group=self.group)
            self.rgb = self.depth = self.raw = None
            from . import lib as L
            rng = (self._base.value, self._base.value + 16 * self.H * self.W)
            if rng in L._PEER_RANGES:
                L._PEER_RANGES.remove(rng)
            if not self.is_root:
                self.lib.ucn_peer_close(self._base)
            dist.barrier(group=self.group)
            if self.is_root:
                self.lib.ucn_peer_free(self._base)
            self._base = None

    def tile(self, ray0: int, n: int):
        return self.rgb[ray0:ray0 + n], self.depth[ray0:ray0 + n]


def render_sharded_p2p(render_into: Callable[[int, int, torch.Tensor, torch.Tensor], None], buf: PeerFrameBuffer,
                       row_align: int = 1):
    """`render_into(ray0, n, out_rgb, out_depth)` renders this rank's row tile directly into the root's frame buffer
    (peer stores from K4); after the barrier the root holds the whole frame.  Returns (rgb, depth) on the root, None
    elsewhere."""
    rank, world = dist.get_rank(buf.group), dist.get_world_size(buf.group)
    ray0, n = row_tile(buf.H, buf.W, rank, world, row_align)
    out_rgb, out_depth = buf.tile(ray0, n)
    render_into(ray0, n, out_rgb, out_depth)
    torch.cuda.synchronize()
    dist.barrier(group=buf.group)
    return (buf.rgb, buf.depth) if buf.is_root else None
