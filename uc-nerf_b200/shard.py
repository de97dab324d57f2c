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
