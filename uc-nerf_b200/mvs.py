"""Encoder-side hot op, the step before the render path: host-side mirror of

    homo_warp             utils/utils.py:1105-1172
    cost_volume_variance  the per-view loop of DepthNet.forward, network/mvs_models.py:603-628

over `ucn_homo_warp` / `ucn_cost_volume_variance` (include/ucnerf.h).  Batch size 1, like every call of the
reference (`train.py` feeds one scene at a time).  The regularisation 3-D CNN, softmax and depth regression that
follow (`mvs_models.py:629-646`) are cuDNN / elementwise torch code and are not re-implemented.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

from . import lib as L


def _proj_host(proj_mat: torch.Tensor) -> np.ndarray:
    p = proj_mat.detach().to("cpu", torch.float32).reshape(-1, 3, 4).contiguous().numpy()
    return np.ascontiguousarray(p, dtype=np.float32)


def homo_warp(src_feat: torch.Tensor, proj_mat: Optional[torch.Tensor], depth_values: Optional[torch.Tensor],
              src_grid: Optional[torch.Tensor] = None, pad: int = 0):
    """src_feat (1, C, H, W); proj_mat (1, 3, 4) = src_proj @ ref_proj_inv; depth_values (1, D, Hp, Wp) or (1, D)
    -> (warped (1, C, D, Hp, Wp), src_grid (1, D, Wp, Hp, 2)); the grid keeps the reference's (W, H)-swapped view
    shape (`utils/utils.py:1155`), whose memory order is [D][Hp][Wp][2]."""
    lib = L.load()
    B, Cn, H, W = src_feat.shape
    if B != 1:
        raise ValueError("homo_warp: batch size 1 only")
    feat = src_feat.detach().to(torch.float32).contiguous()
    Hp, Wp = H + 2 * pad, W + 2 * pad
    if src_grid is None:
        if depth_values.dim() != 4:                                     # (B, D) -> constant planes (:1120-1122)
            depth_values = depth_values[..., None, None].repeat(1, 1, Hp, Wp)
        dv = depth_values.detach().to(torch.float32).contiguous()
        D = dv.shape[1]
        grid_out = torch.empty((1, D, Wp, Hp, 2), device=feat.device, dtype=torch.float32)
        proj = _proj_host(proj_mat)
        args = (proj.ctypes.data_as(C.c_void_p), L.ptr(dv), None)
    else:
        D = src_grid.shape[1]
        grid_out = src_grid
        args = (None, None, L.ptr(src_grid.detach().to(torch.float32).contiguous()))
    warped = torch.empty((1, Cn, D, Hp, Wp), device=feat.device, dtype=torch.float32)
    L.check(lib.ucn_homo_warp(L.ptr(feat), Cn, H, W, args[0], args[1], args[2], D, pad, L.ptr(warped),
                              L.ptr(grid_out) if src_grid is None else None, L.stream_ptr()), "ucn_homo_warp")
    return warped, grid_out


def cost_volume_variance(features: Sequence[torch.Tensor], affine_mat_stage: torch.Tensor,
                         affine_mat_inv_stage: torch.Tensor, depth_values: torch.Tensor, pad: int = 0) -> torch.Tensor:
    """`volume_variance_no_ref` of DepthNet.forward: features = V tensors (1, C, H, W); affine_mat_stage (1 + V, 4, 4)
    with the target view first, affine_mat_inv_stage its inverses; depth_values (1, D, H, W) (padded here with
    `replicate` when pad > 0, `mvs_models.py:597-598`) -> (1, C, D, H + 2 pad, W + 2 pad).

    The reference's in-image mask stack has one slot more than there are source views and the unused slot stays 1
    (`:611`), so the mean divides by (views inside the image + 1): kept."""
    lib = L.load()
    feats = torch.stack([f.detach().to(torch.float32) for f in features])[:, 0].contiguous()      # [V][C][H][W]
    V, Cn, H, W = feats.shape
    if pad > 0:
        depth_values = F.pad(depth_values, (pad, pad, pad, pad), "replicate")
    dv = depth_values.detach().to(torch.float32).contiguous()
    D = dv.shape[1]
    # (affine_mats[i] @ affine_mat_inv)[:3] on the host in fp32, as torch's CPU matmul does it (:612)
    a = affine_mat_stage.detach().to("cpu", torch.float32)
    a_inv = affine_mat_inv_stage.detach().to("cpu", torch.float32)[0:1]
    projs = np.ascontiguousarray(torch.cat([(a[1 + i:2 + i] @ a_inv)[:, :3] for i in range(V)]).numpy(), dtype=np.float32)
    var = torch.empty((1, Cn, D, H + 2 * pad, W + 2 * pad), device=feats.device, dtype=torch.float32)
    L.check(lib.ucn_cost_volume_variance(L.ptr(feats), V, Cn, H, W, projs.ctypes.data_as(C.c_void_p), L.ptr(dv), D, pad, 1,
                                         L.ptr(var), L.stream_ptr()), "ucn_cost_volume_variance")
    return var
