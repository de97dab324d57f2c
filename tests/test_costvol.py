"""Variance cost volume / homo_warp.  CPU: the numpy restatement against volumes produced by the
reference's own DepthNet.forward / homo_warp (tests/golden/cv_*.npz).  GPU (-m gpu): the CUDA path through the C-ABI
against the same goldens and the restatement -- bit-exact (coordinates, nearest texels, masks and the variance are
all reproduced operation for operation)."""
import os
import sys

import numpy as np
import pytest

from oracle import costvol as OC

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden_costvol import CASES, costvol_inputs  # noqa: E402


def _case(name, golden_dir):
    import torch
    kw = CASES[name]
    feats, imgs, mats, invs, depth = costvol_inputs(**kw)
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    V = feats.shape[0]
    projs = np.stack([(torch.from_numpy(mats[i + 1:i + 2]) @ torch.from_numpy(invs[0:1]))[0, :3].numpy() for i in range(V)])
    return feats, mats, invs, depth, projs, kw.get("pad", 0), g


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_cost_volume_matches_reference(name, golden_dir):
    feats, mats, invs, depth, projs, pad, g = _case(name, golden_dir)
    dv = np.pad(depth[0], ((0, 0), (pad, pad), (pad, pad)), mode="edge")
    w, grid = OC.homo_warp(feats[0][0], projs[0], dv, pad)
    assert np.array_equal(bits(grid), bits(g["grid0"][0].reshape(grid.shape)))
    assert np.array_equal(bits(w), bits(g["warped0"][0]))
    assert np.array_equal(bits(OC.cost_volume_variance(feats[:, 0], projs, depth[0], pad)), bits(g["var"][0]))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_cost_volume_bit_exact(name, golden_dir):
    import torch
    from ucnerf_b200 import mvs
    feats, mats, invs, depth, projs, pad, g = _case(name, golden_dir)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dv = T(depth)
    dvp = torch.nn.functional.pad(dv, (pad, pad, pad, pad), "replicate") if pad else dv
    warped, grid = mvs.homo_warp(T(feats[0]), T(projs[0:1]), dvp, pad=pad)
    assert grid.shape == g["grid0"].shape
    assert np.array_equal(bits(grid.cpu().numpy()), bits(g["grid0"]))
    assert np.array_equal(bits(warped.cpu().numpy()), bits(g["warped0"]))
    again, _ = mvs.homo_warp(T(feats[0]), None, None, src_grid=grid, pad=pad)        # the `src_grid=` path
    assert torch.equal(again, warped)
    var = mvs.cost_volume_variance([T(f) for f in feats], T(mats), T(invs), dv, pad=pad)
    assert np.array_equal(bits(var.cpu().numpy()), bits(g["var"]))


@pytest.mark.gpu
def test_cuda_cost_volume_stage3_size_properties():
    """Stage-3 size of config 2 (8 channels, 8 hypotheses, 512x640, 3 views): identical source views at the identity
    warp have zero spread, so var = s2 * (1/4 - 9/16) with the reference's count of (3 inside + 1); a constant plane of
    depth through `(1, D)` depth_values gives the same volume as the expanded tensor; bad arguments fail loudly."""
    import torch
    from ucnerf_b200 import lib as L, mvs
    g = torch.Generator("cuda").manual_seed(9)
    C_, D, H, W, V = 8, 8, 512, 640, 3
    f = torch.randn((1, C_, H, W), device="cuda", generator=g)
    eye = torch.eye(4).repeat(V + 1, 1, 1)
    dv = torch.rand((1, D, H, W), device="cuda", generator=g) + 0.5
    var = mvs.cost_volume_variance([f] * V, eye, eye, dv)
    # interior voxels: all three views inside -> count 1/4: sq*c - (s*c)^2 = 3 f^2 / 4 - 9 f^2 / 16 = 3 f^2 / 16
    want = (3.0 / 16.0) * f[0, :, None, 1:-1, 1:-1] ** 2
    assert torch.allclose(var[0, :, :, 1:-1, 1:-1], want.expand(-1, D, -1, -1), rtol=1e-5, atol=1e-7)
    flat = torch.full((1, D), 2.0, device="cuda")
    w1, g1 = mvs.homo_warp(f, eye[:1, :3], flat)
    w2, g2 = mvs.homo_warp(f, eye[:1, :3], flat[..., None, None].repeat(1, 1, H, W))
    assert torch.equal(w1, w2) and torch.equal(g1, g2)
    assert torch.equal(w1[0, :, 0], f[0])                                # identity warp returns the source
    with pytest.raises(L.UcnError):
        mvs.cost_volume_variance([f] * 9, torch.eye(4).repeat(10, 1, 1), torch.eye(4).repeat(10, 1, 1), dv)
