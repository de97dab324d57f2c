"""CPU: the renderer-side loss terms (ucnerf_b200/losses.py) against values and gradients produced by the reference's
own loss modules (tests/golden/losses.npz <- tests/golden/make_golden_losses.py)."""
import os

import numpy as np
import torch

from ucnerf_b200 import losses


def _inputs():
    rng = np.random.default_rng(4242)
    depth = (1.0 + rng.random((25, 4, 4))).astype(np.float32)
    guide = (1.0 + rng.random((25, 4, 4, 1))).astype(np.float32)
    depth[3] = 2.0
    return depth, guide


def test_patch_losses_match_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "losses.npz"))
    depth, guide = _inputs()
    d1 = torch.from_numpy(depth).clone().requires_grad_(True)
    s = losses.edge_preserving_smoothness(d1, torch.from_numpy(guide))
    s.backward()
    assert abs(s.item() - float(g["smooth"])) <= 1e-6
    assert np.abs(d1.grad.numpy() - g["smooth_grad"]).max() <= 1e-7
    d2 = torch.from_numpy(depth).clone().requires_grad_(True)
    v = losses.scale_invariant_gradient(d2, torch.from_numpy(guide).squeeze(-1))
    v.backward()
    assert abs(v.item() - float(g["scale_inv"])) <= 1e-5 * float(g["scale_inv"])
    assert np.abs(d2.grad.numpy() - g["scale_inv_grad"]).max() <= 1e-4 * np.abs(g["scale_inv_grad"]).max()


def test_render_losses_weights():
    """train.py:185-188: 0.05 * (nerf_depth + smooth) + 0.008 * scale_invariant + 5 * img (cas_mvsnet_loss is encoder-side)."""
    torch.manual_seed(0)
    H, W, N_rays, pn, ps, nd = 32, 40, 64, 2, 4, 16
    R = N_rays + nd
    rgb, tgt, depth = torch.rand(R, 3), torch.rand(R, 3), 1 + torch.rand(R)
    pix = torch.stack((torch.randint(0, H, (R,)), torch.randint(0, W, (R,))))
    sd, sw, dpt = 1 + torch.rand(1, H, W), torch.rand(1, H, W), 1 + torch.rand(1, H, W)
    loss, t = losses.render_losses(rgb, depth, tgt, pix, N_rays, pn, ps, sd, sw, dpt)
    want = 0.05 * t["nerf_depth"] + 0.05 * t["smooth"] + 0.008 * t["scale_invariant"] + 5.0 * t["img"]
    assert abs(loss.item() - want.item()) <= 1e-6
    assert abs(t["img"].item() - ((rgb - tgt) ** 2).mean().item()) <= 1e-7
    assert abs(losses.mse2psnr(torch.tensor(0.01)).item() - 20.0) <= 1e-5
