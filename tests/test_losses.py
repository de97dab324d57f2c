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


# ---- the analytic-gradient restatement (oracle/losses.py) and the fused CUDA kernel (ucn_render_losses) -------------
import pytest  # noqa: E402

from oracle import losses as OL  # noqa: E402


def test_oracle_loss_gradients_match_reference(golden_dir):
    """Closed-form gradients of the two patch terms == the reference modules' autograd gradients."""
    g = np.load(os.path.join(golden_dir, "losses.npz"))
    depth, guide = _inputs()
    v, gr = OL.smooth_term(depth, guide[..., 0])
    assert abs(v - float(g["smooth"])) <= 1e-7 and np.abs(gr - g["smooth_grad"]).max() <= 1e-8
    v, gr = OL.scale_invariant_term(depth, guide[..., 0])
    assert abs(v - float(g["scale_inv"])) <= 1e-6 * float(g["scale_inv"])
    assert np.abs(gr - g["scale_inv_grad"]).max() <= 1e-4 * np.abs(g["scale_inv_grad"]).max()


def _step_inputs(seed, H=32, W=40, N_rays=1024, pn=50, ps=4, nd=1024):
    rng = np.random.default_rng(seed)
    R = N_rays + nd
    rgb, tgt = rng.random((R, 3), dtype=np.float32), rng.random((R, 3), dtype=np.float32)
    depth = (1 + rng.random(R)).astype(np.float32)
    pix = np.stack((rng.integers(0, H, R), rng.integers(0, W, R)))
    sd, sw, dpt = (1 + rng.random((1, H, W))).astype(np.float32), rng.random((1, H, W)).astype(np.float32), \
        (1 + rng.random((1, H, W))).astype(np.float32)
    return rgb, depth, tgt, pix, N_rays, pn, ps, sd, sw, dpt


def test_oracle_render_losses_match_torch_mirror():
    """oracle.render_losses (closed forms, fp64) == losses.render_losses + autograd (pinned to the reference above)."""
    a = _step_inputs(5, N_rays=96, pn=5, nd=40)
    rgb, depth = torch.from_numpy(a[0]).requires_grad_(True), torch.from_numpy(a[1]).requires_grad_(True)
    loss, t = losses.render_losses(rgb, depth, torch.from_numpy(a[2]), torch.from_numpy(a[3]), *a[4:7],
                                   *(torch.from_numpy(x) for x in a[7:]))
    loss.backward()
    terms, d_rgb, d_depth = OL.render_losses(*a)
    assert abs(terms[0] - loss.item()) <= 1e-5 * abs(loss.item())
    assert np.abs(d_rgb - rgb.grad.numpy()).max() <= 1e-6
    assert np.abs(d_depth - depth.grad.numpy()).max() <= 1e-4 * np.abs(d_depth).max()


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", (dict(seed=1), dict(seed=2, N_rays=96, pn=5, nd=40), dict(seed=3, N_rays=200, pn=3, ps=8, nd=7)))
def test_cuda_fused_losses(cfg, golden_dir):
    """ucn_render_losses: value and gradient against the restatement (1e-5 / 1e-4 relative) on the config-4 batch
    (50 4x4 patches + 224 random + 1 024 sparse-depth rays), an odd patch count and 8x8 patches; and against the
    reference-generated goldens for the two patch terms."""
    a = _step_inputs(**cfg)
    T = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
    rgb, depth = T(a[0]).requires_grad_(True), T(a[1]).requires_grad_(True)
    loss, t = losses.render_losses_fused(rgb, depth, T(a[2]), T(a[3]), *a[4:7], *(T(x) for x in a[7:]))
    (2.0 * loss).backward()                                            # the upstream factor must scale the gradients
    terms, d_rgb, d_depth = OL.render_losses(*a)
    got = np.array([loss.item(), t["img"].item(), t["nerf_depth"].item(), t["smooth"].item(), t["scale_invariant"].item()])
    assert np.all(np.abs(got - terms) <= 1e-5 * np.abs(terms)), (got, terms)
    assert np.abs(rgb.grad.cpu().numpy() - 2 * d_rgb).max() <= 1e-4 * np.abs(2 * d_rgb).max()
    assert np.abs(depth.grad.cpu().numpy() - 2 * d_depth).max() <= 1e-4 * np.abs(2 * d_depth).max()
    # the golden patches of the reference's own modules: 25 smoothness + 25 scale-invariant patches on a [1][1][800] map
    g = np.load(os.path.join(golden_dir, "losses.npz"))
    pd, guide = _inputs()
    R = 800 + 8
    dep = np.concatenate([pd.reshape(-1), pd.reshape(-1), np.ones(8, np.float32)])
    dpt = np.concatenate([guide.reshape(-1), guide.reshape(-1)])[None, None].astype(np.float32)
    pix = np.stack((np.zeros(R, np.int64), np.concatenate([np.arange(800), np.zeros(8, np.int64)])))
    z = np.zeros((R, 3), np.float32)
    d = T(dep).requires_grad_(True)
    loss, t = losses.render_losses_fused(T(z), d, T(z), T(pix), 800, 50, 4, T(np.ones_like(dpt[0])), T(np.zeros_like(dpt[0])), T(dpt[0]))
    loss.backward()
    assert abs(t["smooth"].item() - float(g["smooth"])) <= 1e-6
    assert abs(t["scale_invariant"].item() - float(g["scale_inv"])) <= 1e-5 * float(g["scale_inv"])
    gd = d.grad.cpu().numpy()
    assert np.abs(gd[:400] - 0.05 * g["smooth_grad"].reshape(-1)).max() <= 1e-4 * 0.05 * np.abs(g["smooth_grad"]).max()
    assert np.abs(gd[400:800] - 0.008 * g["scale_inv_grad"].reshape(-1)).max() <= 1e-4 * 0.008 * np.abs(g["scale_inv_grad"]).max()
    with pytest.raises(Exception):
        losses.render_losses_fused(T(z).cpu(), d.cpu(), T(z).cpu(), T(pix).cpu(), 800, 50, 4, T(dpt[0]).cpu(), T(dpt[0]).cpu(), T(dpt[0]).cpu())
