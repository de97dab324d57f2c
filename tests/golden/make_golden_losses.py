"""Golden values of the renderer-side loss terms, from the reference's own modules (build container only):
network/mvs_models.py:491 EdgePreservingSmoothnessLoss and utils/loss.py:95 GradientLoss on seeded patches, with the
gradients autograd gives for the depth patches.   Usage: python tests/golden/make_golden_losses.py"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)


def loss_inputs():
    rng = np.random.default_rng(4242)
    depth = (1.0 + rng.random((25, 4, 4))).astype(np.float32)
    guide = (1.0 + rng.random((25, 4, 4, 1))).astype(np.float32)
    depth[3] = 2.0                      # a constant patch: singular alignment system (det == 0) in GradientLoss
    return depth, guide


def main():
    from make_golden import import_reference
    import_reference()
    for name in ("lpips", "skimage"):
        sys.modules.setdefault(name, types.ModuleType(name))
    from network.mvs_models import EdgePreservingSmoothnessLoss
    from utils.loss import GradientLoss
    depth, guide = loss_inputs()
    d1 = torch.from_numpy(depth).clone().requires_grad_(True)
    s = EdgePreservingSmoothnessLoss()(d1, torch.from_numpy(guide))
    s.backward()
    d2 = torch.from_numpy(depth).clone().requires_grad_(True)
    g = GradientLoss()(d2, torch.from_numpy(guide).squeeze(-1), torch.ones_like(d2))
    g.backward()
    np.savez(os.path.join(HERE, "losses.npz"), smooth=s.item(), smooth_grad=d1.grad.numpy(), scale_inv=g.item(),
             scale_inv_grad=d2.grad.numpy())
    print("smooth", s.item(), "scale_inv", g.item())


if __name__ == "__main__":
    main()
