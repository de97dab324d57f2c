"""Golden cost volumes from the reference's own modules (build container only): utils/utils.py:1105 homo_warp and
the variance loop of network/mvs_models.py:589-628 (DepthNet.forward, run with a cost_regularization stub that
captures the variance volume).   Usage: python tests/golden/make_golden_costvol.py"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)


def costvol_inputs(seed, V=3, C=8, H=24, W=32, D=6, pad=0):
    """Random features, SCARED-like intrinsics and small relative poses; matrices [1 + V][4][4] (target first)."""
    rng = np.random.default_rng(seed)
    feats = rng.standard_normal((V, 1, C, H, W)).astype(np.float32)
    imgs = rng.random((1, V, 3, 4 * H, 4 * W)).astype(np.float32)
    K = np.array([[0.9 * W, 0, W / 2 - 0.3, 0], [0, 0.9 * W, H / 2 + 0.2, 0], [0, 0, 1, 0], [0, 0, 0, 1]], np.float32)
    mats, invs = [], []
    for i in range(V + 1):
        ang = 0.04 * rng.standard_normal(3)
        cx, cy, cz = np.cos(ang)
        sx, sy, sz = np.sin(ang)
        Rm = (np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]) @ np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
              @ np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]]))
        w2c = np.eye(4)
        w2c[:3, :3] = Rm
        w2c[:3, 3] = 0.08 * rng.standard_normal(3) * (i > 0)
        A = (K.astype(np.float64) @ w2c)
        mats.append(A.astype(np.float32))
        invs.append(np.linalg.inv(A).astype(np.float32))
    Hp, Wp = H + 2 * pad, W + 2 * pad
    depth = (0.6 + 1.5 * rng.random((1, D, H, W))).astype(np.float32)       # DepthNet pads it itself
    return feats, imgs, np.stack(mats), np.stack(invs), depth


CASES = {"cv_small": dict(seed=11, pad=0), "cv_pad": dict(seed=12, pad=2), "cv_v5": dict(seed=13, V=5, C=8, pad=0)}


def main():
    from make_golden import import_reference
    import_reference()
    for name in ("lpips", "skimage"):
        sys.modules.setdefault(name, types.ModuleType(name))
    from network.mvs_models import DepthNet
    from utils.utils import homo_warp
    for name, kw in CASES.items():
        feats, imgs, mats, invs, depth = costvol_inputs(**kw)
        pad = kw.get("pad", 0)
        V, D = feats.shape[0], depth.shape[1]
        cap = {}

        def cost_reg(v):
            cap["var"] = v.detach().clone()
            return v, torch.zeros((1, 1) + tuple(v.shape[2:]))

        with torch.no_grad():
            DepthNet()([torch.from_numpy(f) for f in feats], torch.from_numpy(mats), torch.from_numpy(invs),
                       torch.from_numpy(depth), D, cost_reg, torch.from_numpy(imgs), pad=pad)
            dv = torch.from_numpy(depth)
            if pad > 0:
                dv = torch.nn.functional.pad(dv, (pad, pad, pad, pad), "replicate")
            proj = (torch.from_numpy(mats[1:2]) @ torch.from_numpy(invs[0:1]))[:, :3]
            warped, grid = homo_warp(torch.from_numpy(feats[0]), proj, dv, pad=pad)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), var=cap["var"].numpy(), warped0=warped.numpy(),
                            grid0=grid.numpy(), proj0=proj.numpy())
        print(name, cap["var"].shape, float(cap["var"].abs().mean()), warped.shape, grid.shape)


if __name__ == "__main__":
    main()
