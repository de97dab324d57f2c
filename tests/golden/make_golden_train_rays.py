"""Golden vectors of the TRAINING sampler, produced by the reference's own `build_rays` (build container only).

Runs utils/utils.py:400-597 verbatim (with_depth=True, as train.py:147 does) on a seeded synthetic scene and stores
the pixel selection it drew together with everything it computed from it: directions, gathered colours, the sorted +
jittered depth candidates, world points and the four NDC tensors.  The selection itself is random (torch.multinomial
/ randint / numpy), so parity is defined on "same pixels, same jitter -> same bits": the tests feed the stored
pixel_coordinates and t_rand into the oracle / CUDA path.  Two CPU-only shims, both name-level:
`Tensor.view` falls back to `reshape` for the non-contiguous patch slices, and torch.rand is captured.

Usage:  python tests/golden/make_golden_train_rays.py        (rewrites tests/golden/train_rays_*.npz)
"""
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

TRAIN_RAY_CASES = {
    # name: forward golden case reused for the scene, N_rays (batch), patch_num, patch_size, sparse-depth rays
    "train_rays_small_v3": ("small_v3", 40, 4, 2, 24),
    "train_rays_s126_v3": ("s126_v3", 64, 2, 4, 16),
}


def train_inputs(base_case, n_depth):
    """Seeded extra inputs of the training sampler: target image, sparse-depth coordinates (row, col)."""
    rng = np.random.default_rng(8000 + base_case["seed"])
    H, W = base_case["H"], base_case["W"]
    target = rng.random((3, H, W), dtype=np.float32)
    coords = np.stack((rng.integers(0, H, n_depth), rng.integers(0, W, n_depth)), 1).astype(np.int64)
    return target, coords


def run_case(U, name, base, N_rays, patch_num, patch_size, n_depth):
    from make_golden import CASES
    from ucnerf_b200 import synthetic
    c = dict(CASES[base])
    H, W, V, S, seed = c["H"], c["W"], c["V"], c["S"], c["seed"]
    sc = synthetic.make_scene(H, W, V, seed=seed, target_motion=c.get("target_motion", 0.0))
    T = torch.from_numpy
    target, coords = train_inputs(c, n_depth)
    imgs = torch.cat([T(target)[None, None], T(sc.imgs)], dim=1)                  # [1, V+1, 3, H, W], target first
    outputs = {}
    for k in range(3):
        D, h, w = sc.depth_values[k].shape[1:]
        outputs[f"stage{k + 1}"] = {"depth_values": T(sc.depth_values[k]), "prob_volume": torch.zeros((1, D, h, w))}
    pose_ref = {"w2cs": T(sc.w2cs).clone(), "intrinsics": T(sc.K).clone(), "c2ws": T(sc.c2ws).clone(),
                "near_fars": T(sc.near_fars)}
    args = SimpleNamespace(patch_num=patch_num, patch_size=patch_size)
    torch.manual_seed(seed)
    np.random.seed(seed)
    real_rand, real_view = torch.rand, torch.Tensor.view
    cap = {}

    def fake_rand(*size, **kw):
        r = real_rand(*size, **{k: v for k, v in kw.items() if k != "device"})
        if len(r.shape) == 2:
            cap["t_rand"] = r.clone()
        return r

    def view_or_reshape(self, *shape):
        try:
            return real_view(self, *shape)
        except RuntimeError:
            return self.reshape(*shape)

    torch.rand, torch.Tensor.view = fake_rand, view_or_reshape
    try:
        with torch.no_grad():
            pts, dirs, colors, ndc, z, rays_o, rays_depths, ndcp, pix = U.build_rays(
                args, imgs, T(sc.confidence), None, T(coords).float(), pose_ref, pose_ref["w2cs"], pose_ref["c2ws"],
                pose_ref["intrinsics"], N_rays, S, with_depth=True, outputs=outputs)
    finally:
        torch.rand, torch.Tensor.view = real_rand, real_view
    R = pix.shape[1]
    assert R == N_rays + n_depth and tuple(cap["t_rand"].shape) == (R, S)
    res = {"meta": np.array([H, W, V, S, N_rays, patch_num, patch_size, n_depth, seed], np.int64), "base": np.array(base),
           "pix": pix.numpy(), "t_rand": cap["t_rand"].numpy(), "z": z.numpy(), "rays_d": dirs.numpy(),
           "pts": pts.numpy(), "colors": colors.numpy(), "rays_o": rays_o.numpy()}
    for k in ("stage1", "stage2", "stage3", "ndc"):
        res[k] = ndc[k].numpy()
    print(f"{name}: R={R} (patches {patch_num}x{patch_size}^2, random {N_rays - patch_num * patch_size ** 2}, depth {n_depth}) "
          f"z[{res['z'].min():.3f},{res['z'].max():.3f}]")
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **res)


def main():
    from make_golden import import_reference
    torch.set_num_threads(2)
    U, M, Rn = import_reference()
    for name, cfg in TRAIN_RAY_CASES.items():
        run_case(U, name, *cfg)


if __name__ == "__main__":
    main()
