"""Shared helpers for the parity tests: load a golden case and rebuild its seeded inputs."""
import os

import numpy as np

from ucnerf_b200 import synthetic

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ("small_v3", "small_v5", "s126_v3", "moved_v3", "s255_v5")


def load_case(name):
    g = dict(np.load(os.path.join(GOLDEN_DIR, f"{name}.npz")))
    H, W, V, S, chunk, seed, stride = (int(x) for x in g["meta"])
    tm = float(g["target_motion"])
    scene = synthetic.make_scene(H, W, V, seed=seed, target_motion=tm)
    weights = synthetic.make_weights(V, seed=seed)
    t_rand = np.random.default_rng(7000 + seed).random((H * W, S), dtype=np.float32)
    return dict(g=g, H=H, W=W, V=V, S=S, chunk=chunk, seed=seed, stride=stride, scene=scene, weights=weights,
                t_rand=t_rand)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


GRAD_CASES = ("grad_small_v3", "grad_small_v5")


def load_grad_case(name):
    """Golden gradients of the reference's autograd (tests/golden/make_golden_grad.py) + the forward case they extend."""
    g = dict(np.load(os.path.join(GOLDEN_DIR, f"{name}.npz")))
    c = load_case(str(g["base"]))
    R = int(g["meta"][4])
    rng = np.random.default_rng(9000 + c["seed"])
    g_rgb = rng.standard_normal((R, 3)).astype(np.float32)
    g_depth = rng.standard_normal((R,)).astype(np.float32)
    c.update(gg=g, R=R, g_rgb=g_rgb, g_depth=g_depth)
    return c


def rel_err(a, b):
    """max |a-b| relative to the largest magnitude of the expected tensor (gradients span decades)."""
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(np.abs(b).max(), 1e-30))


TRAIN_RAY_CASES = ("train_rays_small_v3", "train_rays_s126_v3")


def load_train_rays_case(name):
    """Golden output of the reference's training sampler build_rays (tests/golden/make_golden_train_rays.py)."""
    g = dict(np.load(os.path.join(GOLDEN_DIR, f"{name}.npz")))
    c = load_case(str(g["base"]))
    H, W, V, S, N_rays, patch_num, patch_size, n_depth, seed = (int(x) for x in g["meta"])
    rng = np.random.default_rng(8000 + seed)
    target = rng.random((3, H, W), dtype=np.float32)
    coords = np.stack((rng.integers(0, H, n_depth), rng.integers(0, W, n_depth)), 1).astype(np.int64)
    c.update(tg=g, N_rays=N_rays, patch_num=patch_num, patch_size=patch_size, n_depth=n_depth, target=target, coords=coords)
    return c
