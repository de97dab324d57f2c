"""Shared helpers for the parity tests: load a golden case and rebuild its seeded inputs."""
import os

import numpy as np

from ucnerf_b200 import synthetic

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ("small_v3", "small_v5", "s126_v3", "moved_v3")


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
