"""Seeded synthetic scenes and random-init UC-NeRF weights (numpy only).

One generator feeds every consumer -- the golden-vector script (which hands the
arrays to the *reference* code), the CPU oracle, the CUDA path, bench.py and
smoke() -- so that all of them see bit-identical inputs on any machine.

Shapes follow what the reference's encoder emits and its renderer consumes
(reference: network/mvs_models.py:693-762, utils/utils.py:600-739,
data/scared.py:234-241,462-481):

  depth_values   stage k : [1, D_k, h_k, w_k]      (D,h,w) = (48,H/4,W/4) (32,H/2,W/2) (8,H,W)
  volume feature stage k : [1, 8, D_k, h_k, w_k]
  img_feats              : [V', 1, 8, H, W]
  imgs                   : [1, V', 3, H, W]
  confidence             : [1, H, W]
  poses                  : w2cs/c2ws [V,4,4], intrinsics [V,3,3], near_fars [V,2]  (index 0 = target)
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List

import numpy as np

STAGE_DEPTHS = (48, 32, 8)          # reference: network/mvs_models.py:649 (ndepths)
STAGE_DOWN = (4, 2, 1)              # stage k lives at H/4, H/2, H


@dataclass
class Scene:
    H: int
    W: int
    V: int                           # number of SOURCE views (view_num - 1)
    K: np.ndarray                    # [V+1,3,3]
    c2ws: np.ndarray                 # [V+1,4,4]
    w2cs: np.ndarray                 # [V+1,4,4]
    near_fars: np.ndarray            # [V+1,2]
    depth_values: List[np.ndarray]   # 3 x [1,D,h,w]
    volumes: List[np.ndarray]        # 3 x [1,8,D,h,w]
    img_feats: np.ndarray            # [V,1,8,H,W]
    imgs: np.ndarray                 # [1,V,3,H,W]
    confidence: np.ndarray           # [1,H,W]
    seed: int = 0
    meta: Dict = field(default_factory=dict)

    @property
    def view_num(self) -> int:
        return self.V + 1

    @property
    def feat_dim(self) -> int:       # reference: train.py:31
        return 24 + 12 * self.V + 1


def _bilinear_up(u: np.ndarray, H: int, W: int) -> np.ndarray:
    gh, gw = u.shape
    ys = np.linspace(0, gh - 1, H)
    xs = np.linspace(0, gw - 1, W)
    y0 = np.clip(np.floor(ys).astype(int), 0, gh - 2)
    x0 = np.clip(np.floor(xs).astype(int), 0, gw - 2)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = u[y0][:, x0]
    b = u[y0][:, x0 + 1]
    c = u[y0 + 1][:, x0]
    d = u[y0 + 1][:, x0 + 1]
    return (a * (1 - fy) * (1 - fx) + b * (1 - fy) * fx + c * fy * (1 - fx) + d * fy * fx)


def _roty(deg: float) -> np.ndarray:
    a = np.deg2rad(deg)
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


def make_poses(H: int, W: int, V: int, near: float = 1.2, far: float = 4.0, target_motion: float = 0.0):
    """Small-baseline endoscope-like rig: target at (almost) identity, sources translated/rotated."""
    f = 0.81 * W
    K = np.array([[f, 0, W / 2.0], [0, f, H / 2.0], [0, 0, 1]], dtype=np.float64)
    c2ws = []
    for i in range(V + 1):
        c2w = np.eye(4)
        if i == 0:
            c2w[:3, :3] = _roty(1.7 * target_motion)
            c2w[:3, 3] = [0.02 * target_motion, -0.01 * target_motion, 0.0]
        else:
            sgn = 1.0 if i % 2 else -1.0
            c2w[:3, :3] = _roty(sgn * min(3.0, 1.0 * i))
            c2w[:3, 3] = [sgn * 0.08 * i, 0.03 * i, 0.0]
        c2ws.append(c2w)
    c2ws = np.stack(c2ws)
    w2cs = np.linalg.inv(c2ws)
    Ks = np.repeat(K[None], V + 1, 0)
    nf = np.repeat(np.array([[near, far]]), V + 1, 0)
    return (Ks.astype(np.float32), c2ws.astype(np.float32), w2cs.astype(np.float32), nf.astype(np.float32))


def make_scene(H: int, W: int, V: int = 3, seed: int = 0, target_motion: float = 0.0,
               maps: bool = True) -> Scene:
    assert H % 4 == 0 and W % 4 == 0, "stage-1 maps live at H/4 x W/4"
    rng = np.random.default_rng(seed)
    near, far = 1.2, 4.0
    K, c2ws, w2cs, nf = make_poses(H, W, V, near, far, target_motion)
    di = (far - near) / 48.0
    u = _bilinear_up(rng.random((8, 10)), H, W)
    c = near + (far - near) * (0.3 + 0.4 * u)                        # [H,W] centre depth
    depth_values = []
    for k, (D, ds) in enumerate(zip(STAGE_DEPTHS, STAGE_DOWN)):
        h, w = H // ds, W // ds
        ck = c[::ds, ::ds][:h, :w]
        if k == 0:
            lo = np.full((h, w), near)
            hi = np.full((h, w), far)
        else:
            half = (32 if k == 1 else 4) * di
            lo = np.clip(ck - half, near * 0.5, None)
            hi = np.clip(ck + half, None, far * 1.5)
        t = np.linspace(0.0, 1.0, D)[:, None, None]
        depth_values.append(np.ascontiguousarray((lo[None] * (1 - t) + hi[None] * t)[None].astype(np.float32)))
    volumes, img_feats, imgs, conf = [], None, None, None
    if maps:
        for D, ds in zip(STAGE_DEPTHS, STAGE_DOWN):
            volumes.append(rng.standard_normal((1, 8, D, H // ds, W // ds), dtype=np.float32))
        img_feats = rng.standard_normal((V, 1, 8, H, W), dtype=np.float32)
        imgs = rng.random((1, V, 3, H, W), dtype=np.float32)
        conf = rng.random((1, H, W), dtype=np.float32)
    return Scene(H=H, W=W, V=V, K=K, c2ws=c2ws, w2cs=w2cs, near_fars=nf, depth_values=depth_values,
                 volumes=volumes, img_feats=img_feats, imgs=imgs, confidence=conf, seed=seed,
                 meta=dict(near=near, far=far))


# ----------------------------------------------------------------------------------------------
# Random-init weights with the reference's state_dict names/shapes (network/models.py:87-104).
# Values come from numpy (kaiming-normal-like W, small random b) so that the golden script can
# load the very same numbers into the reference's nn.Module.
# ----------------------------------------------------------------------------------------------

def ucnerf_param_shapes(V: int, D: int = 6, Wd: int = 128, in_pts: int = 63, in_views: int = 27,
                        skips=(4,)) -> Dict[str, tuple]:
    shp: Dict[str, tuple] = {}
    for i in range(D):
        if i == 0:
            k = in_pts
        elif (i - 1) in skips:
            k = Wd + in_pts
        else:
            k = Wd
        shp[f"nerf.pts_linears.{i}.weight"] = (Wd, k)
        shp[f"nerf.pts_linears.{i}.bias"] = (Wd,)
    def lin(name, o, i):
        shp[f"nerf.{name}.weight"] = (o, i)
        shp[f"nerf.{name}.bias"] = (o,)
    lin("pts_bias_depth_fine", Wd, 24 + 4 * V)
    lin("pts_bias_confidence", Wd, 8 * V)
    lin("pts_bias_confidence_1", 1, 1)                  # unused by forward (Q9), kept for state_dict parity
    lin("views_linears.0", Wd // 2, in_views + Wd)
    lin("view_confi_linears.0", Wd // 2, in_views + Wd)
    lin("feature_linear", Wd, Wd)
    lin("feature_linear_1", Wd, Wd)                     # unused
    lin("confi_linear", Wd, Wd)                         # unused
    lin("alpha_linear", 1, Wd // 2)
    lin("alpha_linear_1", 1, Wd)
    lin("rgb_linear", 3, Wd // 2)
    lin("confi_rgb_linear", 3, Wd)
    return shp


def make_weights(V: int, seed: int = 0, **kw) -> Dict[str, np.ndarray]:
    rng = np.random.default_rng(1000 + seed)
    out = {}
    for name, shape in ucnerf_param_shapes(V, **kw).items():
        if name.endswith("weight"):
            fan_in = shape[1]
            out[name] = (rng.standard_normal(shape) * np.sqrt(2.0 / fan_in)).astype(np.float32)
        else:
            out[name] = (rng.standard_normal(shape) * 0.05).astype(np.float32)
    return out
