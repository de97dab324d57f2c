"""ctypes front-end of the CPU oracle (oracle/libucnerf_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs -- never by the product package.  Function names mirror the reference functions they restate
(utils/utils.py: build_rays_test, index_point_feature, build_color_volume; network/renderer.py:
gen_pts_feats, run_network_mvs, raw2outputs, rendering).  Inputs are synthetic.Scene objects and
numpy arrays in the reference's own layouts.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libucnerf_oracle.so")
_lib = None

_fp = C.POINTER(C.c_float)


class _Scene(C.Structure):
    _fields_ = [("H", C.c_int32), ("W", C.c_int32), ("V", C.c_int32), ("S", C.c_int32),
                ("D", C.c_int32 * 3), ("h", C.c_int32 * 3), ("w", C.c_int32 * 3),
                ("K", C.c_float * 9), ("c2w", C.c_float * 16), ("w2c", C.c_float * 16),
                ("near", C.c_float), ("far", C.c_float),
                ("dir_rot0", C.c_float * 9), ("dir_rot1", C.c_float * 9),
                ("dir_switch", C.c_int32), ("_pad", C.c_int32),
                ("depth", _fp * 3), ("vol", _fp * 3), ("conf", _fp), ("imgs", _fp), ("feats", _fp),
                ("src_w2c", _fp), ("src_K", _fp)]


class _Weights(C.Structure):
    _fields_ = [("V", C.c_int32), ("in_pts", C.c_int32), ("in_views", C.c_int32), ("width", C.c_int32),
                ("depth", C.c_int32), ("skip", C.c_int32),
                ("pts_w", _fp * 8), ("pts_b", _fp * 8)] + [
        (n + s, _fp) for n in ("film", "fb", "feature", "views", "vconf", "rgb", "alpha", "crgb", "alpha1")
        for s in ("_w", "_b")]


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc only; no reference sources involved)."""
    srcs = [os.path.join(_HERE, f) for f in ("ucnerf_oracle.c", "ucnerf_oracle_mlp.c", "ucnerf_oracle.h", "Makefile")]
    stale = os.path.exists(_LIB_PATH) and any(os.path.getmtime(f) > os.path.getmtime(_LIB_PATH) for f in srcs)
    if force or stale or not os.path.exists(_LIB_PATH):
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.uco_num_threads.restype = C.c_int
    return _lib


def _p(a: Optional[np.ndarray]):
    if a is None:
        return C.cast(None, _fp)
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"], (a.dtype, a.flags)
    return a.ctypes.data_as(_fp)


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def linspace01(n: int) -> np.ndarray:
    """torch.linspace(0., 1., n) on CPU, restated (ATen RangeFactories: symmetric two-sided fill;
    reference call site utils/utils.py:394).  Checked against torch.linspace in tests/test_oracle.py."""
    if n == 1:
        return np.zeros(1, np.float32)
    step = np.float64(np.float32(1.0) / np.float32(n - 1))       # fp32 step, widened
    i = np.arange(n)
    # ATen's vectorised CPU fill evaluates start + step*i / end - step*(n-1-i) with ONE rounding
    # (FMA); the products are exact in float64, so rounding the float64 sum once reproduces it.
    lo = (step * i).astype(np.float32)
    hi = (1.0 - step * (n - 1 - i)).astype(np.float32)
    return np.where(i < n // 2, lo, hi).astype(np.float32)


class OracleScene:
    """Holds numpy arrays alive and exposes the C struct."""

    def __init__(self, scene, S: int, dir_switch: int = 0, quirk_q4: bool = True):
        assert S % 3 == 0, "N_samples must be a multiple of 3 (reference crashes otherwise, Q1)"
        self.scene, self.S = scene, S
        s = _Scene()
        s.H, s.W, s.V, s.S = scene.H, scene.W, scene.V, S
        self._keep = []
        for k in range(3):
            dv = _f32(scene.depth_values[k][0])
            s.D[k], s.h[k], s.w[k] = dv.shape
            s.depth[k] = _p(dv)
            self._keep.append(dv)
            if scene.volumes:
                vol = _f32(scene.volumes[k][0])
                assert vol.shape == (8,) + dv.shape
                s.vol[k] = _p(vol)
                self._keep.append(vol)
        s.K[:] = scene.K[0].reshape(-1).tolist()
        s.c2w[:] = scene.c2ws[0].reshape(-1).tolist()
        s.w2c[:] = scene.w2cs[0].reshape(-1).tolist()
        s.near, s.far = float(scene.near_fars[0, 0]), float(scene.near_fars[0, 1])
        # Q4: first `dir_switch` rays rotate into the target frame, the rest into source view 1's.
        s.dir_rot0[:] = scene.w2cs[0][:3, :3].reshape(-1).tolist()
        s.dir_rot1[:] = (scene.w2cs[1] if quirk_q4 else scene.w2cs[0])[:3, :3].reshape(-1).tolist()
        s.dir_switch = dir_switch
        if scene.imgs is not None:
            for name, arr in (("conf", scene.confidence[0]), ("imgs", scene.imgs[0]),
                              ("feats", scene.img_feats[:, 0]), ("src_w2c", scene.w2cs[1:]),
                              ("src_K", scene.K[1:])):
                a = _f32(arr)
                setattr(s, name, _p(a))
                self._keep.append(a)
        self.c = s

    @property
    def F(self):
        return 24 + 12 * self.scene.V + 1


class OracleWeights:
    def __init__(self, sd: Dict[str, np.ndarray], V: int):
        w = _Weights()
        w.V, w.in_pts, w.in_views, w.width, w.depth, w.skip = V, 63, 27, 128, 6, 4
        self._keep = []

        def g(name):
            a = _f32(sd[name])
            self._keep.append(a)
            return _p(a)
        for i in range(6):
            w.pts_w[i], w.pts_b[i] = g(f"nerf.pts_linears.{i}.weight"), g(f"nerf.pts_linears.{i}.bias")
        for field, name in (("film", "pts_bias_depth_fine"), ("fb", "pts_bias_confidence"),
                            ("feature", "feature_linear"), ("views", "views_linears.0"),
                            ("vconf", "view_confi_linears.0"), ("rgb", "rgb_linear"),
                            ("alpha", "alpha_linear"), ("crgb", "confi_rgb_linear"),
                            ("alpha1", "alpha_linear_1")):
            setattr(w, field + "_w", g(f"nerf.{name}.weight"))
            setattr(w, field + "_b", g(f"nerf.{name}.bias"))
        assert sd["nerf.pts_linears.0.weight"].shape == (128, 63)
        assert sd["nerf.pts_linears.5.weight"].shape == (128, 191)
        self.c = w


def chunk_pixels(H: int, W: int, chunk: int = -1, idx: int = -1):
    """Row-major meshgrid slice of get_rays_mvs (utils/utils.py:249-257) as float (ys, xs)."""
    n = H * W
    lo, hi = (0, n) if chunk <= 0 else (idx * chunk, min(n, (idx + 1) * chunk))
    g = np.arange(lo, hi)
    return (g // W).astype(np.float32), (g % W).astype(np.float32)


def build_rays_test(osc: OracleScene, pix_y, pix_x, t_rand, with_pts=True, with_ndc=True):
    """utils/utils.py:600-739. Returns dict(z, rays_d, pts, ndc{stage1,stage2,stage3,ndc})."""
    R, S = len(pix_y), osc.S
    pix_y, pix_x, t_rand = _f32(pix_y), _f32(pix_x), _f32(t_rand).reshape(R, S)
    t_vals = linspace01(S // 3)
    z = np.empty((R, S), np.float32)
    rd = np.empty((R, 3), np.float32)
    pts = np.empty((R, S, 3), np.float32) if with_pts else None
    ndc = np.empty((4, R, S, 3), np.float32) if with_ndc else None
    lib().uco_raygen_sample(C.byref(osc.c), _p(pix_y), _p(pix_x), R, _p(t_vals), _p(t_rand), _p(z), _p(rd),
                            _p(pts), _p(ndc))
    out = dict(z=z, rays_d=rd, pts=pts)
    if with_ndc:
        out["ndc_all"] = ndc
        out["ndc"] = {k: ndc[i] for i, k in enumerate(("stage1", "stage2", "stage3", "ndc"))}
    return out


def gen_pts_feats(osc: OracleScene, pts, ndc_all, with_idx=False):
    """network/renderer.py:177-212 (+ utils/utils.py:742-799, 833-893). -> feat [R,S,F] (, idx)."""
    R, S = pts.shape[:2]
    feat = np.empty((R, S, osc.F), np.float32)
    idx = np.empty((R, S, 11 + 3 * osc.scene.V), np.int32) if with_idx else None
    lib().uco_gather(C.byref(osc.c), _p(_f32(pts)), _p(_f32(ndc_all)), R, _p(feat),
                     idx.ctypes.data_as(C.POINTER(C.c_int32)) if with_idx else None)
    return (feat, idx) if with_idx else feat


def dir_feature(osc: OracleScene, rays_d, ray_offset=0):
    R = rays_d.shape[0]
    ang = np.empty((R, 3), np.float32)
    lib().uco_dir_feature(C.byref(osc.c), _p(_f32(rays_d)), R, ray_offset, _p(ang))
    return ang


def embed(x, L):
    x = _f32(x).reshape(-1, 3)
    out = np.empty((x.shape[0], 3 + 6 * L), np.float32)
    for i in range(x.shape[0]):
        lib().uco_embed(_p(x[i]), L, _p(out[i]))
    return out


def run_network(ow: OracleWeights, ndc_pts, feat, angle):
    """run_network_mvs + UCNeRF.forward (network/renderer.py:78-106, network/models.py:138-184)."""
    R, S = feat.shape[:2]
    raw = np.empty((R, S, 4), np.float32)
    lib().uco_mlp(C.byref(ow.c), _p(_f32(ndc_pts)), _p(_f32(feat)), _p(_f32(angle)), R, S, _p(raw))
    return raw


def raw2outputs(raw, z):
    """network/renderer.py:109-140 -> (rgb_map [R,3], depth_map [R])."""
    R, S = z.shape
    rgb = np.empty((R, 3), np.float32)
    depth = np.empty((R,), np.float32)
    lib().uco_composite(_p(_f32(raw)), _p(_f32(z)), R, S, _p(rgb), _p(depth))
    return rgb, depth


def raw2outputs_bwd(raw, z, g_rgb, g_depth):
    """Backward of raw2outputs (network/renderer.py:25-36,125-131) -> d_raw [R,S,4]."""
    R, S = z.shape
    d_raw = np.empty((R, S, 4), np.float32)
    lib().uco_composite_bwd(_p(_f32(raw)), _p(_f32(z)), R, S, _p(_f32(g_rgb)), _p(_f32(g_depth)), _p(d_raw))
    return d_raw


def gen_pts_feats_bwd(osc: OracleScene, pts, ndc_all, d_feat):
    """Backward of gen_pts_feats w.r.t. the sampled tensors -> (d_vol[3] [8,D,h,w], d_img_feats [V,8,H,W], d_conf [H,W])."""
    R = pts.shape[0]
    c = osc.c
    dv = [np.zeros((8, c.D[k], c.h[k], c.w[k]), np.float32) for k in range(3)]
    dfe = np.zeros((c.V, 8, c.H, c.W), np.float32)
    dco = np.zeros((c.H, c.W), np.float32)
    lib().uco_gather_bwd(C.byref(c), _p(_f32(pts)), _p(_f32(ndc_all)), R, _p(_f32(d_feat)), _p(dv[0]), _p(dv[1]),
                         _p(dv[2]), _p(dfe), _p(dco))
    return dv, dfe, dco


def render_rays(osc: OracleScene, ow: OracleWeights, ray0: int, R: int, t_rand):
    """Whole path for rays [ray0, ray0+R) (chained C driver; used as CPU baseline)."""
    t_rand = _f32(t_rand).reshape(R, osc.S)
    t_vals = linspace01(osc.S // 3)
    rgb = np.empty((R, 3), np.float32)
    depth = np.empty((R,), np.float32)
    lib().uco_render_rays(C.byref(osc.c), C.byref(ow.c), ray0, R, _p(t_vals), _p(t_rand), _p(rgb), _p(depth))
    return rgb, depth


def ray_marcher(rays: np.ndarray, S: int, lindisp: bool = False, perturb: float = 0.0, rnd: Optional[np.ndarray] = None):
    """data/ray_utils.py:152-197 -> (xyz [R,S,3], z [R,S])."""
    rays = _f32(rays)
    R = rays.shape[0]
    xyz, z = np.empty((R, S, 3), np.float32), np.empty((R, S), np.float32)
    rnd = _f32(rnd) if rnd is not None else None
    lib().uco_ray_marcher(_p(rays), R, S, _p(linspace01(S)), _p(rnd), C.c_float(perturb), int(lindisp), _p(xyz), _p(z))
    return xyz, z


def num_threads() -> int:
    return int(lib().uco_num_threads())


def set_num_threads(n: int) -> int:
    """OpenMP threads of the oracle (torchrun exports OMP_NUM_THREADS=1; the timed CPU legs want every core)."""
    lib().uco_set_num_threads(int(n))
    return num_threads()
