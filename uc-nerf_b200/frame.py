"""Device-resident packed state for one target view, packed MLP weights, and the whole-frame entry
point ``render_frame`` that replaces the reference's per-chunk Python loop (train.py:251-275).

PyTorch is used for device memory and streams only; all arithmetic happens in libucnerf.so.
"""
from __future__ import annotations

import ctypes as C
import weakref
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import lib as L

STAGES = ("stage1", "stage2", "stage3")


_HOST_COPIES: Dict = {}


def _host_copy(x: torch.Tensor) -> np.ndarray:
    """Host copy of a small device tensor (camera matrices, near / far) that the kernels take BY VALUE.  The
    device->host read behind it synchronises the stream, and a training step that rebuilds its frame every step pays
    eight of them; the copy is therefore cached per tensor OBJECT: key = identity of the base tensor (weak reference,
    so a new tensor at a recycled address never hits) + view geometry, valid while `_version` is unchanged (every
    in-place write bumps it).  The cached array is read-only."""
    base = x._base if x._base is not None else x
    key = (id(base), x.storage_offset(), tuple(x.shape), tuple(x.stride()), x.dtype)
    hit = _HOST_COPIES.get(key)
    if hit is not None and hit[0]() is base and hit[1] == x._version:
        return hit[2]
    a = x.detach().to("cpu", torch.float32).contiguous().numpy()
    a.setflags(write=False)
    if len(_HOST_COPIES) >= 256:
        for k in [k for k, v in _HOST_COPIES.items() if v[0]() is None]:     # dead tensors first ...
            del _HOST_COPIES[k]
        while len(_HOST_COPIES) >= 256:                                      # ... then the oldest entries, one by one
            _HOST_COPIES.pop(next(iter(_HOST_COPIES)))
    _HOST_COPIES[key] = (weakref.ref(base), x._version, a)
    return a


def _host(x, shape=None) -> np.ndarray:
    if isinstance(x, torch.Tensor):
        x = _host_copy(x) if x.is_cuda else x.detach().to(torch.float32).numpy()
    a = np.ascontiguousarray(np.asarray(x, dtype=np.float32))
    return a if shape is None else a.reshape(shape)


def host_scalar(x) -> float:
    """float(x) without a stream synchronisation when x is a device tensor whose host copy is cached."""
    if isinstance(x, torch.Tensor) and x.is_cuda:
        return float(_host_copy(x).reshape(-1)[0])
    return float(x)


def _dev(x, device) -> torch.Tensor:
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    return x.detach().to(device=device, dtype=torch.float32).contiguous()


def fresh_seed() -> int:
    """A new 63-bit Philox key from torch's CPU generator (so torch.manual_seed controls the sequence): the default of
    every sampling entry point when the caller gives neither `t_rand` nor `seed`, mirroring the reference's fresh
    torch.rand per call (utils/utils.py:572,705)."""
    return int(torch.randint(0, 2 ** 62, (1,), device="cpu").item())


_T_VALS: Dict = {}


def linspace01(n: int, device) -> torch.Tensor:
    """t_vals exactly as the reference builds them: torch.linspace on the CPU, then .to(device)
    (utils/utils.py:394).  Cached per (n, device) and read-only: the pageable host->device copy behind .to()
    synchronises the stream, which would drain the GPU once per frame of a sweep."""
    device = torch.device(device)
    if device.type == "cuda" and device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    key = (int(n), str(device))
    t = _T_VALS.get(key)
    if t is None:
        t = _T_VALS[key] = torch.linspace(0., 1., steps=n).to(device)
    return t


STAGE_DOWN = (4, 2, 1)          # stage k maps live at H/4, H/2, H (network/mvs_models.py:649)


def stage_band(H: int, r0: int, r1: int, ds: int, halo: int = 2):
    """Rows [a, b) of a stage map (H/ds rows) that the rays of image rows [r0, r1) can touch.  The volumes live in the
    target's own frustum (train.py:235: world_to_ref is the target camera), so a sample of pixel row y projects back to
    row y: grid_sample(align_corners=False) reads stage rows floor(y*h/(H-1) - 0.5) and +1, i.e. within one row of
    y/ds; `halo` = 2 leaves a spare row on each side."""
    h = H // ds
    return max(0, r0 // ds - halo), min(h, -(-r1 // ds) + halo)


class PackedFrame:
    """Cameras (by value) + packed maps (device) for one target view == one ``ucn_frame``.

    Row-band mode (`set_row_band`, ray-tile sharding): only the rows of the target-frustum maps (depth hypotheses,
    volumes, confidence) that the rays of image rows [r0, r1) touch are packed -- from full maps or from maps that
    were already cut to the band (e.g. by a banded host->device copy, HostSweep).  The frame keeps the FULL map
    heights (border clamping is unchanged) and stores band pointers shifted back by the band's first row, so the
    kernels index absolute rows.  Source images / features are always whole: projections into the source views are
    not row-local."""

    def __init__(self, device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise L.UcnError("PackedFrame needs a CUDA device: there is no CPU path")
        L.load()
        self.c = L.Frame()
        self._keep: Dict[str, torch.Tensor] = {}
        self._t_vals: Optional[torch.Tensor] = None
        self.band = None                       # (r0, r1) image rows this frame can render, None = all
        self._halo = 2

    def set_row_band(self, r0: int, r1: int, halo: int = 2):
        """Restrict the frame to image rows [r0, r1) (call after set_target, before pack_*)."""
        c = self.c
        if not (0 <= r0 < r1 <= c.H):
            raise L.UcnError(f"row band [{r0}, {r1}) outside the {c.H}-row image")
        m = np.array(c.c2w[:], np.float64).reshape(4, 4) @ np.array(c.w2c[:], np.float64).reshape(4, 4)
        if np.abs(m - np.eye(4)).max() > 1e-4:
            raise L.UcnError("row bands need world_to_ref to be the target's own camera (inverse of tgt_to_world), as in "
                             "the reference's eval loop (train.py:235,247): the volumes are then pixel-aligned with the rays")
        self.band, self._halo = (int(r0), int(r1)), int(halo)
        return self

    def _rows(self, k_ds: int, full_h: int, src_h: int):
        """(source row offset, first band row a, band rows) for a stage map given with src_h rows."""
        if self.band is None:
            return 0, 0, src_h
        a, b = stage_band(self.c.H, self.band[0], self.band[1], k_ds, self._halo)
        if src_h == full_h:
            return a, a, b - a
        if src_h == b - a:
            return 0, a, b - a
        raise L.UcnError(f"stage map has {src_h} rows: expected the full {full_h} or the band's {b - a}")

    # ---- cameras ----------------------------------------------------------------------------
    def set_target(self, H, W, S, intrinsic, c2w, w2c, near, far, dir_rot0=None, dir_rot1=None, dir_switch=None):
        if S % 3 != 0 or S <= 0 or S > L.UCN_MAX_S:
            raise L.UcnError(f"N_samples={S}: must be a positive multiple of 3 (<= {L.UCN_MAX_S}); the reference "
                             "itself fails for other values (utils/utils.py:666,707-712)")
        c = self.c
        c.H, c.W, c.S = int(H), int(W), int(S)
        K = _host(intrinsic)
        if K.ndim == 3:                      # reference: intrinsic.mean(0) when batched (utils/utils.py:638)
            K = K.mean(0)
        c.K[:] = K.reshape(-1).tolist()
        c.c2w[:] = _host(c2w, (4, 4)).reshape(-1).tolist()
        w2c = _host(w2c, (4, 4))
        c.w2c[:] = w2c.reshape(-1).tolist()
        c.near, c.far = host_scalar(near), host_scalar(far)
        r0 = w2c[:3, :3] if dir_rot0 is None else _host(dir_rot0)[:3, :3]
        r1 = r0 if dir_rot1 is None else _host(dir_rot1)[:3, :3]
        c.dir_rot0[:] = np.ascontiguousarray(r0).reshape(-1).tolist()
        c.dir_rot1[:] = np.ascontiguousarray(r1).reshape(-1).tolist()
        c.dir_switch = int(2 ** 31 - 1 if dir_switch is None else dir_switch)
        self._t_vals = linspace01(S // 3, self.device)
        return self

    def set_sources(self, w2cs, intrinsics):
        w2cs, Ks = _host(w2cs), _host(intrinsics)
        V = w2cs.shape[0]
        if not 1 <= V <= L.UCN_MAX_V:
            raise L.UcnError(f"{V} source views: supported range is 1..{L.UCN_MAX_V}")
        self.c.V = V
        self.c.src_w2c[:V * 16] = w2cs.reshape(-1).tolist()
        self.c.src_K[:V * 9] = Ks.reshape(-1).tolist()
        return self

    # ---- maps ------------------------------------------------------------------------------
    def _alloc(self, name, *shape):
        t = torch.empty(shape, device=self.device, dtype=torch.float32)
        self._keep[name] = t
        return t

    def pack_depth_values(self, depth_values: Sequence):
        """outputs['stageK']['depth_values'] [1,D,h,w] -> [h,w,2] (first/last plane)."""
        lib, st = L.load(), L.stream_ptr()
        for k, dv in enumerate(depth_values):
            dv = _dev(dv, self.device)
            dv = dv[0] if dv.dim() == 4 else dv
            D, h, w = dv.shape
            full_h = h if self.band is None else self.c.H // STAGE_DOWN[k]
            src0, a, rows = self._rows(STAGE_DOWN[k], full_h, h)
            self.c.D[k], self.c.h[k], self.c.w[k] = D, full_h, w
            dst = self._alloc(f"nearfar{k}", rows, w, 2)
            L.check(lib.ucn_pack_nearfar_rows(L.ptr(dv.contiguous()), L.ptr(dst), D, h, w, src0, rows, st), "ucn_pack_nearfar")
            self.c.nearfar[k] = dst.data_ptr() - a * w * 2 * 4
        return self

    def pack_volumes(self, volumes: Sequence):
        """outputs['stageK']['volume_feature_no_ref'] [1,8,D,h,w] -> [h,w,D,8] (depth innermost)."""
        lib, st = L.load(), L.stream_ptr()
        for k, vol in enumerate(volumes):
            vol = _dev(vol, self.device)
            vol = vol[0] if vol.dim() == 5 else vol
            Cn, D, h, w = vol.shape
            if Cn != 8:
                raise L.UcnError("feature volumes must have 8 channels")
            full_h = h if self.band is None else self.c.H // STAGE_DOWN[k]
            if self.c.D[k] and (self.c.D[k], self.c.h[k], self.c.w[k]) != (D, full_h, w):
                raise L.UcnError("volume / depth_values shape mismatch")
            src0, a, rows = self._rows(STAGE_DOWN[k], full_h, h)
            self.c.D[k], self.c.h[k], self.c.w[k] = D, full_h, w
            dst = self._alloc(f"vol{k}", rows, w, D, 8)
            L.check(lib.ucn_pack_volume_rows(L.ptr(vol.contiguous()), L.ptr(dst), D, h, w, src0, rows, st), "ucn_pack_volume")
            self.c.vol[k] = dst.data_ptr() - a * w * D * 8 * 4
        return self

    def pack_sources(self, imgs, img_feat, confidence):
        """imgs [1,V,3,H,W] -> [V,H,W,4]; img_feat [V,1,8,H,W] -> [V,H,W,8]; confidence [1,H,W] as is."""
        lib, st = L.load(), L.stream_ptr()
        imgs = _dev(imgs, self.device)
        imgs = imgs[0] if imgs.dim() == 5 else imgs
        V, C3, H, W = imgs.shape
        feat = _dev(img_feat, self.device).reshape(V, 8, H, W)
        conf = _dev(confidence, self.device)
        conf = conf.reshape(conf.shape[-2], W)
        _, conf_a, _ = self._rows(1, H, conf.shape[0])          # whole map, or already cut to the band's rows
        if conf.shape[0] == H:
            conf_a = 0
        rgb = self._alloc("rgb", V, H, W, 4)
        ft = self._alloc("feat", V, H, W, 8)
        L.check(lib.ucn_pack_channels_last(L.ptr(imgs.contiguous()), L.ptr(rgb), V, C3, H * W, 4, st), "pack rgb")
        L.check(lib.ucn_pack_channels_last(L.ptr(feat), L.ptr(ft), V, 8, H * W, 8, st), "pack feat")
        self._keep["conf"] = conf
        self.c.rgb, self.c.feat, self.c.conf = rgb.data_ptr(), ft.data_ptr(), conf.data_ptr() - conf_a * W * 4
        return self

    @property
    def t_vals(self) -> torch.Tensor:
        return self._t_vals

    @property
    def F(self) -> int:
        return 24 + 12 * self.c.V + 1

    @property
    def map_bytes(self) -> int:
        return sum(t.numel() * 4 for t in self._keep.values())

    # ---- convenience ----------------------------------------------------------------------
    @classmethod
    def from_scene(cls, scene, S: int, device="cuda", dir_switch: Optional[int] = None, quirk_q4: bool = True):
        """Build from a synthetic.Scene (numpy, reference layouts)."""
        f = cls(device)
        f.set_target(scene.H, scene.W, S, scene.K[0], scene.c2ws[0], scene.w2cs[0], scene.near_fars[0, 0],
                     scene.near_fars[0, 1], dir_rot0=scene.w2cs[0],
                     dir_rot1=scene.w2cs[1] if quirk_q4 else scene.w2cs[0],
                     dir_switch=1024 if dir_switch is None else dir_switch)   # opt.py:65 default chunk
        f.set_sources(scene.w2cs[1:], scene.K[1:])
        f.pack_depth_values(scene.depth_values)
        if scene.volumes:
            f.pack_volumes(scene.volumes)
            f.pack_sources(scene.imgs, scene.img_feats, scene.confidence)
        return f


_SD_FIELDS = (("film", "pts_bias_depth_fine"), ("fb", "pts_bias_confidence"), ("feature", "feature_linear"),
              ("views", "views_linears.0"), ("vconf", "view_confi_linears.0"), ("rgb", "rgb_linear"),
              ("alpha", "alpha_linear"), ("crgb", "confi_rgb_linear"), ("alpha1", "alpha_linear_1"))


class PackedWeights:
    """fp32 + bf16 kernel packs of a UC-NeRF state_dict (reference key names, network/models.py:87-104)."""

    def __init__(self, state_dict: Dict, V: int, device="cuda"):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise L.UcnError("PackedWeights needs a CUDA device: there is no CPU path")
        lib = L.load()
        sd = {k: _dev(v, self.device) for k, v in state_dict.items()}
        exp = {"nerf.pts_linears.0.weight": (128, 63), "nerf.pts_linears.5.weight": (128, 191),
               "nerf.pts_bias_depth_fine.weight": (128, 24 + 4 * V), "nerf.pts_bias_confidence.weight": (128, 8 * V),
               "nerf.views_linears.0.weight": (64, 155)}
        for k, shp in exp.items():
            if tuple(sd[k].shape) != shp:
                raise L.UcnError(f"{k}: shape {tuple(sd[k].shape)} != {shp}; only D=6, W=128, multires=10/4, skips=[4] "
                                 "(the reference defaults, opt.py:40-44,90-92) are supported")
        c = L.StateDict()
        c.V = V
        for i in range(6):
            c.pts_w[i] = sd[f"nerf.pts_linears.{i}.weight"].data_ptr()
            c.pts_b[i] = sd[f"nerf.pts_linears.{i}.bias"].data_ptr()
        for field, name in _SD_FIELDS:
            setattr(c, field + "_w", sd[f"nerf.{name}.weight"].data_ptr())
            setattr(c, field + "_b", sd[f"nerf.{name}.bias"].data_ptr())
        self.f32 = torch.empty(lib.ucn_weights_f32_bytes(V) // 4, device=self.device, dtype=torch.float32)
        self.bf16 = torch.empty(lib.ucn_weights_bf16_bytes(V), device=self.device, dtype=torch.uint8)
        self.bf16x3 = torch.empty(lib.ucn_weights_bf16x3_bytes(V), device=self.device, dtype=torch.uint8)
        L.check(lib.ucn_pack_weights(C.byref(c), L.ptr(self.f32), L.ptr(self.bf16), L.ptr(self.bf16x3), L.stream_ptr()),
                "ucn_pack_weights")
        torch.cuda.current_stream().synchronize()      # sd temporaries may be freed after this
        self.c = L.Weights()
        self.c.V, self.c.f32, self.c.bf16, self.c.bf16x3 = V, self.f32.data_ptr(), self.bf16.data_ptr(), self.bf16x3.data_ptr()
        self.V = V


class FrameRenderer:
    """Whole-frame renderer: owns the workspace, runs K1->K2->K3->K4 on the current stream."""

    def __init__(self, frame: PackedFrame, weights: PackedWeights, mode: str = "bf16", max_rays: Optional[int] = None):
        self.frame, self.weights, self.mode = frame, weights, L.MODES[mode]
        self.lib = L.load()
        L.check(self.lib.ucn_check_device(), "ucn_check_device")
        self.max_rays = int(max_rays or frame.c.H * frame.c.W)
        self.ws_bytes = int(self.lib.ucn_frame_workspace_bytes(C.byref(frame.c), self.max_rays, self.mode))
        self.ws = torch.empty(self.ws_bytes, device=frame.device, dtype=torch.uint8)

    def render(self, ray0: int = 0, n_rays: Optional[int] = None, t_rand: Optional[torch.Tensor] = None,
               seed: Optional[int] = None, out_rgb: Optional[torch.Tensor] = None, out_depth: Optional[torch.Tensor] = None,
               timers: Optional[dict] = None):
        """`t_rand` [n, S] injects the jitter; otherwise the in-kernel Philox stream keyed by (`seed`, global ray id,
        sample) is used -- `seed=None` draws a fresh seed per call (frame.fresh_seed)."""
        f = self.frame
        if t_rand is None and seed is None:
            seed = fresh_seed()
        seed = int(seed or 0)
        total = f.c.H * f.c.W
        n = total - ray0 if n_rays is None else int(n_rays)
        if f.band is not None and n > 0 and not (f.band[0] * f.c.W <= ray0 and ray0 + n <= f.band[1] * f.c.W):
            raise L.UcnError(f"rays [{ray0}, {ray0 + n}) leave the frame's row band {f.band}")
        rgb = out_rgb if out_rgb is not None else torch.empty((n, 3), device=f.device, dtype=torch.float32)
        depth = out_depth if out_depth is not None else torch.empty((n,), device=f.device, dtype=torch.float32)
        for lo in range(0, n, self.max_rays):
            m = min(self.max_rays, n - lo)
            tr = None if t_rand is None else t_rand[lo:lo + m]
            if timers is None:
                L.check(self.lib.ucn_render_rays(C.byref(f.c), C.byref(self.weights.c), ray0 + lo, m, L.ptr(f.t_vals),
                                                 L.ptr(tr), int(seed), self.mode, L.ptr(self.ws), self.ws_bytes,
                                                 L.ptr(rgb[lo:lo + m]), L.ptr(depth[lo:lo + m]), L.stream_ptr()),
                        "ucn_render_rays")
            else:
                self._render_staged(ray0 + lo, m, tr, seed, rgb[lo:lo + m], depth[lo:lo + m], timers)
        return rgb, depth

    def _carve(self, m):
        """Workspace views with the same order ucn_render_rays uses (z, rays_d, angle, ndc, feat, raw)."""
        S, F = self.frame.c.S, self.frame.F
        out, o = [], 0
        n_tiles = (m * S + 127) // 128
        tile_floats = max(F * 128, int(self.lib.ucn_operand_tile_bytes(self.frame.c.V)) // 4)   # either interchange layout
        for shape in ((m, S), (m, 3), (m, 3), (m, S, 3), (n_tiles, tile_floats), (m, S, 4)):
            nbytes = 4 * int(np.prod(shape))
            out.append(self.ws[o:o + nbytes].view(torch.float32).view(*shape))
            o += (nbytes + 255) & ~255
        return out

    def _render_staged(self, ray0, m, t_rand, seed, rgb, depth, timers):
        """Same five launches as ucn_render_rays, issued one by one so that CUDA events on the launching
        stream can bracket each kernel (bench.py's roofline measurement)."""
        f, lib, st = self.frame, self.lib, L.stream_ptr()
        z, rays_d, angle, ndc, feat, raw = self._carve(m)

        def timed(name, fn):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            timers.setdefault(name, []).append((e0, e1))
        timed("raygen", lambda: L.check(lib.ucn_raygen_sample(C.byref(f.c), None, ray0, m, L.ptr(f.t_vals), L.ptr(t_rand),
                                                              int(seed), L.ptr(z), L.ptr(rays_d), None, None, st), "K1"))
        timed("dir", lambda: L.check(lib.ucn_dir_feature(C.byref(f.c), L.ptr(rays_d), m, ray0, L.ptr(angle), st), "dir"))
        tiled = self.mode == L.MLP_BF16X3_TC          # same choices as ucn_render_rays
        if tiled and lib.ucn_mlp_x3_variant(0) == 2 and f.c.V <= 7:
            gather, mlp = lib.ucn_gather_operands_from_z, lib.ucn_mlp_forward_operands
        elif tiled:
            gather, mlp = lib.ucn_gather_tiles_from_z, lib.ucn_mlp_forward_tiles
        else:
            gather, mlp = lib.ucn_gather_from_z, lib.ucn_mlp_forward
        timed("gather", lambda: L.check(gather(C.byref(f.c), L.ptr(z), None, ray0, m, L.ptr(feat), L.ptr(ndc), st), "K2"))
        timed("mlp", lambda: L.check(mlp(C.byref(self.weights.c), L.ptr(ndc), L.ptr(feat), L.ptr(angle),
                                         m, f.c.S, L.ptr(raw), self.mode, st), "K3"))
        timed("composite", lambda: L.check(lib.ucn_composite(L.ptr(raw), L.ptr(z), m, f.c.S, 0, L.ptr(rgb), L.ptr(depth),
                                                             None, None, st), "K4"))


def render_frame(frame: PackedFrame, weights: PackedWeights, mode: str = "bf16", t_rand=None, seed: Optional[int] = None,
                 ray0: int = 0, n_rays: Optional[int] = None, max_rays: Optional[int] = None):
    """One call per target view: (rgb [R,3], depth [R]) for rays ray0..ray0+R of the row-major pixel grid."""
    return FrameRenderer(frame, weights, mode, max_rays).render(ray0, n_rays, t_rand, seed)


class HostSweep:
    """Novel-view sweep from HOST maps (BASELINE config 3 shape: one set of encoder outputs per target view).

    ``render(i, host_maps, next_host_maps)`` renders view i from pinned host tensors and returns the image in a
    pinned host buffer; ``submit`` / ``result`` split the call so that a sweep can keep two views in flight.  The maps of view i+1 are copied host->device on a side stream while view i renders
    (two staging sets), so the PCIe transfer of every view is inside the sweep but off its critical path; the
    result is copied device->host and the call synchronises, as the reference's eval loop does per chunk
    (train.py:274-275).  `make_frame(dev_maps)` builds the PackedFrame (cameras + ucn_pack_* of the staged maps).

    Ray-tile sharding (one process per GPU): `ray0` / `n_rays` select this rank's row tile.  With
    `band_keys`, the host->device copy of those map groups (the target-frustum maps: depth hypotheses, volumes,
    confidence) carries only the rows the tile can touch (`stage_band`, one strided DMA per map: ucn_h2d_rows) --
    `make_frame` then receives band-shaped tensors and must call `PackedFrame.set_row_band` before packing.  With
    `peer` (two shard.PeerFrameBuffer, double buffered), K4 stores the tile straight into the root's frame buffer
    over NVLink; one 4-byte all-reduce per view is the barrier, after which the root copies the whole frame
    (16 B / ray) to its pinned host buffer.  `result(i)` returns that buffer on the root ([3n | n] floats, see
    `split_frame`) and None elsewhere."""

    def __init__(self, weights: PackedWeights, make_frame, n_rays: int, mode: str = "bf16", device="cuda",
                 max_rays: Optional[int] = None, ray0: int = 0, frame_hw=None, band_keys=(), peer=None, group=None):
        self.device = torch.device(device)
        self.weights, self.make_frame, self.mode, self.n_rays = weights, make_frame, mode, int(n_rays)
        self.ray0, self.peer, self.group = int(ray0), peer, group
        self.max_rays = int(max_rays or n_rays)         # rays per launch group (bounds the workspace)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.staging = [None, None]
        self.copied = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [torch.cuda.Event(), torch.cuda.Event()]
        self.pending = [None, None]                     # id of the host map set staged (or in flight) per slot
        self.renderer: Optional[FrameRenderer] = None
        self.band_keys, self.rows = tuple(band_keys), None
        if self.band_keys:
            H, W = frame_hw
            if self.ray0 % W or self.n_rays % W:
                raise L.UcnError("banded copies need a tile of whole image rows")
            self.frame_hw, self.rows = (int(H), int(W)), (self.ray0 // W, (self.ray0 + self.n_rays) // W)
        if peer is None:
            self.rgb = torch.empty((self.n_rays, 3), device=self.device)
            self.depth = torch.empty((self.n_rays,), device=self.device)
            self.out_dev = [torch.empty((self.n_rays, 4), device=self.device) for _ in range(2)]
            self.out_host = [torch.empty((self.n_rays, 4), dtype=torch.float32).pin_memory() for _ in range(2)]
        else:
            import torch.distributed as dist
            self._dist = dist
            self.is_root = peer[0].is_root
            self.flag = torch.zeros(1, device=self.device)
            total = peer[0].H * peer[0].W
            self.out_host = [torch.empty((4 * total,), dtype=torch.float32).pin_memory() for _ in range(2)] if self.is_root else None
        self.done = [torch.cuda.Event(), torch.cuda.Event()]
        self._frames = [None, None]
        for e in self.free:
            e.record(torch.cuda.current_stream(self.device))

    @staticmethod
    def _walk(maps):
        for k, v in maps.items():
            for j, t in enumerate(v if isinstance(v, (list, tuple)) else [v]):
                yield k, j, t

    def _band_of(self, t):
        """(first row, rows) of the band of host map `t` ([..., h, w], h = H/ds) for this rank's tile."""
        H = self.frame_hw[0]
        h = t.shape[-2]
        if H % h:
            raise L.UcnError(f"map with {h} rows is not a stage map of a {H}-row frame")
        a, b = stage_band(H, self.rows[0], self.rows[1], H // h)
        return a, b - a

    def _staging_like(self, k, t):
        if k in self.band_keys:
            shape = list(t.shape)
            shape[-2] = self._band_of(t)[1]
            return torch.empty(shape, device=self.device)
        return torch.empty(t.shape, device=self.device)

    def _enqueue(self, slot, host_maps, tag):
        if self.staging[slot] is None:
            self.staging[slot] = {k: ([self._staging_like(k, t) for t in v] if isinstance(v, (list, tuple))
                                      else self._staging_like(k, v)) for k, v in host_maps.items()}
            # the caching allocator may hand out blocks whose previous owner still has kernels in flight on the main
            # stream (safe for main-stream users only): the copy stream's first write must come after them
            self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))
        lib = L.load()
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.free[slot])            # the previous user of this staging set is done
            st = C.c_void_p(self.copy_stream.cuda_stream)
            for k, j, t in self._walk(host_maps):
                dst = self.staging[slot][k]
                dst = dst[j] if isinstance(dst, list) else dst
                if k in self.band_keys:
                    a, rows = self._band_of(t)
                    h, w = t.shape[-2:]
                    if t.is_cuda or not t.is_contiguous() or t.dtype != torch.float32:
                        raise L.UcnError("banded copies take contiguous fp32 HOST maps")
                    L.check(lib.ucn_h2d_rows(L.ptr(dst), C.c_void_p(t.data_ptr()), t.numel() // (h * w), h, w, a, rows, st),
                            "ucn_h2d_rows")
                else:
                    dst.copy_(t, non_blocking=True)
            self.copied[slot].record(self.copy_stream)
        self.pending[slot] = tag

    def h2d_bytes(self, host_maps) -> int:
        """Bytes one `_enqueue` moves over PCIe for `host_maps` (banded groups count their band only)."""
        n = 0
        for k, _, t in self._walk(host_maps):
            n += (t.numel() // t.shape[-2] * self._band_of(t)[1] if k in self.band_keys else t.numel()) * 4
        return n

    def submit(self, i: int, host_maps, next_host_maps=None, seed: Optional[int] = None):
        """Enqueue view i (host->device copy if not prefetched, K0 packing, K1..K4, device->host copy of the image) and
        return without synchronising; `result(i)` waits for it.  Submitting view i+1 before collecting view i keeps the
        GPU busy while the host builds the next frame (at most two views in flight: two staging sets, two images)."""
        slot = i & 1
        main = torch.cuda.current_stream(self.device)
        if self.pending[slot] != i:
            self._enqueue(slot, host_maps, i)
        if next_host_maps is not None:
            self._enqueue(slot ^ 1, next_host_maps, i + 1)          # overlaps this view's kernels
        main.wait_event(self.copied[slot])
        fr = self.make_frame(self.staging[slot])
        if self.renderer is None:
            self.renderer = FrameRenderer(fr, self.weights, self.mode, max_rays=self.max_rays)
        self.renderer.frame = fr
        self._frames[slot] = fr                                     # keeps the packed maps alive until the view is collected
        if self.peer is None:
            self.renderer.render(self.ray0, self.n_rays, seed=seed, out_rgb=self.rgb, out_depth=self.depth)
            self.free[slot].record(main)
            torch.cat((self.rgb, self.depth[:, None]), dim=1, out=self.out_dev[slot])
            self.out_host[slot].copy_(self.out_dev[slot], non_blocking=True)
        else:
            buf = self.peer[slot]
            out_rgb, out_depth = buf.tile(self.ray0, self.n_rays)   # K4's stores land in the root's HBM
            self.renderer.render(self.ray0, self.n_rays, seed=seed, out_rgb=out_rgb, out_depth=out_depth)
            self.free[slot].record(main)
            # stream-ordered barrier: completes once every rank's tile of view i is stored.  It also orders the root's
            # read-out of view i-1 (earlier in the root's stream) before any rank's stores of view i+1 into that buffer.
            self._dist.all_reduce(self.flag, group=self.group)
            if self.is_root:
                self.out_host[slot].copy_(buf.raw, non_blocking=True)
        self.done[slot].record(main)
        self.pending[slot] = None

    def result(self, i: int) -> torch.Tensor:
        """Pinned host image of view i (valid until view i+2 is submitted): [n_rays, 4] = (rgb, depth); in peer mode the
        root's whole frame as [3n | n] floats (`split_frame`), None on the other ranks."""
        self.done[i & 1].synchronize()
        if self.peer is not None and not self.is_root:
            return None
        return self.out_host[i & 1]

    @staticmethod
    def split_frame(flat: torch.Tensor):
        """[3n | n] -> (rgb [n,3], depth [n]) views of a peer-mode result."""
        n = flat.numel() // 4
        return flat[:3 * n].view(n, 3), flat[3 * n:]

    def render(self, i: int, host_maps, next_host_maps=None, seed: Optional[int] = None):
        """submit + result: one view, synchronous (the reference's eval loop syncs per chunk, train.py:274-275)."""
        self.submit(i, host_maps, next_host_maps, seed)
        return self.result(i)
