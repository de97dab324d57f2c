"""Drop-in mirrors of the reference's ray / sampling / gather helpers, backed by libucnerf.so.

Same names, argument order, return tuples and tensor shapes as the reference functions:

    utils/utils.py:217   get_rays_mvs            utils/utils.py:600   build_rays_test
    utils/utils.py:323   get_ndc_coordinate      utils/utils.py:393   sample_points_uniform
    utils/utils.py:742   build_color_volume      utils/utils.py:833   index_point_feature
    utils/utils.py:400   build_rays (training sampler; selection in PyTorch, sampling on K1)
    data/ray_utils.py:12 get_ray_directions      data/ray_utils.py:32 get_rays
    data/ray_utils.py:152 ray_marcher

Everything runs on the tensors' CUDA device; there is no CPU path (``use_cpu=True`` raises).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from . import lib as L
from . import train_path as _tp
import weakref

from .frame import PackedFrame, _host, fresh_seed, host_scalar, linspace01

# key -> (weak references to the base tensors of the key, PackedFrame).  A PackedFrame holds cameras by value and its own
# packed copies of the maps, so nothing keeps the caller's tensors alive: the key is therefore the IDENTITY of each
# tensor object (id of its base + view geometry + `_version`), and a hit additionally requires every weak reference to
# be alive and to be that very object -- a later frame whose fresh tensors land on recycled addresses (or recycled
# ids) can never hit an earlier frame's entry.
_FRAME_CACHE: Dict[tuple, tuple] = {}
_CACHE_MAX = 8


def _key(*tensors):
    key, bases = [], []
    for t in tensors:
        if isinstance(t, torch.Tensor):
            base = t._base if t._base is not None else t
            key.append((id(base), t._version, t.storage_offset(), tuple(t.shape), tuple(t.stride()), t.dtype))
            bases.append(base)
        else:
            key.append(t)
    return tuple(key), bases


def _cache_get(key, bases):
    hit = _FRAME_CACHE.get(key)
    if hit is None:
        return None
    refs, val = hit
    if len(refs) == len(bases) and all(r() is b for r, b in zip(refs, bases)):
        return val
    del _FRAME_CACHE[key]                # an id was recycled by a new tensor object: stale entry
    return None


def _cache_put(key, bases, val):
    for k in [k for k, (refs, _) in _FRAME_CACHE.items() if any(r() is None for r in refs)]:
        del _FRAME_CACHE[k]              # entries whose tensors are gone can never hit again
    if len(_FRAME_CACHE) >= _CACHE_MAX:
        _FRAME_CACHE.pop(next(iter(_FRAME_CACHE)))
    _FRAME_CACHE[key] = ([weakref.ref(b) for b in bases], val)
    return val


_INV_SCALE: Dict[tuple, torch.Tensor] = {}


def _inv_scale(W: int, H: int, device) -> torch.Tensor:
    """torch.tensor([W-1, H-1]).to(device) of the reference (utils/utils.py:585,726), cached per (W, H, device): the
    pageable host->device copy behind it synchronises the stream once per call otherwise.  Read-only by contract."""
    key = (int(W), int(H), str(device))
    t = _INV_SCALE.get(key)
    if t is None:
        t = _INV_SCALE[key] = torch.tensor([W - 1, H - 1]).to(device)
    return t


def _need_cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise L.UcnError(f"{what}: tensors must be on a CUDA device (no CPU fallback exists)")


def sample_points_uniform(device, near, far, num_samples, num_rays):
    """utils/utils.py:393-397 (elementwise torch ops; kept for API completeness, K1 fuses it)."""
    t_vals = linspace01(num_samples, device).view(1, num_samples).expand(num_rays, num_samples)
    return near * (1.0 - t_vals) + far * t_vals


def _rays_frame(H, W, S, tgt_to_world, world_to_ref, intrinsic, near, far, outputs) -> PackedFrame:
    dvs = [outputs[s]["depth_values"] for s in ("stage1", "stage2", "stage3")]
    tkey, bases = _key(tgt_to_world, world_to_ref, intrinsic, *dvs)
    key = ("rays", H, W, S, host_scalar(near), host_scalar(far)) + tkey
    fr = _cache_get(key, bases)
    if fr is None:
        fr = PackedFrame(tgt_to_world.device)
        fr.set_target(H, W, S, intrinsic, tgt_to_world, world_to_ref, near, far)
        fr.pack_depth_values(dvs)
        _cache_put(key, bases, fr)
    return fr


def build_rays_test(H, W, tgt_to_world, world_to_ref, intrinsic, near_fars_ref, near_fars, N_samples, pad=0,
                    ref_idx=0, use_cpu=False, chunk=-1, idx=-1, outputs=None, t_rand: Optional[torch.Tensor] = None,
                    seed: Optional[int] = None):
    """utils/utils.py:600-739.  Returns (ray_coordinate_world [R,S,3], ray_dir_world [R,3],
    ray_coordinate_ref {stage1,stage2,stage3,ndc: [R,S,3]}, depth_candidates [R,S], rays_os [R,3], ndc_parameters).

    ``t_rand`` ([R,S] uniforms) is the one extension: the reference draws ``torch.rand`` internally
    (:705); pass it to reproduce a given draw, omit it for the in-kernel Philox stream keyed by ``seed``.  With
    neither, a fresh 64-bit seed is drawn from torch's CPU generator on every call -- like the reference, two calls
    never reuse a jitter pattern, and ``torch.manual_seed`` makes the sequence reproducible."""
    if use_cpu:
        raise L.UcnError("build_rays_test(use_cpu=True): no CPU path in this implementation")
    _need_cuda(tgt_to_world, "build_rays_test")
    device = tgt_to_world.device
    S = int(N_samples)
    near, far = near_fars_ref[ref_idx, 0], near_fars_ref[ref_idx, 1]
    fr = _rays_frame(int(H), int(W), S, tgt_to_world, world_to_ref, intrinsic, host_scalar(near), host_scalar(far), outputs)
    total = H * W
    lo, hi = (0, total) if chunk <= 0 else (idx * chunk, min(total, (idx + 1) * chunk))
    R = hi - lo
    z = torch.empty((R, S), device=device)
    rays_d = torch.empty((R, 3), device=device)
    pts = torch.empty((R, S, 3), device=device)
    ndc = torch.empty((4, R, S, 3), device=device)
    if t_rand is not None:
        t_rand = t_rand.to(device=device, dtype=torch.float32).contiguous()
        assert tuple(t_rand.shape) == (R, S)
    elif seed is None:
        seed = fresh_seed()
    L.check(L.load().ucn_raygen_sample(C.byref(fr.c), None, lo, R, L.ptr(fr.t_vals), L.ptr(t_rand), int(seed or 0),
                                       L.ptr(z), L.ptr(rays_d), L.ptr(pts), L.ptr(ndc), L.stream_ptr()),
            "ucn_raygen_sample")
    rays_o = tgt_to_world[:3, -1].clone().reshape(1, 3).expand(R, -1)
    inv_scale = _inv_scale(W, H, device)
    ndc_parameters = {"w2c_ref": world_to_ref, "intrinsic_ref": intrinsic, "inv_scale": inv_scale, "near": near,
                      "far": far, "pad": pad}
    rays_ndc = {"stage1": ndc[0], "stage2": ndc[1], "stage3": ndc[2], "ndc": ndc[3]}
    return pts, rays_d, rays_ndc, z, rays_o, ndc_parameters


def get_rays_mvs(H, W, intrinsic, c2w, N=1024, isRandom=True, is_precrop_iters=False, chunk=-1, idx=-1,
                 with_mask=False, photo_confidence=None):
    """utils/utils.py:217-271, eval branch (isRandom=False): -> (rays_o [3], rays_d [R,3], pixel_coordinates [2,R])."""
    if isRandom:
        raise L.UcnError("get_rays_mvs(isRandom=True): ray *selection* stays in PyTorch; use get_rays_for_pixels")
    total = H * W
    lo, hi = (0, total) if chunk <= 0 else (idx * chunk, min(total, (idx + 1) * chunk))
    g = torch.arange(lo, hi, device=c2w.device)
    ys, xs = torch.div(g, W, rounding_mode="floor").float(), (g % W).float()
    return get_rays_for_pixels(H, W, intrinsic, c2w, ys, xs)


def get_rays_for_pixels(H, W, intrinsic, c2w, ys, xs):
    """Shared tail of get_rays_mvs / get_rays_mvs_coord (utils/utils.py:259-271, 304-320) for explicit pixels."""
    _need_cuda(c2w, "get_rays")
    device, R = c2w.device, ys.numel()
    fr = PackedFrame(device)
    eye = torch.eye(4)
    fr.set_target(H, W, 3, intrinsic, c2w, eye, 0.0, 1.0)
    # K1 needs near/far maps; a 1x1 dummy per stage keeps the kernel shape-agnostic for this ray-only use
    dummy = [torch.zeros((1, 2, max(H // d, 1), max(W // d, 1)), device=device) for d in (4, 2, 1)]
    fr.pack_depth_values(dummy)
    pix = torch.stack((ys.float(), xs.float())).contiguous()
    z = torch.empty((R, 3), device=device)
    rays_d = torch.empty((R, 3), device=device)
    tr = torch.zeros((R, 3), device=device)
    L.check(L.load().ucn_raygen_sample(C.byref(fr.c), L.ptr(pix), 0, R, L.ptr(fr.t_vals), L.ptr(tr), 0, L.ptr(z),
                                       L.ptr(rays_d), None, None, L.stream_ptr()), "ucn_raygen_sample")
    return c2w[:3, -1].clone(), rays_d, pix


def get_ray_directions(H, W, focal, center=None):
    """data/ray_utils.py:12-29: camera-frame directions [H,W,3] (x right, y down, z forward)."""
    device = focal.device if isinstance(focal, torch.Tensor) else "cuda"
    cx, cy = (W / 2, H / 2) if center is None else center
    K = torch.tensor([[float(focal[0]) if hasattr(focal, "__len__") else float(focal), 0, cx],
                      [0, float(focal[1]) if hasattr(focal, "__len__") else float(focal), cy], [0, 0, 1]])
    _, d, _ = get_rays_mvs(H, W, K, torch.eye(4, device=device), isRandom=False)
    return d.reshape(H, W, 3)


def get_rays(directions, c2w):
    """data/ray_utils.py:32-53: rotate camera-frame directions into the world; -> (rays_o [N,3], rays_d [N,3])."""
    rays_d = directions.reshape(-1, 3) @ c2w[:3, :3].T
    rays_o = c2w[:3, 3].expand(rays_d.shape)
    return rays_o, rays_d


def ray_marcher(rays, N_samples=64, lindisp=False, perturb=0, bbox_3D=None, t_rand: Optional[torch.Tensor] = None):
    """data/ray_utils.py:152-197.  rays [N,8] = (o, d, near, far) -> (xyz [N,S,3], rays_o, rays_d, z_vals [N,S]).
    ``t_rand`` ([N,S] uniforms) replaces the reference's internal ``torch.rand`` when perturb > 0."""
    _need_cuda(rays, "ray_marcher")
    rays = rays.to(torch.float32).contiguous()
    R, S = rays.shape[0], int(N_samples)
    xyz = torch.empty((R, S, 3), device=rays.device)
    z = torch.empty((R, S), device=rays.device)
    if perturb > 0 and t_rand is None:
        t_rand = torch.rand((R, S), device=rays.device)
    if t_rand is not None:
        t_rand = t_rand.to(device=rays.device, dtype=torch.float32).contiguous()
    L.check(L.load().ucn_ray_marcher(L.ptr(rays), R, S, L.ptr(linspace01(S, rays.device)), L.ptr(t_rand), float(perturb),
                                     int(bool(lindisp)), L.ptr(xyz), L.ptr(z), L.stream_ptr()), "ucn_ray_marcher")
    return xyz, rays[:, 0:3], rays[:, 3:6], z


def get_ndc_coordinate(w2c_ref, intrinsic_ref, point_samples, inv_scale, near_far=None, sample_2d=False):
    """utils/utils.py:323-373.  point_samples [R,S,3] -> [R,S,3] (sample_2d) or the 4-entry dict."""
    _need_cuda(point_samples, "get_ndc_coordinate")
    R, S = point_samples.shape[:2]
    pts = point_samples.to(torch.float32).contiguous()
    out = torch.empty((R, S, 3), device=pts.device)
    w2c = (C.c_float * 16)(*_host(torch.as_tensor(w2c_ref)).reshape(-1).tolist())
    K = (C.c_float * 9)(*_host(torch.as_tensor(intrinsic_ref)).reshape(-1).tolist())
    Wm1, Hm1 = (int(v) for v in torch.as_tensor(inv_scale).reshape(-1).tolist())
    L.check(L.load().ucn_project_points(w2c, K, Hm1 + 1, Wm1 + 1, L.ptr(pts), R * S, L.ptr(out), L.stream_ptr()),
            "ucn_project_points")
    if sample_2d:
        return out
    xy, pz = out[..., :2], out[..., 2:]
    res = {}
    for name, nk, fk in (("stage1", "near_1", "far_1"), ("stage2", "near_2", "far_2"), ("stage3", "near_3", "far_3"),
                         ("ndc", "near", "far")):
        res[name] = torch.cat([xy, (pz - near_far[nk]) / (near_far[fk] - near_far[nk])], dim=-1)
    return res


def _maps_frame(volume_feature, imgs, img_feat, confidence, pose_ref) -> PackedFrame:
    """Packed maps for (volumes, images, features, confidence, source cameras); any group may be None."""
    vols = None if volume_feature is None else [volume_feature[s]["volume_feature_no_ref"]
                                                for s in ("stage1", "stage2", "stage3")]
    cams = () if pose_ref is None else (pose_ref["w2cs"], pose_ref["intrinsics"])
    tkey, bases = _key(imgs, img_feat, confidence, *cams, *(vols or ()))
    key = ("maps",) + tkey
    fr = _cache_get(key, bases)
    if fr is None:
        ref = imgs if imgs is not None else confidence
        fr = PackedFrame(ref.device)
        H, W = ref.shape[-2:]
        fr.c.H, fr.c.W, fr.c.S = int(H), int(W), 3
        if vols is not None:
            fr.pack_volumes(vols)
        if imgs is not None:
            fr.set_sources(pose_ref["w2cs"], pose_ref["intrinsics"])
            if img_feat is None:
                img_feat = torch.zeros((imgs.shape[1], 1, 8, H, W), device=imgs.device)
            if confidence is None:
                confidence = torch.zeros((1, H, W), device=imgs.device)
            fr.pack_sources(imgs, img_feat, confidence)
        else:
            fr.c.V = 1
            conf = confidence.to(torch.float32).reshape(H, W).contiguous()
            fr._keep["conf"] = conf
            fr.c.conf = conf.data_ptr()
        _cache_put(key, bases, fr)
    return fr


def gather_features(volume_feature, imgs, img_feat, confidence, pose_ref, rays_pts, rays_ndc, with_idx=False,
                    parts=L.PART_ALL):
    """Fused index_point_feature + build_color_volume (+ assembly, renderer.py:191-207) -> input_feat [R,S,F]."""
    _need_cuda(rays_pts if rays_pts is not None else rays_ndc["stage1"], "gen_pts_feats")
    fr = _maps_frame(volume_feature, imgs, img_feat, confidence, pose_ref)
    if rays_pts is None:
        rays_pts = rays_ndc["stage1"]
    if rays_ndc is None:
        rays_ndc = {k: rays_pts for k in ("stage1", "stage2", "stage3")}
    R, S = rays_pts.shape[:2]
    if parts == L.PART_ALL and not with_idx and volume_feature is not None and imgs is not None:
        vols = [volume_feature[s]["volume_feature_no_ref"] for s in ("stage1", "stage2", "stage3")]
        if _tp.wants_grad(*vols, img_feat, confidence):
            # training step: K2 with its backward K2b (gradients for the volumes, FPN features, confidence)
            return _tp.GatherFn.apply(fr, rays_pts, rays_ndc["stage1"], rays_ndc["stage2"], rays_ndc["stage3"], *vols,
                                      img_feat, confidence)
    feat = torch.empty((R, S, fr.F), device=rays_pts.device)
    idx = torch.empty((R, S, 11 + 3 * fr.c.V), device=rays_pts.device, dtype=torch.int32) if with_idx else None
    args = [rays_pts, rays_ndc["stage1"], rays_ndc["stage2"], rays_ndc["stage3"]]
    args = [a.to(torch.float32).contiguous() for a in args]
    L.check(L.load().ucn_gather(C.byref(fr.c), *[L.ptr(a) for a in args], R * S, parts, L.ptr(feat), L.ptr(idx),
                                L.stream_ptr()), "ucn_gather")
    return (feat, idx) if with_idx else feat


def index_point_feature(volume_feature, volume_confidence, ray_coordinate_ref, chunk=-1):
    """utils/utils.py:833-893 -> (out_features [R,S,24], confidence_sampled [R,S])."""
    feat = gather_features(volume_feature, None, None, volume_confidence, None, None, ray_coordinate_ref,
                           parts=L.PART_VOLUMES)
    return feat[..., :24], feat[..., -1]


def build_color_volume(point_samples, pose_ref, imgs, near_far=None, img_feat=None, confidence=None, downscale=1.0,
                       with_mask=False):
    """utils/utils.py:742-799 -> colors [R,S,4V (+8V when img_feat is given)]."""
    V = imgs.shape[1]
    feat = gather_features(None, imgs, img_feat, None, pose_ref, point_samples, None, parts=L.PART_VIEWS)
    return feat[..., 24:24 + (12 if img_feat is not None else 4) * V]


# ---- training sampler -------------------------------------------------------------------

_SELECT_WS: Dict[tuple, torch.Tensor] = {}


def select_training_pixels(H, W, confidence, coords, N_rays, patch_num, patch_size, generator=None,
                           is_precrop_iters=False, seed: Optional[int] = None):
    """Ray SELECTION of build_rays (utils/utils.py:450-479,169-215,217-247,274-320): patch_num/2 patches drawn by
    confidence-multinomial, patch_num/2 by (1 - confidence), uniform random pixels up to N_rays, then the sparse-depth
    coordinates.  -> pixel_coordinates [2, R] (row, col), float32.

    One launch (`ucn_select_training_pixels`: exponential-race sampling without replacement -- what torch.multinomial
    runs itself -- patch placement, uniform pixels and the coordinates; keyed by `seed`, default a fresh one per call).
    With a `generator` the selection is drawn by torch instead (same distribution, ~45 small launches): the
    reference's per-patch Python loop as one vectorised index computation, reproducible with a torch.Generator."""
    device = confidence.device
    ps, half = int(patch_size), int(patch_num) // 2
    if N_rays < 2 * half * ps * ps:
        raise L.UcnError("batch_size must be >= patch_num * patch_size^2 (utils/utils.py:468-474, quirk Q20)")
    if generator is None and confidence.is_cuda and half <= 64 and ((H * W + 2047) // 2048) * half <= 2048:
        lib = L.load()
        conf = confidence.detach().reshape(-1).to(torch.float32).contiguous()
        n_rand = int(N_rays) - 2 * half * ps * ps
        crd = None if coords is None else coords.to(device=device, dtype=torch.int64).contiguous()
        n_depth = 0 if crd is None else int(crd.shape[0])
        key = (str(device), int(lib.ucn_select_workspace_bytes(H, W, int(patch_num))))
        ws = _SELECT_WS.get(key)
        if ws is None:
            ws = _SELECT_WS[key] = torch.zeros(key[1], dtype=torch.uint8, device=device)
        pix = torch.empty((2, 2 * half * ps * ps + n_rand + n_depth), device=device, dtype=torch.float32)
        crop = int(bool(is_precrop_iters) and float(torch.rand((1,))) > 0.3)            # utils/utils.py:241
        L.check(lib.ucn_select_training_pixels(L.ptr(conf), int(H), int(W), n_rand, int(patch_num), ps, crop, L.ptr(crd),
                                               n_depth, int(fresh_seed() if seed is None else seed), L.ptr(ws), L.ptr(pix),
                                               L.stream_ptr()), "ucn_select_training_pixels")
        return pix
    conf = confidence.detach().reshape(-1).float()
    grid = torch.arange(ps, device=device)
    pix = []
    for weights in (conf, 1.0 - conf):
        sel = torch.multinomial(weights, half, generator=generator)
        pi = torch.clamp(torch.div(sel, W, rounding_mode="floor") // ps, 0, H // ps - 2)
        pj = torch.clamp((sel % W) // ps, 0, W // ps - 2)
        si = torch.randint(0, ps, (half,), device=device, generator=generator)
        sj = torch.randint(0, ps, (half,), device=device, generator=generator)
        rows = (pi * ps + si)[:, None, None] + grid[None, :, None]            # [half, ps, 1]
        cols = (pj * ps + sj)[:, None, None] + grid[None, None, :]            # [half, 1, ps]
        pix.append(torch.stack((rows.expand(-1, ps, ps).reshape(-1), cols.expand(-1, ps, ps).reshape(-1))))
    n_rand = N_rays - 2 * half * ps * ps
    if is_precrop_iters and float(torch.rand((1,))) > 0.3:                   # centre crop, utils/utils.py:241-246
        xs = torch.randint(W // 6, W - W // 6, (n_rand,), device=device, generator=generator)
        ys = torch.randint(H // 6, H - H // 6, (n_rand,), device=device, generator=generator)
    else:
        xs = torch.randint(0, W, (n_rand,), device=device, generator=generator)
        ys = torch.randint(0, H, (n_rand,), device=device, generator=generator)
    pix.append(torch.stack((ys, xs)))
    if coords is not None:
        pix.append(coords.to(device).t().long())                              # coords [N,2] = (row, col)
    return torch.cat(pix, dim=1).float()


def build_rays(args, imgs, mvs_confidence, sparse_depths, coords, pose_ref, w2cs, c2ws, intrinsics, N_rays, N_samples,
               pad=0, is_precrop_iters=False, ref_idx=0, with_depth=False, outputs=None,
               pixel_coordinates: Optional[torch.Tensor] = None, t_rand: Optional[torch.Tensor] = None,
               seed: Optional[int] = None):
    """utils/utils.py:400-597 (training sampler).  Returns the reference's 9-tuple
    (ray_coordinate_world [R,S,3], ray_dir_world [R,3], colors [R,3], points_all {stage1,stage2,stage3,ndc},
    depth_candidates [R,S], rays_os [R,3], rays_depths (None), ndc_parameters, pixel_coordinates_int [2,R]).

    Selection stays in PyTorch (`select_training_pixels`); everything after it -- ray directions, per-pixel stage
    near/far, 3 x S/3 sweeps, sort, jitter, world points, the four NDC tensors -- is K1's explicit-pixel mode, the
    same arithmetic as `build_rays_test`.  Extensions: `pixel_coordinates` injects a selection (parity tests),
    `t_rand` injects the jitter the reference draws with torch.rand (:572); without `t_rand` the in-kernel Philox stream
    is keyed by `seed`, and with neither a fresh seed is drawn per call from torch's CPU generator (every training step
    resamples its depths, as the reference does)."""
    if not with_depth:
        raise L.UcnError("build_rays(with_depth=False) is dead in the reference (depth_candidate undefined, "
                         "utils/utils.py:575); train.py:147 always passes with_depth=True")
    _need_cuda(imgs, "build_rays")
    device = imgs.device
    N, V, C3, H, W = imgs.shape
    S = int(N_samples)
    if pixel_coordinates is None:
        pixel_coordinates = select_training_pixels(H, W, mvs_confidence, coords, N_rays, args.patch_num, args.patch_size,
                                                   is_precrop_iters=is_precrop_iters)
    pix = pixel_coordinates.to(device=device, dtype=torch.float32).contiguous()
    R = pix.shape[1]
    w2c_ref, intrinsic_ref = pose_ref['w2cs'][ref_idx], pose_ref['intrinsics'][ref_idx]
    near, far = pose_ref['near_fars'][ref_idx, 0], pose_ref['near_fars'][ref_idx, 1]
    fr = _rays_frame(int(H), int(W), S, c2ws[0], w2c_ref, intrinsics[0], host_scalar(near), host_scalar(far), outputs)
    z = torch.empty((R, S), device=device)
    rays_d = torch.empty((R, 3), device=device)
    pts = torch.empty((R, S, 3), device=device)
    ndc = torch.empty((4, R, S, 3), device=device)
    if t_rand is not None:
        t_rand = t_rand.to(device=device, dtype=torch.float32).contiguous()
        assert tuple(t_rand.shape) == (R, S)
    elif seed is None:
        seed = fresh_seed()
    L.check(L.load().ucn_raygen_sample(C.byref(fr.c), L.ptr(pix), 0, R, L.ptr(fr.t_vals), L.ptr(t_rand), int(seed or 0),
                                       L.ptr(z), L.ptr(rays_d), L.ptr(pts), L.ptr(ndc), L.stream_ptr()),
            "ucn_raygen_sample")
    pix_int = pix.long()
    colors = imgs[0, 0][:, pix_int[0], pix_int[1]].permute(1, 0)              # :497-501, target view = index 0
    rays_o = c2ws[0][:3, -1].clone().reshape(1, 3).expand(R, -1)
    ndc_parameters = {"w2c_ref": w2c_ref, "intrinsic_ref": intrinsic_ref,
                      "inv_scale": _inv_scale(W, H, device), "near": near, "far": far, "pad": pad}
    points_all = {"stage1": ndc[0], "stage2": ndc[1], "stage3": ndc[2], "ndc": ndc[3]}
    return pts, rays_d, colors, points_all, z, rays_o, None, ndc_parameters, pix_int
