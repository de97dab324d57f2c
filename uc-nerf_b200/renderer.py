"""Drop-in mirror of the reference's per-ray renderer (network/renderer.py), backed by libucnerf.so.

    network/renderer.py:25   raw2alpha        network/renderer.py:163  gen_dir_feature
    network/renderer.py:57   batchify         network/renderer.py:177  gen_pts_feats
    network/renderer.py:78   run_network_mvs  network/renderer.py:215  rendering
    network/renderer.py:109  raw2outputs      network/renderer.py:6    depth2dist

Same names, argument order and return values.  Reference quirks are reproduced on purpose (SURVEY.md
Q2-Q4, Q19): `dists` is ignored by the compositing, activations live in the MLP, and ``rendering``
strips the target pose from ``pose_ref`` IN PLACE on its first call.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as L
from . import rays as _rays
from . import train_path as _tp
from .frame import PackedFrame, _host


def depth2dist(z_vals, cos_angle):
    """network/renderer.py:6-15 (its result is unused by the reference's compositing, Q3)."""
    dists = z_vals[..., 1:] - z_vals[..., :-1]
    dists = torch.cat([dists, torch.full_like(dists[..., :1], 1e10)], -1)
    return dists * cos_angle.unsqueeze(-1)


def _composite(raw, z_vals, white_bkgd, want_weights):
    if not raw.is_cuda:
        raise L.UcnError("raw2outputs: tensors must be on a CUDA device (no CPU fallback exists)")
    R, S = z_vals.shape
    if not want_weights and _tp.wants_grad(raw):      # training step: K4 with its backward K4b (train.py:147-190)
        rgb, depth = _tp.CompositeFn.apply(raw, z_vals, white_bkgd)
        return rgb, depth, None, None
    raw4 = raw[..., :4].to(torch.float32).contiguous()
    z = z_vals.to(torch.float32).contiguous()
    rgb = torch.empty((R, 3), device=raw.device)
    depth = torch.empty((R,), device=raw.device)
    acc = torch.empty((R,), device=raw.device)
    weights = torch.empty((R, S), device=raw.device) if want_weights else None
    L.check(L.load().ucn_composite(L.ptr(raw4), L.ptr(z), R, S, int(bool(white_bkgd)), L.ptr(rgb), L.ptr(depth),
                                   L.ptr(acc), L.ptr(weights), L.stream_ptr()), "ucn_composite")
    return rgb, depth, acc, weights


def raw2alpha(sigma, dist, net_type):
    """network/renderer.py:25-36 -> (alpha, weights, alpha_softmax); `dist` is ignored, as in the reference."""
    R, S = sigma.shape
    raw = torch.zeros((R, S, 4), device=sigma.device)
    raw[..., 3] = sigma
    _, _, _, weights = _composite(raw, torch.zeros_like(sigma), False, True)
    return 1. - torch.exp(-sigma), weights, torch.softmax(sigma, 1)


def raw2outputs(raw, z_vals, dists, white_bkgd=False, net_type='v2'):
    """network/renderer.py:109-140 -> (rgb_map, disp_map, acc_map, weights, depth_map, weights, var)."""
    rgb_map, depth_map, acc_map, weights = _composite(raw, z_vals, white_bkgd, True)
    var = torch.var(weights, dim=1)
    disp_map = 1. / torch.max(1e-10 * torch.ones_like(depth_map), depth_map / acc_map)
    return rgb_map, disp_map, acc_map, weights, depth_map, weights, var


def gen_dir_feature(w2c_ref, rays_dir):
    """network/renderer.py:163-174: rays_dir @ w2c_ref[:3,:3].T (the caller normalises first, :236-237)."""
    return _dir_feature(w2c_ref, rays_dir, normalize=False)


def _dir_feature(w2c_ref, rays_dir, normalize=True):
    if not rays_dir.is_cuda:
        raise L.UcnError("gen_dir_feature: tensors must be on a CUDA device (no CPU fallback exists)")
    if not normalize:      # reference signature: already-normalised directions; one tiny matmul, not a hot op
        return rays_dir @ w2c_ref[:3, :3].t()
    fr = PackedFrame(rays_dir.device)
    rot = _host(w2c_ref)[:3, :3]
    fr.c.dir_rot0[:] = rot.reshape(-1).tolist()
    fr.c.dir_rot1[:] = rot.reshape(-1).tolist()
    fr.c.dir_switch = 2 ** 31 - 1
    R = rays_dir.shape[0]
    angle = torch.empty((R, 3), device=rays_dir.device)
    L.check(L.load().ucn_dir_feature(C.byref(fr.c), L.ptr(rays_dir.to(torch.float32).contiguous()), R, 0,
                                     L.ptr(angle), L.stream_ptr()), "ucn_dir_feature")
    return angle


def gen_pts_feats(imgs, volume_feature, rays_pts, pose_ref, rays_ndc, feat_dim, img_feat=None, near_fars=None,
                  img_downscale=1.0, use_color_volume=False, net_type='v0', confidence=None):
    """network/renderer.py:177-212 -> input_feat [R,S,feat_dim] (volumes 24 | rgb+mask 4V | feats 8V | conf 1)."""
    if use_color_volume:
        raise L.UcnError("use_color_volume=True is not part of the UC-NeRF hot path (opt.py:34 default False)")
    feat = _rays.gather_features(volume_feature, imgs, img_feat, confidence, pose_ref, rays_pts, rays_ndc)
    if feat.shape[-1] != feat_dim:
        raise L.UcnError(f"feat_dim={feat_dim} does not match 24+12*V+1={feat.shape[-1]} (train.py:31)")
    return feat


def batchify(fn, chunk):
    """network/renderer.py:57-75 (slices dim 0 = rays, Q19).  Chunking is a no-op need for the fused kernel,
    kept for API parity."""
    if chunk is None:
        return fn

    def ret(inputs, alpha_only):
        if alpha_only:
            raise L.UcnError("alpha_only path is dead in the reference (forward_alpha is broken, Q8)")
        return torch.cat([fn(inputs[i:i + chunk]) for i in range(0, inputs.shape[0], chunk)], 0)
    return ret


def run_network_mvs(pts, viewdirs, alpha_feat, fn, embed_fn, embeddirs_fn, netchunk=1024):
    """network/renderer.py:78-106.  With this package's UCNeRF and Embedders the [R,S,63+F+27] input is never
    materialised: PE, concat and the MLP run inside one kernel."""
    from .models import UCNeRF, _is_default_embedders
    if isinstance(fn, UCNeRF) and viewdirs is not None and viewdirs.dim() == 2 and alpha_feat is not None \
            and _is_default_embedders(embed_fn, embeddirs_fn):
        return fn.forward_fused(pts, viewdirs, alpha_feat)
    # generic composition (foreign embedders / foreign network): same tensor plumbing as the reference
    if embed_fn is not None:
        pts = embed_fn(pts)
    if alpha_feat is not None:
        pts = torch.cat((pts, alpha_feat), dim=-1)
    if viewdirs is not None:
        if viewdirs.dim() != 3:
            viewdirs = viewdirs[:, None].expand(-1, pts.shape[1], -1)
        if embeddirs_fn is not None:
            viewdirs = embeddirs_fn(viewdirs)
        pts = torch.cat([pts, viewdirs], -1)
    outputs_flat = batchify(fn, netchunk)(pts, viewdirs is None)
    return torch.reshape(outputs_flat, list(pts.shape[:-1]) + [outputs_flat.shape[-1]])


def rendering(args, pose_ref, rays_pts, rays_ndc, depth_candidates, rays_dir, volume_feature=None, imgs=None,
              network_fn=None, near_fars=None, img_feat=None, network_query_fn=None, white_bkgd=False,
              confidence=None, **kwargs):
    """network/renderer.py:215-255 -> (rgb_map [R,3], depth_map [R])."""
    # view-dir feature in the frame of whatever pose is first AT CALL TIME (Q4), then strip in place
    angle = _dir_feature(pose_ref['w2cs'][0], rays_dir, normalize=True) if pose_ref is not None else \
        rays_dir / torch.norm(rays_dir, dim=-1, keepdim=True)
    if pose_ref['w2cs'].shape[0] == args.view_num:
        pose_ref['w2cs'] = pose_ref['w2cs'][1:]
        pose_ref['intrinsics'] = pose_ref['intrinsics'][1:]
    input_feat = gen_pts_feats(imgs, volume_feature, rays_pts, pose_ref, rays_ndc, args.feat_dim, img_feat, near_fars,
                               args.img_downscale, args.use_color_volume, args.net_type, confidence=confidence)
    raw = network_query_fn(rays_ndc['ndc'], angle, input_feat, network_fn)
    rgb_map, depth_map, _, _ = _composite(raw, depth_candidates, white_bkgd, False)
    return rgb_map, depth_map
