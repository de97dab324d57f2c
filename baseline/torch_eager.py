"""PyTorch-eager restatement of the reference's eval chunk loop, for timing on the SAME B200 (bench.py
`gpu_eager_baseline`; SURVEY.md 8(d) calls this loop "the bar to beat").

/root/reference does not travel to the GPU box, so this file restates its op sequence with the library calls the
reference makes -- per 1024-ray chunk (train.py:251-275):

    build_rays_test       utils/utils.py:600-739   meshgrid rays, 3 x S/3 sweeps, torch.sort, jitter, 4 x NDC
    index_point_feature   utils/utils.py:833-893   4 x F.grid_sample (3 volumes + confidence), align_corners=False, border
    build_color_volume    utils/utils.py:742-799   per source view: project, 2 x F.grid_sample align_corners=True, mask
    run_network_mvs       network/renderer.py:78-106   PE, cat, netchunk=1024 batches of 13 nn.Linear (fp32, TF32 off)
    raw2outputs           network/renderer.py:109-140  exp, cumprod, weighted sums
    rgb.cpu(), depth.cpu()                          train.py:274-275 (optional: `sync_per_chunk`)

It is a BASELINE only: nothing under uc-nerf_b200/ imports it, it is never the checker and never the product.
`bench.py` checks it against the oracle on one window (<= 1e-4) before timing it.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


def _pe(x, n_freq):
    fb = (2.0 ** torch.linspace(0.0, n_freq - 1, steps=n_freq)).to(x.device)
    sc = (x.unsqueeze(-2) * fb.view(*[1] * (x.dim() - 1), -1, 1)).reshape(*x.shape[:-1], -1)
    return torch.cat((x, torch.sin(sc), torch.cos(sc)), dim=-1)


class EagerMLP(nn.Module):
    """network/models.py:74-184 with the reference's parameter names (loads the same state_dict)."""

    def __init__(self, V, Wd=128, D=6, in_pts=63, in_views=27):
        super().__init__()
        self.V = V
        self.pts_linears = nn.ModuleList([nn.Linear(in_pts, Wd)] + [
            nn.Linear(Wd + in_pts if i == 4 else Wd, Wd) for i in range(D - 1)])
        self.pts_bias_depth_fine = nn.Linear(24 + 4 * V, Wd)
        self.pts_bias_confidence = nn.Linear(8 * V, Wd)
        self.pts_bias_confidence_1 = nn.Linear(1, 1)
        self.views_linears = nn.ModuleList([nn.Linear(in_views + Wd, Wd // 2)])
        self.view_confi_linears = nn.ModuleList([nn.Linear(in_views + Wd, Wd // 2)])
        self.feature_linear, self.feature_linear_1, self.confi_linear = nn.Linear(Wd, Wd), nn.Linear(Wd, Wd), nn.Linear(Wd, Wd)
        self.alpha_linear, self.alpha_linear_1 = nn.Linear(Wd // 2, 1), nn.Linear(Wd, 1)
        self.rgb_linear, self.confi_rgb_linear = nn.Linear(Wd // 2, 3), nn.Linear(Wd, 3)
        self.in_pts, self.in_views = in_pts, in_views

    def forward(self, x):
        V = self.V
        pts, feats, views = torch.split(x, [self.in_pts, x.shape[-1] - self.in_pts - self.in_views, self.in_views], dim=-1)
        film = self.pts_bias_depth_fine(torch.cat([feats[..., :24], feats[..., 24:24 + 4 * V]], dim=-1))
        fb = self.pts_bias_confidence(feats[..., 24 + 4 * V:24 + 12 * V])
        unc = 1 - feats[..., -1:]
        h = pts
        for i, lin in enumerate(self.pts_linears):
            h = F.relu(lin(h) * film)
            if i == 4:
                h = torch.cat([pts, h], -1)
        base_c, base_a = self.confi_rgb_linear(h), self.alpha_linear_1(h)
        h1 = torch.cat([self.feature_linear(h * fb), views], -1)
        ad_c = self.rgb_linear(F.relu(self.views_linears[0](h1)))
        ad_a = self.alpha_linear(F.relu(self.view_confi_linears[0](h1)))
        rgb = torch.sigmoid(base_c * (1 - unc) + ad_c * unc)
        alpha = torch.relu(ad_a * (1 - unc) + base_a * unc)
        return torch.cat([rgb, alpha], -1)


class EagerPath:
    """Device tensors in the reference's layouts + the chunk loop."""

    def __init__(self, scene, weights, S, device, chunk=1024, netchunk=1024):
        T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        self.H, self.W, self.V, self.S, self.chunk, self.netchunk, self.dev = scene.H, scene.W, scene.V, S, chunk, netchunk, device
        self.dv = [T(a) for a in scene.depth_values]
        self.vol = [T(a) for a in scene.volumes]
        self.img_feats, self.imgs, self.conf = T(scene.img_feats), T(scene.imgs), T(scene.confidence)
        self.K, self.c2ws, self.w2cs, self.near_fars = T(scene.K), T(scene.c2ws), T(scene.w2cs), T(scene.near_fars)
        self.mlp = EagerMLP(scene.V).to(device).eval()
        self.mlp.load_state_dict({k[len("nerf."):]: torch.from_numpy(v) for k, v in weights.items()})
        ys, xs = torch.meshgrid(torch.linspace(0, self.H - 1, self.H), torch.linspace(0, self.W - 1, self.W), indexing="ij")
        self.ys, self.xs = ys.reshape(-1), xs.reshape(-1)                 # host, like the reference (moved per chunk)

    # utils/utils.py:323-373
    def _ndc(self, w2c, K, pts, inv_scale, nf=None):
        R, S = pts.shape[:2]
        p = torch.matmul(pts.reshape(-1, 3), w2c[:3, :3].t()) + w2c[:3, 3:].reshape(1, 3)
        p[p[:, 2].abs() < 1e-4, 2] = 1e-4
        q = p @ K.t()
        q[:, :2] = (q[:, :2] / q[:, -1:] + 0.0) / inv_scale.reshape(1, 2)
        q = q.view(R, S, 3)
        if nf is None:
            return q
        out = {}
        for name, (n_, f_) in nf.items():
            out[name] = torch.cat([q[..., :2], (q[..., 2:] - n_) / (f_ - n_)], dim=-1)
        return out

    def build_rays(self, idx, t_rand=None):
        H, W, S, dev = self.H, self.W, self.S, self.dev
        ys, xs = self.ys[idx * self.chunk:(idx + 1) * self.chunk].to(dev), self.xs[idx * self.chunk:(idx + 1) * self.chunk].to(dev)
        K, c2w = self.K[0], self.c2ws[0]
        dirs = torch.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], torch.ones_like(xs)], -1)
        rays_d = dirs @ c2w[:3, :3].t()
        R = rays_d.shape[0]
        rays_o = c2w[:3, -1].clone().reshape(1, 3).expand(R, -1)
        inv_scale = torch.tensor([W - 1, H - 1]).to(dev)
        pix = torch.stack((ys, xs)).long()
        pix4, pix2 = torch.div(pix, 4, rounding_mode="trunc").long(), torch.div(pix, 2, rounding_mode="trunc").long()
        n = S // 3
        nf, zs = {}, []
        for k, p in ((0, pix4), (1, pix2), (2, pix)):
            near = self.dv[k][0, 0, p[0], p[1]].unsqueeze(1)
            far = self.dv[k][0, -1, p[0], p[1]].unsqueeze(1)
            t = torch.linspace(0., 1., steps=n).view(1, n).expand(R, n).to(dev)
            zs.append(near.expand(-1, n) * (1. - t) + far.expand(-1, n) * t)
            nf[f"stage{k + 1}"] = (near.expand(-1, S).unsqueeze(-1), far.expand(-1, S).unsqueeze(-1))
        nf["ndc"] = (self.near_fars[0, 0], self.near_fars[0, 1])
        z, _ = torch.sort(torch.cat(zs, dim=-1))
        mids = .5 * (z[..., 1:] + z[..., :-1])
        upper, lower = torch.cat([mids, z[..., -1:]], -1), torch.cat([z[..., :1], mids], -1)
        if t_rand is None:
            t_rand = torch.rand(z.shape, device=dev)
        z = lower + (upper - lower) * t_rand
        pts = rays_o.unsqueeze(1) + z.unsqueeze(-1) * rays_d.unsqueeze(1)
        ndc = self._ndc(self.w2cs[0], K, pts, inv_scale, nf)
        return pts, rays_d, ndc, z

    def render_chunk(self, idx, t_rand=None, first_chunk=False):
        H, W, V, S, dev = self.H, self.W, self.V, self.S, self.dev
        pts, rays_d, ndc, z = self.build_rays(idx, t_rand)
        R = pts.shape[0]
        cos = torch.norm(rays_d, dim=-1)
        # quirk Q4: the first chunk of a frame rotates by the target's w2c, later chunks by source view 1's
        rot = self.w2cs[0] if first_chunk else self.w2cs[1]
        angle = (rays_d / cos.unsqueeze(-1)) @ rot[:3, :3].t()
        feat = torch.empty((R, S, 24 + 12 * V + 1), device=dev)
        grids = [ndc[f"stage{k + 1}"].view(-1, 1, R, S, 3) * 2 - 1.0 for k in range(3)]
        gs = lambda vol, g: F.grid_sample(vol, g, align_corners=False, mode="bilinear",
                                          padding_mode="border")[:, :, 0].permute(2, 3, 0, 1).squeeze()
        conf_s = gs(self.conf.reshape(1, 1, 1, H, W), grids[2])
        feat[..., :24] = torch.cat([gs(self.vol[k], grids[k]) for k in range(3)], dim=-1)
        inv_scale = torch.tensor([W - 1, H - 1]).to(dev)
        colors = torch.empty((R, S, 4 * V), device=dev)
        all_feats = torch.tensor(()).to(dev)
        for v in range(V):
            g = self._ndc(self.w2cs[1 + v], self.K[1 + v].clone(), pts, inv_scale)[None][..., :2] * 2.0 - 1.0
            data = F.grid_sample(self.imgs[:, v], g, align_corners=True, mode="bilinear", padding_mode="border")
            all_feats = torch.cat((all_feats, F.grid_sample(self.img_feats[v], g, align_corners=True, mode="bilinear",
                                                            padding_mode="border")), dim=1)
            m = (g > -1.0) * (g < 1.0)
            data = torch.cat((data, (m[..., 0] * m[..., 1]).float().unsqueeze(1)), dim=1)
            colors[..., 4 * v:4 * v + 4] = data[0].permute(1, 2, 0)
        feat[..., 24:-1] = torch.cat([colors, all_feats[0].permute(1, 2, 0)], dim=-1)
        feat[..., -1:] = conf_s.unsqueeze(-1)
        x = torch.cat((_pe(ndc["ndc"], 10), feat), dim=-1)
        x = torch.cat([x, _pe(angle[:, None].expand(-1, S, -1), 4)], -1)
        raw = torch.cat([self.mlp(x[i:i + self.netchunk]) for i in range(0, R, self.netchunk)], 0)
        sigma = raw[..., 3]
        alpha = 1. - torch.exp(-sigma)
        Tr = torch.cumprod(torch.cat([torch.ones(R, 1).to(dev), 1. - alpha + 1e-10], -1), -1)[:, :-1]
        w = alpha * Tr
        return torch.sum(w[..., None] * raw[..., :3], -2), torch.sum(w * z, -1)


def time_chunks(path: EagerPath, chunk_ids, sync_per_chunk: bool, warm: int = 2):
    """Seconds per chunk (CUDA events + host clock, whichever is larger) over `chunk_ids`."""
    import time
    with torch.no_grad():
        for i in range(warm):
            path.render_chunk(chunk_ids[i % len(chunk_ids)])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for c in chunk_ids:
            rgb, depth = path.render_chunk(c, first_chunk=(c == 0))
            if sync_per_chunk:
                rgb, depth = rgb.cpu(), depth.cpu()
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    return max(wall, e0.elapsed_time(e1) * 1e-3) / len(chunk_ids)
