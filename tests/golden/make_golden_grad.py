"""Golden GRADIENTS of the per-ray path, produced by the reference's own autograd (build container only).

Runs the reference's `build_rays_test` + `rendering` (network/renderer.py:215-255) on the first chunk of a seeded
synthetic scene with autograd enabled -- the same graph `train.py:147-163` builds with `build_rays`, whose sampling
maths is identical (utils/utils.py:510-573) and carries no gradient (depth hypotheses are detached,
network/mvs_models.py:716-717) -- and back-propagates  loss = sum(rgb * G_rgb) + sum(depth * G_depth)  with seeded
cotangents G.  Stored: d loss / d {raw, input_feat, every UCNeRF parameter, the three feature volumes, img_feats,
confidence}.  Inputs and cotangents are regenerated from the seed (`grad_cotangents`).

Usage:  python tests/golden/make_golden_grad.py        (rewrites tests/golden/grad_*.npz)
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

GRAD_CASES = {
    # name: (forward golden case whose scene/weights/jitter are reused, rays taken from the start of the frame)
    "grad_small_v3": ("small_v3", 256),
    "grad_small_v5": ("small_v5", 128),
}


def grad_cotangents(seed, R):
    rng = np.random.default_rng(9000 + seed)
    return rng.standard_normal((R, 3)).astype(np.float32), rng.standard_normal((R,)).astype(np.float32)


def run_case(U, M, Rn, name, base, R):
    from make_golden import CASES
    from ucnerf_b200 import synthetic
    c = CASES[base]
    H, W, V, S, seed = c["H"], c["W"], c["V"], c["S"], c["seed"]
    sc = synthetic.make_scene(H, W, V, seed=seed, target_motion=c.get("target_motion", 0.0))
    wts = synthetic.make_weights(V, seed=seed)
    T = torch.from_numpy
    args = SimpleNamespace(view_num=V + 1, feat_dim=sc.feat_dim, img_downscale=1.0, use_color_volume=False,
                           net_type="v0", netchunk=1 << 20)
    embed_fn, ic = M.get_embedder(10, 0)
    embeddirs, icv = M.get_embedder(4, 0)
    model = M.UCNeRF(D=6, W=128, input_ch_pts=ic, skips=[4], input_ch_views=icv, input_ch_feat=args.feat_dim,
                     view_num=V + 1)
    model.load_state_dict({k: T(v) for k, v in wts.items()})
    model.train()
    vols = [T(sc.volumes[k]).clone().requires_grad_(True) for k in range(3)]
    img_feats = T(sc.img_feats).clone().requires_grad_(True)
    conf = T(sc.confidence).clone().requires_grad_(True)
    outputs = {f"stage{k + 1}": {"depth_values": T(sc.depth_values[k]), "volume_feature_no_ref": vols[k]}
               for k in range(3)}
    outputs["stage3"]["img_feats"] = img_feats
    near_fars = T(sc.near_fars)
    pose_ref = {"w2cs": T(sc.w2cs).clone(), "intrinsics": T(sc.K).clone(), "c2ws": T(sc.c2ws).clone(),
                "near_fars": near_fars}
    t_rand = np.random.default_rng(7000 + seed).random((H * W, S), dtype=np.float32)
    real_rand = torch.rand
    torch.rand = lambda *a, **k: T(t_rand[:R]).clone()
    cap = {}
    orig_gen = Rn.gen_pts_feats

    def spy_gen(*a, **kw):
        f = orig_gen(*a, **kw)
        f.retain_grad()
        cap["feat"] = f
        return f

    def query(pts, dirs, feats, fn):
        r = Rn.run_network_mvs(pts, dirs, feats, fn, embed_fn=embed_fn, embeddirs_fn=embeddirs, netchunk=1 << 20)
        r.retain_grad()
        cap["raw"] = r
        return r

    Rn.gen_pts_feats = spy_gen
    try:
        with torch.no_grad():
            pts, dirs, ndc, z, o, ndcp = U.build_rays_test(H, W, pose_ref["c2ws"][0], T(sc.w2cs[0]), T(sc.K[0]),
                                                           near_fars, near_fars[-1], S, chunk=R, idx=0, outputs=outputs)
        rgb, depth = Rn.rendering(args, pose_ref, pts, ndc, z, dirs, outputs, T(sc.imgs), near_fars=near_fars[0],
                                  img_feat=img_feats, confidence=conf, ndc_parameters=ndcp, network_fn=model,
                                  network_query_fn=query)
    finally:
        torch.rand = real_rand
        Rn.gen_pts_feats = orig_gen
    g_rgb, g_depth = grad_cotangents(seed, R)
    loss = (rgb * T(g_rgb)).sum() + (depth * T(g_depth)).sum()
    loss.backward()
    res = {"meta": np.array([H, W, V, S, R, seed], np.int64), "base": np.array(base),
           "rgb": rgb.detach().numpy(), "depth": depth.detach().numpy(),
           "d_raw": cap["raw"].grad.numpy(), "d_feat": cap["feat"].grad.numpy(),
           "d_vol0": vols[0].grad.numpy(), "d_vol1": vols[1].grad.numpy(), "d_vol2": vols[2].grad.numpy(),
           "d_img_feats": img_feats.grad.numpy(), "d_conf": conf.grad.numpy()}
    unused = []
    for k, p in model.named_parameters():
        if p.grad is None:
            unused.append(k)
        else:
            res["p:" + k] = p.grad.numpy()
    print(f"{name}: R={R} loss={loss.item():.4f} |d_raw|max={np.abs(res['d_raw']).max():.3e} "
          f"|d_feat|max={np.abs(res['d_feat']).max():.3e} params with grad {sum(k.startswith('p:') for k in res)} "
          f"unused {sorted(unused)}")
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **res)


def main():
    from make_golden import import_reference
    torch.set_num_threads(4)
    U, M, Rn = import_reference()
    for name, (base, R) in GRAD_CASES.items():
        run_case(U, M, Rn, name, base, R)


if __name__ == "__main__":
    main()
