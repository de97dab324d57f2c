"""Generate golden vectors by EXECUTING THE REFERENCE'S OWN CODE (build container only).

The reference (wrld/UC-NeRF, /root/reference) is pure Python/PyTorch, ships no tests and no golden
vectors, so the parity pin for oracle/ is the reference run itself: this script imports
the reference behind name-only stub modules, feeds it the seeded synthetic
scenes of ucnerf_b200.synthetic, runs its eval loop verbatim --

    utils/utils.py:600  build_rays_test      network/renderer.py:215  rendering
    (-> get_rays_mvs, get_ndc_coordinate, index_point_feature, build_color_volume,
        run_network_mvs, UCNeRF.forward, raw2outputs)

-- chunk by chunk with the shipped loop structure (train.py:251-275, incl. the in-place pose_ref strip,
quirk Q4), and stores the outputs.  Inputs are NOT stored: they are regenerated from the seed.
/root/reference does not exist on the GPU box; only the .npz files travel.

Usage:  python tests/golden/make_golden.py [case ...]  (rewrites tests/golden/<case>.npz; default: all cases)
"""
import os
import sys
import types
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("UCNERF_REFERENCE", "/root/reference")

CASES = {
    # name: H, W, V(source views), S, chunk, seed, rays kept for the feature dump (stride)
    "small_v3": dict(H=24, W=32, V=3, S=12, chunk=256, seed=1, stride=4),
    "small_v5": dict(H=16, W=24, V=5, S=15, chunk=128, seed=2, stride=4),
    "s126_v3": dict(H=16, W=20, V=3, S=126, chunk=160, seed=3, stride=16),
    "moved_v3": dict(H=16, W=20, V=3, S=12, chunk=100, seed=4, stride=4, target_motion=1.0),
    # BASELINE config 5's sample / view counts (S=255 -> 85-long stage sweeps, V'=5 -> F=85) on a tiny frame, two
    # chunks so that the dir-frame switch (quirk Q4) falls inside
    "s255_v5": dict(H=8, W=12, V=5, S=255, chunk=48, seed=5, stride=8),
}


def import_reference():
    def mk(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    def create_meshgrid(height, width, normalized_coordinates=True, device=None, dtype=None):
        xs = torch.linspace(0, width - 1, width, device=device, dtype=dtype or torch.float32)
        ys = torch.linspace(0, height - 1, height, device=device, dtype=dtype or torch.float32)
        if normalized_coordinates:
            xs = (xs / (width - 1) - 0.5) * 2
            ys = (ys / (height - 1) - 0.5) * 2
        return torch.stack(torch.meshgrid([xs, ys], indexing="ij"), -1).permute(1, 0, 2).unsqueeze(0)

    k = mk("kornia", create_meshgrid=create_meshgrid)
    k.utils = mk("kornia.utils", create_meshgrid=create_meshgrid)
    mk("warmup_scheduler", GradualWarmupScheduler=object)
    mk("tkinter", X="x")
    mk("inplace_abn", InPlaceABN=torch.nn.BatchNorm2d)
    mk("imp")
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
    sys.path.insert(0, REF)
    import utils.utils as U            # noqa  (order matters, Q16)
    import network.models as M         # noqa
    import network.renderer as Rn      # noqa
    torch.autograd.set_detect_anomaly(False)
    return U, M, Rn


def run_case(U, M, Rn, name, H, W, V, S, chunk, seed, stride, target_motion=0.0):
    from ucnerf_b200 import synthetic
    sc = synthetic.make_scene(H, W, V, seed=seed, target_motion=target_motion)
    wts = synthetic.make_weights(V, seed=seed)
    T = torch.from_numpy
    view_num = V + 1
    args = SimpleNamespace(view_num=view_num, feat_dim=sc.feat_dim, img_downscale=1.0,
                           use_color_volume=False, net_type="v0", netchunk=1024)
    embed_fn, ic = M.get_embedder(10, 0)
    embeddirs, icv = M.get_embedder(4, 0)
    model = M.UCNeRF(D=6, W=128, input_ch_pts=ic, skips=[4], input_ch_views=icv,
                     input_ch_feat=args.feat_dim, view_num=view_num)
    sd = model.state_dict()
    assert set(sd.keys()) == set(wts.keys()), set(sd.keys()) ^ set(wts.keys())
    for k_, v_ in wts.items():
        assert tuple(sd[k_].shape) == v_.shape, (k_, sd[k_].shape, v_.shape)
    model.load_state_dict({k_: T(v_) for k_, v_ in wts.items()})
    model.eval()
    query = lambda pts, dirs, feats, fn: Rn.run_network_mvs(pts, dirs, feats, fn, embed_fn=embed_fn,
                                                            embeddirs_fn=embeddirs, netchunk=1024)
    outputs = {}
    for k_ in range(3):
        outputs[f"stage{k_ + 1}"] = {"depth_values": T(sc.depth_values[k_]),
                                     "volume_feature_no_ref": T(sc.volumes[k_])}
    outputs["stage3"]["img_feats"] = T(sc.img_feats)
    near_fars = T(sc.near_fars)
    pose_ref = {"w2cs": T(sc.w2cs).clone(), "intrinsics": T(sc.K).clone(), "c2ws": T(sc.c2ws).clone(),
                "near_fars": near_fars}
    imgs = T(sc.imgs)
    conf = T(sc.confidence)
    n_rays = H * W
    n_chunks = n_rays // chunk + int(n_rays % chunk > 0)
    rng = np.random.default_rng(7000 + seed)
    t_rand_all = rng.random((n_rays, S), dtype=np.float32)

    real_rand = torch.rand
    state = {"lo": 0}

    def fake_rand(*size, **kw):      # build_rays_test: torch.rand(depth_candidate.shape, device=device)
        shape = tuple(size[0]) if len(size) == 1 and not isinstance(size[0], int) else tuple(size)
        r = T(t_rand_all[state["lo"]:state["lo"] + shape[0]]).clone()
        assert tuple(r.shape) == shape, (r.shape, shape)
        return r

    out = {k_: [] for k_ in ("z", "rays_d", "pts", "stage1", "stage2", "stage3", "ndc", "rgb", "depth")}
    feats, raws, angles = [], [], []
    captured = {}
    orig_gen, orig_query = Rn.gen_pts_feats, query

    def spy_gen(*a, **kw):
        f = orig_gen(*a, **kw)
        captured["feat"] = f.detach().clone()
        return f

    def spy_query(pts, dirs, feats_, fn):
        captured["angle"] = dirs.detach().clone()
        r = orig_query(pts, dirs, feats_, fn)
        captured["raw"] = r.detach().clone()
        return r

    Rn.gen_pts_feats = spy_gen
    torch.rand = fake_rand
    try:
        with torch.no_grad():
            for idx in range(n_chunks):
                state["lo"] = idx * chunk
                pts, dirs, ndc, z, o, ndcp = U.build_rays_test(
                    H, W, pose_ref["c2ws"][0], T(sc.w2cs[0]), T(sc.K[0]), near_fars, near_fars[-1], S,
                    chunk=chunk, idx=idx, outputs=outputs)
                rgb, depth = Rn.rendering(args, pose_ref, pts, ndc, z, dirs, outputs, imgs,
                                          near_fars=near_fars[0], img_feat=outputs["stage3"]["img_feats"],
                                          confidence=conf, ndc_parameters=ndcp, network_fn=model,
                                          network_query_fn=spy_query)
                for k_, v_ in (("z", z), ("rays_d", dirs), ("pts", pts), ("rgb", rgb), ("depth", depth)):
                    out[k_].append(v_.numpy().copy())
                for k_ in ("stage1", "stage2", "stage3", "ndc"):
                    out[k_].append(ndc[k_].numpy().copy())
                feats.append(captured["feat"].numpy())
                raws.append(captured["raw"].numpy())
                angles.append(captured["angle"].numpy())
    finally:
        torch.rand = real_rand
        Rn.gen_pts_feats = orig_gen
    res = {k_: np.concatenate(v_, 0) for k_, v_ in out.items()}
    feat = np.concatenate(feats, 0)
    res["feat_sub"] = feat[::stride].copy()
    res["raw"] = np.concatenate(raws, 0)
    res["angle"] = np.concatenate(angles, 0)
    res["meta"] = np.array([H, W, V, S, chunk, seed, stride], np.int64)
    res["target_motion"] = np.float32(target_motion)
    sig = res["raw"][..., 3]
    print(f"{name}: rays={n_rays} chunks={n_chunks} sigma>0 {np.mean(sig > 0):.3f} mean sigma {sig.mean():.3f} "
          f"rgb[{res['rgb'].min():.3f},{res['rgb'].max():.3f}] depth[{res['depth'].min():.3f},{res['depth'].max():.3f}]")
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **res)


def main():
    torch.set_num_threads(4)
    U, M, Rn = import_reference()
    # side pins used by the oracle's arithmetic model
    lin = {str(n): torch.linspace(0., 1., n).numpy() for n in (1, 2, 3, 4, 5, 14, 30, 42, 43, 85, 86)}
    np.savez_compressed(os.path.join(HERE, "torch_linspace01.npz"), **lin)
    only = sys.argv[1:]
    for name, cfg in CASES.items():
        if not only or name in only:
            run_case(U, M, Rn, name, **cfg)


if __name__ == "__main__":
    main()
