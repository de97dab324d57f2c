"""GPU (-m gpu): the CUDA path, called through the reference-shaped Python API over the C-ABI, against
(a) the CPU oracle on the same seeded inputs and (b) golden vectors produced by the reference's own code.

Gates (BASELINE.json north_star / SURVEY.md §8d):
  sample positions (z, pts, 4x ndc) and gather indices/masks ... bit-exact
  feature rows ............................................... <= 1e-5 abs
  fp32 path rgb / depth ...................................... <= 1e-4 abs
  bf16 tcgen05 path .......................................... <= 1e-3 abs per-pixel rgb, <= 0.05 dB PSNR
"""
import copy
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from helpers import CASES, bits, load_case
from oracle import oracle as O

pytestmark = pytest.mark.gpu

DEV = "cuda"


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def ref_inputs(c):
    """The dicts/tensors the reference's eval loop passes around (train.py:238-272), on the GPU."""
    sc = c["scene"]
    outputs = {f"stage{k + 1}": {"depth_values": T(sc.depth_values[k]), "volume_feature_no_ref": T(sc.volumes[k])}
               for k in range(3)}
    outputs["stage3"]["img_feats"] = T(sc.img_feats)
    pose_ref = {"w2cs": T(sc.w2cs), "intrinsics": T(sc.K), "c2ws": T(sc.c2ws), "near_fars": T(sc.near_fars)}
    return outputs, pose_ref, T(sc.imgs), T(sc.confidence)


def make_model(c, mode):
    from ucnerf_b200 import models
    m = models.UCNeRF(D=6, W=128, input_ch_pts=63, skips=[4], input_ch_views=27, input_ch_feat=c["scene"].feat_dim,
                      view_num=c["V"] + 1, mode=mode)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in c["weights"].items()})
    return m.to(DEV).eval()          # eval loop (train.py:251-279); the training path needs train() (tests/test_gpu_grad.py)


def make_query(mode="fp32"):
    from ucnerf_b200 import models, renderer
    embed_fn, _ = models.get_embedder(10, 0)
    embeddirs_fn, _ = models.get_embedder(4, 0)
    return lambda pts, dirs, feats, fn: renderer.run_network_mvs(pts, dirs, feats, fn, embed_fn=embed_fn,
                                                                 embeddirs_fn=embeddirs_fn, netchunk=1024)


def psnr(a, b):
    return -10.0 * np.log10(np.mean((a - b) ** 2) + 1e-20)


@pytest.mark.parametrize("name", CASES)
def test_build_rays_test_bit_exact(name):
    from ucnerf_b200 import rays
    c = load_case(name)
    g, sc = c["g"], c["scene"]
    outputs, pose_ref, _, _ = ref_inputs(c)
    H, W, S, chunk = c["H"], c["W"], c["S"], c["chunk"]
    osc = O.OracleScene(sc, S)
    n_chunks = (H * W + chunk - 1) // chunk
    for idx in range(n_chunks):
        lo, hi = idx * chunk, min(H * W, (idx + 1) * chunk)
        pts, dirs, ndc, z, rays_o, ndcp = rays.build_rays_test(
            H, W, pose_ref["c2ws"][0], pose_ref["w2cs"][0], pose_ref["intrinsics"][0], pose_ref["near_fars"],
            pose_ref["near_fars"][-1], S, chunk=chunk, idx=idx, outputs=outputs, t_rand=T(c["t_rand"][lo:hi]))
        py, px = O.chunk_pixels(H, W, chunk, idx)
        o = O.build_rays_test(osc, py, px, c["t_rand"][lo:hi])
        for got, want_o, want_g in ((z, o["z"], g["z"][lo:hi]), (pts, o["pts"], g["pts"][lo:hi]),
                                    (dirs, o["rays_d"], g["rays_d"][lo:hi])):
            got = got.cpu().numpy()
            assert np.array_equal(bits(got), bits(want_o))
            assert np.array_equal(bits(got), bits(want_g))
        for k in ("stage1", "stage2", "stage3", "ndc"):
            got = ndc[k].cpu().numpy()
            assert np.array_equal(bits(got), bits(o["ndc"][k])), k
            assert np.array_equal(bits(got), bits(g[k][lo:hi])), k
        assert rays_o.shape == (hi - lo, 3) and torch.equal(rays_o[0], pose_ref["c2ws"][0][:3, 3])
        assert set(ndcp) == {"w2c_ref", "intrinsic_ref", "inv_scale", "near", "far", "pad"}


@pytest.mark.parametrize("name", CASES)
def test_gather_indices_and_features(name):
    from ucnerf_b200 import rays, renderer
    c = load_case(name)
    g, sc = c["g"], c["scene"]
    outputs, pose_ref, imgs, conf = ref_inputs(c)
    pose_src = {"w2cs": pose_ref["w2cs"][1:], "intrinsics": pose_ref["intrinsics"][1:]}
    osc = O.OracleScene(sc, c["S"])
    ndc = {k: T(g[k]) for k in ("stage1", "stage2", "stage3", "ndc")}
    pts = T(g["pts"])
    feat, idx = rays.gather_features(outputs, imgs, outputs["stage3"]["img_feats"], conf, pose_src, pts, ndc, with_idx=True)
    ndc_all = np.stack([g[k] for k in ("stage1", "stage2", "stage3", "ndc")])
    feat_o, idx_o = O.gen_pts_feats(osc, g["pts"], ndc_all, with_idx=True)
    assert np.array_equal(idx.cpu().numpy(), idx_o)                       # floor indices + masks: bit-exact
    feat = feat.cpu().numpy()
    assert np.abs(feat - feat_o).max() <= 1e-5
    assert np.abs(feat[::c["stride"]] - g["feat_sub"]).max() <= 1e-5      # vs the reference's gen_pts_feats
    # the two reference functions on their own
    vf, cs = rays.index_point_feature(outputs, conf.reshape(1, 1, 1, c["H"], c["W"]), ndc)
    assert np.array_equal(vf.cpu().numpy(), feat[..., :24]) and np.array_equal(cs.cpu().numpy(), feat[..., -1])
    cv = rays.build_color_volume(pts, pose_src, imgs, img_feat=outputs["stage3"]["img_feats"], with_mask=True)
    assert np.array_equal(cv.cpu().numpy(), feat[..., 24:-1])
    f2 = renderer.gen_pts_feats(imgs, outputs, pts, pose_src, ndc, sc.feat_dim, outputs["stage3"]["img_feats"],
                                confidence=conf)
    assert np.array_equal(f2.cpu().numpy(), feat)


def test_gather_matches_torch_grid_sample():
    """Cross-check against the library op the reference calls (F.grid_sample on this GPU)."""
    import torch.nn.functional as F
    from ucnerf_b200 import rays
    c = load_case("small_v3")
    g = c["g"]
    outputs, pose_ref, imgs, conf = ref_inputs(c)
    ndc = {k: T(g[k]) for k in ("stage1", "stage2", "stage3", "ndc")}
    vf, cs = rays.index_point_feature(outputs, conf.reshape(1, 1, 1, c["H"], c["W"]), ndc)
    R, S = g["z"].shape
    for k, name in enumerate(("stage1", "stage2", "stage3")):
        grid = ndc[name].view(-1, 1, R, S, 3) * 2 - 1.0
        want = F.grid_sample(outputs[name]["volume_feature_no_ref"], grid, align_corners=False, mode="bilinear",
                             padding_mode="border")[:, :, 0].permute(2, 3, 0, 1).squeeze()
        assert (vf[..., 8 * k:8 * k + 8] - want).abs().max().item() <= 1e-5


@pytest.mark.parametrize("name", CASES)
def test_mlp_fp32_and_composite(name):
    from ucnerf_b200 import renderer
    c = load_case(name)
    g = c["g"]
    model = make_model(c, "fp32")
    ow = O.OracleWeights(c["weights"], c["V"])
    R, S = g["z"].shape
    feat_o = O.gen_pts_feats(O.OracleScene(c["scene"], S), g["pts"],
                             np.stack([g[k] for k in ("stage1", "stage2", "stage3", "ndc")]))
    raw = make_query()(T(g["ndc"]), T(g["angle"]), T(feat_o), model).cpu().numpy()
    raw_o = O.run_network(ow, g["ndc"], feat_o, g["angle"])
    assert np.abs(raw - raw_o).max() <= 1e-4
    assert np.abs(raw - g["raw"]).max() <= 1e-4
    # UCNeRF.forward on the reference's concatenated [PE | feat | PE] input
    from ucnerf_b200 import models
    e0, _ = models.get_embedder(10, 0)
    e1, _ = models.get_embedder(4, 0)
    sub = slice(0, 40)
    x = torch.cat([e0(T(g["ndc"][sub])), T(feat_o[sub]), e1(T(g["angle"][sub])[:, None].expand(-1, S, -1))], -1)
    assert (model(x).cpu().numpy() - g["raw"][sub]).max() <= 1e-4
    rgb, disp, acc, w, depth, w2, var = renderer.raw2outputs(T(g["raw"]), T(g["z"]), None)
    assert np.abs(rgb.cpu().numpy() - g["rgb"]).max() <= 1e-4
    assert np.abs(depth.cpu().numpy() - g["depth"]).max() <= 1e-4
    rgb_o, depth_o = O.raw2outputs(g["raw"], g["z"])
    assert np.abs(rgb.cpu().numpy() - rgb_o).max() <= 1e-5
    assert np.abs(acc.cpu().numpy() - w.sum(1).cpu().numpy()).max() <= 1e-5


def _reference_loop(c, mode):
    """The reference's eval loop verbatim (train.py:251-275) on the drop-in API."""
    from ucnerf_b200 import rays, renderer
    outputs, pose_ref, imgs, conf = ref_inputs(c)
    H, W, S, chunk, V = c["H"], c["W"], c["S"], c["chunk"], c["V"]
    args = SimpleNamespace(view_num=V + 1, feat_dim=c["scene"].feat_dim, img_downscale=1.0, use_color_volume=False,
                           net_type="v0", netchunk=1024)
    model, query = make_model(c, mode), make_query()
    near_fars = pose_ref["near_fars"]
    tgt_to_world, world_to_ref, intrinsic = pose_ref["c2ws"][0], pose_ref["w2cs"][0], pose_ref["intrinsics"][0]
    rgbs, depths = [], []
    for idx in range((H * W + chunk - 1) // chunk):
        lo, hi = idx * chunk, min(H * W, (idx + 1) * chunk)
        pts, dirs, ndc, z, rays_o, ndcp = rays.build_rays_test(H, W, tgt_to_world, world_to_ref, intrinsic, near_fars,
                                                               near_fars[-1], S, chunk=chunk, idx=idx, outputs=outputs,
                                                               t_rand=T(c["t_rand"][lo:hi]))
        rgb, depth = renderer.rendering(args, pose_ref, pts, ndc, z, dirs, outputs, imgs, near_fars=near_fars[0],
                                        img_feat=outputs["stage3"]["img_feats"], confidence=conf, ndc_parameters=ndcp,
                                        network_fn=model, network_query_fn=query)
        rgbs.append(rgb.cpu())
        depths.append(depth.cpu())
    assert pose_ref["w2cs"].shape[0] == V                     # stripped in place, like the reference (Q4)
    return torch.cat(rgbs).numpy(), torch.cat(depths).numpy()


@pytest.mark.parametrize("name", CASES)
def test_drop_in_eval_loop_fp32(name):
    c = load_case(name)
    rgb, depth = _reference_loop(c, "fp32")
    assert np.abs(rgb - c["g"]["rgb"]).max() <= 1e-4
    assert np.abs(depth - c["g"]["depth"]).max() <= 1e-4


@pytest.mark.parametrize("name", CASES)
def test_whole_frame_fp32(name):
    from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights
    c = load_case(name)
    fr = PackedFrame.from_scene(c["scene"], c["S"], DEV, dir_switch=c["chunk"])
    pw = PackedWeights(c["weights"], c["V"], DEV)
    rgb, depth = FrameRenderer(fr, pw, "fp32").render(t_rand=T(c["t_rand"]))
    assert np.abs(rgb.cpu().numpy() - c["g"]["rgb"]).max() <= 1e-4
    assert np.abs(depth.cpu().numpy() - c["g"]["depth"]).max() <= 1e-4
    # ragged: sub-range of rays with a small workspace == the same rays of the full render, bit-identical
    n = c["H"] * c["W"]
    r2, d2 = FrameRenderer(fr, pw, "fp32", max_rays=97).render(ray0=33, n_rays=n - 70, t_rand=T(c["t_rand"][33:n - 37]))
    assert torch.equal(r2, rgb[33:n - 37]) and torch.equal(d2, depth[33:n - 37])


@pytest.mark.parametrize("name", CASES)
def test_whole_frame_bf16_tensor_core(name):
    from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights
    c = load_case(name)
    g = c["g"]
    fr = PackedFrame.from_scene(c["scene"], c["S"], DEV, dir_switch=c["chunk"])
    pw = PackedWeights(c["weights"], c["V"], DEV)
    rgb, depth = FrameRenderer(fr, pw, "bf16").render(t_rand=T(c["t_rand"]))
    rgb = rgb.cpu().numpy()
    assert np.abs(rgb - g["rgb"]).max() <= 1e-3
    pseudo_gt = np.random.default_rng(5).random(g["rgb"].shape).astype(np.float32)
    assert abs(psnr(rgb, pseudo_gt) - psnr(g["rgb"], pseudo_gt)) <= 0.05


def test_edge_cases():
    """Empty ray sets, out-of-frustum samples (mask 0, border clamp), explicit-pixel ray lists."""
    from ucnerf_b200 import rays
    from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights
    c = load_case("small_v3")
    sc = copy.deepcopy(c["scene"])
    sc.w2cs[1, 0, 3] += 3.0                       # push source view 1 far sideways: most samples project outside
    sc.c2ws[1] = np.linalg.inv(sc.w2cs[1])
    S = c["S"]
    osc = O.OracleScene(sc, S)
    py, px = O.chunk_pixels(c["H"], c["W"])
    o = O.build_rays_test(osc, py, px, c["t_rand"])
    feat_o, idx_o = O.gen_pts_feats(osc, o["pts"], o["ndc_all"], with_idx=True)
    assert (idx_o[..., 13] == 0).mean() > 0.5     # the shifted view really is mostly masked out
    fr = PackedFrame.from_scene(sc, S, DEV)
    outputs = {f"stage{k + 1}": {"volume_feature_no_ref": T(sc.volumes[k])} for k in range(3)}
    pose_src = {"w2cs": T(sc.w2cs[1:]), "intrinsics": T(sc.K[1:])}
    feat, idx = rays.gather_features(outputs, T(sc.imgs), T(sc.img_feats), T(sc.confidence), pose_src, T(o["pts"]),
                                     {k: T(v) for k, v in o["ndc"].items()}, with_idx=True)
    assert np.array_equal(idx.cpu().numpy(), idx_o)
    assert np.abs(feat.cpu().numpy() - feat_o).max() <= 1e-5
    pw = PackedWeights(c["weights"], c["V"], DEV)
    rgb, depth = FrameRenderer(fr, pw, "fp32").render(ray0=5, n_rays=0)
    assert rgb.shape == (0, 3) and depth.shape == (0,)
    # explicit pixel list == the same pixels of the grid
    ys, xs = T(py[40:90]), T(px[40:90])
    _, d, pix = rays.get_rays_for_pixels(c["H"], c["W"], T(sc.K[0]), T(sc.c2ws[0]), ys, xs)
    assert np.array_equal(bits(d.cpu().numpy()), bits(o["rays_d"][40:90]))


def test_philox_jitter_statistics_and_sortedness():
    """No t_rand given: in-kernel Philox jitter.  z stays inside its stratum, so z is sorted per ray;
    a different seed changes the draw; the same seed reproduces it bit-for-bit."""
    from ucnerf_b200 import rays
    c = load_case("s126_v3")
    outputs, pose_ref, _, _ = ref_inputs(c)
    a = (c["H"], c["W"], pose_ref["c2ws"][0], pose_ref["w2cs"][0], pose_ref["intrinsics"][0], pose_ref["near_fars"],
         pose_ref["near_fars"][-1], c["S"])
    z1 = rays.build_rays_test(*a, outputs=outputs, seed=1)[3]
    z1b = rays.build_rays_test(*a, outputs=outputs, seed=1)[3]
    z2 = rays.build_rays_test(*a, outputs=outputs, seed=2)[3]
    assert torch.equal(z1, z1b) and not torch.equal(z1, z2)
    assert bool((z1[:, 1:] >= z1[:, :-1]).all())
    # jitter uniformity: reconstruct u from the un-jittered strata
    z0 = rays.build_rays_test(*a, outputs=outputs, t_rand=torch.zeros_like(z1))[3]
    zh = rays.build_rays_test(*a, outputs=outputs, t_rand=torch.ones_like(z1))[3]
    u = ((z1 - z0) / (zh - z0).clamp_min(1e-12))[(zh - z0) > 1e-6]
    assert 0.45 < u.mean().item() < 0.55 and 0.0 <= u.min().item() and u.max().item() <= 1.0 + 1e-4


def _size_properties(H, W, V, S, seed, max_rays, shards):
    """BASELINE-config shapes: the CUDA path at full size, checked by (a) the oracle on three 256-ray windows -- first
    rows, the window that contains the dir-frame switch at ray 1024 (quirk Q4), last rows -- in fp32 (1e-4 rgb / depth)
    and bf16 (1e-3 rgb), and (b) size-independent properties: the frame rendered as `shards` row tiles (what
    `bench.py --gpus N` does, shard.row_tile) is bit-identical to the whole (launch-grouped) frame, z is sorted, depth
    lies inside [near, far] of the widest stage."""
    from ucnerf_b200 import shard, synthetic
    from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights
    sc = synthetic.make_scene(H, W, V, seed=seed)
    wts = synthetic.make_weights(V, seed=seed)
    fr = PackedFrame.from_scene(sc, S, DEV, dir_switch=1024)
    pw = PackedWeights(wts, V, DEV)
    n = H * W
    t_rand = torch.rand((n, S), device=DEV, generator=torch.Generator(DEV).manual_seed(3))
    osc, ow = O.OracleScene(sc, S, dir_switch=1024), O.OracleWeights(wts, V)
    windows = (0, 900, n - 256)
    want = {r0: O.render_rays(osc, ow, r0, 256, t_rand[r0:r0 + 256].cpu().numpy()) for r0 in windows}
    # bf16 (tcgen05): whole frame, launch-grouped
    ren = FrameRenderer(fr, pw, "bf16", max_rays=max_rays)
    rgb, depth = ren.render(t_rand=t_rand)
    for r0 in windows:
        err = np.abs(rgb[r0:r0 + 256].cpu().numpy() - want[r0][0]).max()
        assert err <= 1e-3, ("bf16", r0, err)
    assert bool(torch.isfinite(rgb).all()) and bool(torch.isfinite(depth).all())
    assert 0.0 <= float(rgb.min()) and float(rgb.max()) <= 1.0
    tiles = []
    for p in range(shards):
        r0, m = shard.row_tile(H, W, p, shards)
        tiles.append(ren.render(r0, m, t_rand=t_rand[r0:r0 + m]))
    assert torch.equal(torch.cat([t[0] for t in tiles]), rgb) and torch.equal(torch.cat([t[1] for t in tiles]), depth)
    del tiles
    # fp32: the windows only (ray0 / n_rays of the same frame), gate 1e-4 on rgb and depth
    ren32 = FrameRenderer(fr, pw, "fp32", max_rays=256)
    for r0 in windows:
        r32, d32 = ren32.render(r0, 256, t_rand=t_rand[r0:r0 + 256])
        assert np.abs(r32.cpu().numpy() - want[r0][0]).max() <= 1e-4, ("fp32 rgb", r0)
        assert np.abs(d32.cpu().numpy() - want[r0][1]).max() <= 1e-4, ("fp32 depth", r0)


def test_config1_size_properties():
    """BASELINE config 1: 320x256, 3 source views, 126 samples."""
    _size_properties(256, 320, 3, 126, seed=11, max_rays=256 * 320 // 2, shards=8)


def test_config2_size_properties():
    """BASELINE config 2 (the shape the headline number is quoted on): 640x512, 3 source views, 126 samples."""
    _size_properties(512, 640, 3, 126, seed=12, max_rays=512 * 640 // 2, shards=8)


def test_config5_size_properties():
    """BASELINE config 5: 1280x1024, 5 source views, 255 samples (85-long stage sweeps, F = 85, two-pass img-feature
    staging in K3, 8 scan groups in K4), launch-grouped like bench.py --workload c5."""
    per_ray = 255 * (4 * (24 + 12 * 5 + 1) + 32)
    _size_properties(1024, 1280, 5, 255, seed=13, max_rays=int(16e9 // per_ray) // 1280 * 1280, shards=8)


@pytest.mark.parametrize("name", ("lin", "lin_jit", "disp_jit"))
def test_ray_marcher_bit_exact(name, golden_dir):
    """ucn_ray_marcher (data/ray_utils.py:152-197) vs the oracle and the reference-generated golden, bit for bit."""
    import sys
    sys.path.insert(0, golden_dir)
    from make_golden_marcher import marcher_inputs
    from ucnerf_b200 import rays as R
    g = np.load(f"{golden_dir}/ray_marcher.npz")
    rays, rnd, c = marcher_inputs(name)
    xyz, o, d, z = R.ray_marcher(T(rays), N_samples=c["S"], lindisp=c["lindisp"], perturb=c["perturb"], t_rand=T(rnd))
    xyz_o, z_o = O.ray_marcher(rays, c["S"], c["lindisp"], c["perturb"], rnd)
    assert np.array_equal(bits(z.cpu().numpy()), bits(z_o)) and np.array_equal(bits(z_o), bits(g[name + "_z"]))
    assert np.array_equal(bits(xyz.cpu().numpy()), bits(xyz_o)) and np.array_equal(bits(xyz_o), bits(g[name + "_xyz"]))
    assert np.array_equal(o.cpu().numpy(), rays[:, 0:3]) and np.array_equal(d.cpu().numpy(), rays[:, 3:6])


def test_host_sweep_matches_frame_renderer():
    """frame.HostSweep (pinned host maps -> pinned host image, next view's maps prefetched on a side stream) returns
    exactly what FrameRenderer returns for device-resident maps, view after view."""
    from ucnerf_b200 import synthetic
    from ucnerf_b200.frame import FrameRenderer, HostSweep, PackedFrame, PackedWeights
    H, W, V, S = 24, 40, 3, 126
    pw = PackedWeights(synthetic.make_weights(V, seed=3), V, DEV)
    scenes = [synthetic.make_scene(H, W, V, seed=30 + i) for i in range(3)]

    def host_maps(sc):
        return {"dv": [torch.from_numpy(a).pin_memory() for a in sc.depth_values],
                "vol": [torch.from_numpy(a).pin_memory() for a in sc.volumes],
                "feat": torch.from_numpy(sc.img_feats).pin_memory(), "imgs": torch.from_numpy(sc.imgs).pin_memory(),
                "conf": torch.from_numpy(sc.confidence).pin_memory()}
    maps = [host_maps(sc) for sc in scenes]
    cur = {"sc": None}

    def make_frame(dev_maps):
        sc = cur["sc"]
        fr = PackedFrame(DEV)
        fr.set_target(H, W, S, sc.K[0], sc.c2ws[0], sc.w2cs[0], sc.near_fars[0, 0], sc.near_fars[0, 1],
                      dir_rot0=sc.w2cs[0], dir_rot1=sc.w2cs[1], dir_switch=1024)
        fr.set_sources(sc.w2cs[1:], sc.K[1:])
        return fr.pack_depth_values(dev_maps["dv"]).pack_volumes(dev_maps["vol"]).pack_sources(
            dev_maps["imgs"], dev_maps["feat"], dev_maps["conf"])
    sweep = HostSweep(pw, make_frame, H * W, "bf16", DEV)
    for i, sc in enumerate(scenes):
        cur["sc"] = sc
        out = sweep.render(i, maps[i], maps[i + 1] if i + 1 < len(maps) else None, seed=7).clone()
        rgb, depth = FrameRenderer(PackedFrame.from_scene(sc, S, DEV, dir_switch=1024), pw, "bf16").render(seed=7)
        assert torch.equal(out[:, :3], rgb.cpu()) and torch.equal(out[:, 3], depth.cpu()), i
    # pipelined use: view i+1 is submitted before view i is collected (two views in flight); same images
    sweep2 = HostSweep(pw, make_frame, H * W, "bf16", DEV)
    want = []
    for sc in scenes:
        rgb, depth = FrameRenderer(PackedFrame.from_scene(sc, S, DEV, dir_switch=1024), pw, "bf16").render(seed=7)
        want.append(torch.cat((rgb, depth[:, None]), 1).cpu())
    cur["sc"] = scenes[0]
    sweep2.submit(0, maps[0], maps[1], seed=7)
    for i in range(len(scenes)):
        if i + 1 < len(scenes):
            cur["sc"] = scenes[i + 1]
            sweep2.submit(i + 1, maps[i + 1], maps[i + 2] if i + 2 < len(scenes) else None, seed=7)
        assert torch.equal(sweep2.result(i), want[i]), i


def test_host_copy_cache_follows_the_tensor():
    """Camera tensors reach the kernels by value through a cached host copy (frame._host_copy): the cache is keyed by
    tensor identity + view geometry and invalidated by `_version`, so in-place writes, views and new tensors (possibly at
    a recycled device address) must all be seen."""
    from ucnerf_b200 import frame
    t = torch.eye(4, device=DEV)
    assert frame._host(t)[0, 0] == 1.0 and frame._host(t) is frame._host(t)          # second read: cached, no sync
    t[0, 0] = 5.0
    assert frame._host(t)[0, 0] == 5.0                                               # in-place write bumps _version
    assert frame._host(t[1])[1] == 1.0
    t.mul_(2.0)
    assert frame._host(t[1])[1] == 2.0 and frame.host_scalar(t[0, 0]) == 10.0        # views share the version counter
    with pytest.raises(ValueError):
        frame._host_copy(t)[0, 0] = 0.0                                              # cached arrays are read-only
    for k in range(4):                                                               # fresh tensors, recycled storage
        u = torch.full((4, 4), float(k), device=DEV)
        assert frame._host(u).sum() == 16.0 * k
        del u


def test_frame_cache_follows_tensor_identity():
    """rays._FRAME_CACHE must never serve frame A's packed cameras / near-far maps to frame B: every frame of an eval
    loop brings FRESH tensors (`_version` 0) that the caching allocator places at recycled addresses.  Render A, free
    it, build a same-shaped B, and compare B with an uncached evaluation."""
    from ucnerf_b200 import rays
    ca, cb = load_case("small_v3"), load_case("moved_v3")
    assert (ca["H"], ca["W"]) != (cb["H"], cb["W"])            # different goldens; B' below has A's shapes
    from ucnerf_b200 import synthetic
    scenes = [synthetic.make_scene(ca["H"], ca["W"], ca["V"], seed=100 + i, target_motion=0.5 * i) for i in range(4)]
    S = ca["S"]
    tr = T(ca["t_rand"])
    outs, ptrs = [], set()
    for sc in scenes:                                          # one frame at a time; previous tensors are freed
        c = dict(ca, scene=sc)
        outputs, pose_ref, _, _ = ref_inputs(c)
        ptrs.add(outputs["stage1"]["depth_values"].data_ptr())
        z = rays.build_rays_test(c["H"], c["W"], pose_ref["c2ws"][0], pose_ref["w2cs"][0], pose_ref["intrinsics"][0],
                                 pose_ref["near_fars"], pose_ref["near_fars"][-1], S, outputs=outputs, t_rand=tr)[3]
        outs.append(z.cpu().numpy())
        del outputs, pose_ref, z
    assert len(ptrs) < len(scenes)                             # the allocator did recycle addresses
    for sc, got in zip(scenes, outs):
        rays._FRAME_CACHE.clear()
        py, px = O.chunk_pixels(ca["H"], ca["W"])
        want = O.build_rays_test(O.OracleScene(sc, S), py, px, ca["t_rand"])["z"]
        assert np.array_equal(bits(got), bits(want))
    # same tensor objects again -> cache hit, same result; in-place edit -> miss (version bump)
    c = dict(ca, scene=scenes[0])
    outputs, pose_ref, _, _ = ref_inputs(c)
    a = (c["H"], c["W"], pose_ref["c2ws"][0], pose_ref["w2cs"][0], pose_ref["intrinsics"][0], pose_ref["near_fars"],
         pose_ref["near_fars"][-1], S)
    n0 = len(rays._FRAME_CACHE)
    z1 = rays.build_rays_test(*a, outputs=outputs, t_rand=tr)[3]
    z2 = rays.build_rays_test(*a, outputs=outputs, t_rand=tr)[3]
    assert torch.equal(z1, z2) and len(rays._FRAME_CACHE) == n0 + 1
    outputs["stage3"]["depth_values"].mul_(1.01)
    z3 = rays.build_rays_test(*a, outputs=outputs, t_rand=tr)[3]
    assert not torch.equal(z1, z3)


def test_default_seed_is_fresh_per_call():
    """No t_rand, no seed: every call draws a new Philox key from torch's CPU generator (the reference draws a fresh
    torch.rand per call, utils/utils.py:572,705); torch.manual_seed reproduces the sequence."""
    from ucnerf_b200 import rays
    from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights
    c = load_case("small_v3")
    outputs, pose_ref, _, _ = ref_inputs(c)
    a = (c["H"], c["W"], pose_ref["c2ws"][0], pose_ref["w2cs"][0], pose_ref["intrinsics"][0], pose_ref["near_fars"],
         pose_ref["near_fars"][-1], c["S"])
    torch.manual_seed(123)
    z1 = rays.build_rays_test(*a, outputs=outputs)[3]
    z2 = rays.build_rays_test(*a, outputs=outputs)[3]
    torch.manual_seed(123)
    z1b = rays.build_rays_test(*a, outputs=outputs)[3]
    assert not torch.equal(z1, z2) and torch.equal(z1, z1b)
    ren = FrameRenderer(PackedFrame.from_scene(c["scene"], c["S"], DEV), PackedWeights(c["weights"], c["V"], DEV), "fp32")
    r1, r2 = ren.render()[0], ren.render()[0]
    assert not torch.equal(r1, r2)
    assert torch.equal(ren.render(seed=5)[0], ren.render(seed=5)[0])


def test_wrong_device_tensor_is_rejected():
    """Kernels launch on the current device's stream: a tensor of another device must raise, not fault."""
    from ucnerf_b200 import lib as L
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    t = torch.zeros(4, device="cuda:1")
    with pytest.raises(L.UcnError):
        L.ptr(t)
    with torch.cuda.device(1):
        assert L.ptr(t).value == t.data_ptr()


def test_row_band_frame_matches_whole_frame():
    """Ray-tile sharding packs / uploads only the row band of the target-frustum maps a tile touches
    (PackedFrame.set_row_band, HostSweep band_keys -> ucn_h2d_rows + ucn_pack_*_rows).  Every tile of a banded
    frame must be bit-identical to the same rays of the whole frame, from resident maps and from host maps."""
    from ucnerf_b200 import lib as L, shard, synthetic
    from ucnerf_b200.frame import FrameRenderer, HostSweep, PackedFrame, PackedWeights
    H, W, V, S = 48, 40, 3, 126
    sc = synthetic.make_scene(H, W, V, seed=40, target_motion=0.5)
    pw = PackedWeights(synthetic.make_weights(V, seed=40), V, DEV)
    whole = FrameRenderer(PackedFrame.from_scene(sc, S, DEV, dir_switch=1024), pw, "bf16").render(seed=9)
    host = {"dv": [torch.from_numpy(a).pin_memory() for a in sc.depth_values],
            "vol": [torch.from_numpy(a).pin_memory() for a in sc.volumes],
            "feat": torch.from_numpy(sc.img_feats).pin_memory(), "imgs": torch.from_numpy(sc.imgs).pin_memory(),
            "conf": torch.from_numpy(sc.confidence).pin_memory()}
    resident = {k: ([t.to(DEV) for t in v] if isinstance(v, list) else v.to(DEV)) for k, v in host.items()}
    for world in (3, 4):
        for rank in range(world):
            ray0, n = shard.row_tile(H, W, rank, world)
            rows = (ray0 // W, (ray0 + n) // W)

            def make_frame(maps):
                fr = PackedFrame(DEV)
                fr.set_target(H, W, S, sc.K[0], sc.c2ws[0], sc.w2cs[0], sc.near_fars[0, 0], sc.near_fars[0, 1],
                              dir_rot0=sc.w2cs[0], dir_rot1=sc.w2cs[1], dir_switch=1024)
                fr.set_sources(sc.w2cs[1:], sc.K[1:])
                fr.set_row_band(*rows)
                return fr.pack_depth_values(maps["dv"]).pack_volumes(maps["vol"]).pack_sources(maps["imgs"], maps["feat"], maps["conf"])
            fr = make_frame(resident)                                   # full maps in HBM, band packed
            assert fr.map_bytes < PackedFrame.from_scene(sc, S, DEV).map_bytes
            rgb, depth = FrameRenderer(fr, pw, "bf16", max_rays=n).render(ray0, n, seed=9)
            assert torch.equal(rgb, whole[0][ray0:ray0 + n]) and torch.equal(depth, whole[1][ray0:ray0 + n]), (world, rank)
            with pytest.raises(L.UcnError):                             # rays outside the band are refused
                FrameRenderer(fr, pw, "bf16", max_rays=n).render(0, H * W, seed=9)
            sweep = HostSweep(pw, make_frame, n, "bf16", DEV, ray0=ray0, frame_hw=(H, W), band_keys=("dv", "vol", "conf"))
            full_bytes = sum(t.numel() * 4 for _, _, t in HostSweep._walk(host))
            assert sweep.h2d_bytes(host) < full_bytes
            out = sweep.render(0, host, None, seed=9)
            assert torch.equal(out[:, :3], whole[0][ray0:ray0 + n].cpu()) and torch.equal(out[:, 3], whole[1][ray0:ray0 + n].cpu())
    # a target that is NOT the reference camera cannot be banded
    fr = PackedFrame(DEV)
    fr.set_target(H, W, S, sc.K[0], sc.c2ws[0], sc.w2cs[1], sc.near_fars[0, 0], sc.near_fars[0, 1])
    with pytest.raises(L.UcnError):
        fr.set_row_band(0, 8)


@pytest.mark.parametrize("name", ("small_v3", "small_v5", "s126_v3"))
def test_operand_tiles_match_feature_rows(name):
    """K2's operand tiles (ucn_gather_operands_from_z: bf16 hi/lo K-chunks the tensor-core MLP copies verbatim) carry
    exactly the split of the fp32 feature rows K2 writes in its other layouts: hi + lo reproduces every channel to
    2^-16 relative, hi == bf16(x) bit for bit, ones column and zero padding in place, confidence fp32 verbatim."""
    import ctypes as C
    from ucnerf_b200 import lib as L
    from ucnerf_b200.frame import PackedFrame
    c = load_case(name)
    V, S = c["V"], c["S"]
    fr = PackedFrame.from_scene(c["scene"], S, DEV, dir_switch=c["chunk"])
    lib = L.load()
    R = 37                                                            # ragged: the last tile is partial
    z = T(c["g"]["z"][:R])
    n = R * S
    F = fr.F
    nt = (n + 127) // 128
    rows = torch.zeros((nt, F, 128), device=DEV)
    ndc = torch.empty((n, 3), device=DEV)
    L.check(lib.ucn_gather_tiles_from_z(C.byref(fr.c), L.ptr(z), None, 0, R, L.ptr(rows), L.ptr(ndc), L.stream_ptr()), "tiles")
    tb = int(lib.ucn_operand_tile_bytes(V))
    rec = torch.zeros((nt, tb), dtype=torch.uint8, device=DEV)
    ndc2 = torch.empty((n, 3), device=DEV)
    L.check(lib.ucn_gather_operands_from_z(C.byref(fr.c), L.ptr(z), None, 0, R, L.ptr(rec), L.ptr(ndc2), L.stream_ptr()), "operands")
    assert torch.equal(ndc, ndc2)
    kf, kb = 24 + 4 * V, 8 * V
    cf, cb = 3 + (V + 1) // 2, V                                      # stored chunks per half (film, img)
    assert tb == (2 * cf + 2 * cb) * 2048 + 512
    o = 0

    def take(chunks):
        nonlocal o
        t = rec[:, o:o + chunks * 2048].contiguous().view(torch.bfloat16).view(nt, chunks, 128, 8)
        o += chunks * 2048
        return t.permute(0, 1, 3, 2).reshape(nt, chunks * 8, 128).float()      # [tile][k][row]
    fh, fl, ih, il = take(cf), take(cf), take(cb), take(cb)
    conf = rec[:, o:].contiguous().view(torch.float32)
    valid = (torch.arange(nt * 128, device=DEV) < n).view(nt, 1, 128).float()
    for hi, lo, x, k in ((fh, fl, rows[:, :kf], kf), (ih, il, rows[:, kf:kf + kb], kb)):
        assert torch.equal(hi[:, :k], x.to(torch.bfloat16).float())   # hi is the bf16 rounding of the fp32 channel
        assert ((hi + lo)[:, :k] - x).abs().max().item() <= 2.0 ** -16 * max(1.0, x.abs().max().item())
        if hi.shape[1] > k:                                           # V odd: the ones column (bias) shares the last film chunk
            assert torch.equal(hi[:, k], valid[:, 0]) and float(lo[:, k].abs().max()) == 0.0
            assert float(hi[:, k + 1:].abs().max()) == 0.0 and float(lo[:, k + 1:].abs().max()) == 0.0
    assert torch.equal(conf, rows[:, F - 1])

def test_round1_x3_kernel_still_meets_the_gate():
    """ucn_mlp_x3_variant(1): the round-1 kernel pair (tile-transposed fp32 rows) stays selectable and parity-green."""
    from ucnerf_b200 import lib as L
    from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights
    lib = L.load()
    assert lib.ucn_mlp_x3_variant(0) in (1, 2)
    prev = lib.ucn_mlp_x3_variant(0)
    try:
        out = {}
        for variant in (1, 2):
            assert lib.ucn_mlp_x3_variant(variant) == variant
            for name in ("s126_v3", "small_v5"):
                c = load_case(name)
                fr = PackedFrame.from_scene(c["scene"], c["S"], DEV, dir_switch=c["chunk"])
                pw = PackedWeights(c["weights"], c["V"], DEV)
                rgb, _ = FrameRenderer(fr, pw, "bf16").render(t_rand=T(c["t_rand"]))
                assert np.abs(rgb.cpu().numpy() - c["g"]["rgb"]).max() <= 1e-3, (variant, name)
                out[(variant, name)] = rgb
        for name in ("s126_v3", "small_v5"):                          # same arithmetic, different schedule
            assert (out[(1, name)] - out[(2, name)]).abs().max().item() <= 2e-4
    finally:
        lib.ucn_mlp_x3_variant(prev)
