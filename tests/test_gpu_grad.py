"""GPU (-m gpu): the training path (backward kernels K4b / K3b / K2b through the C-ABI and the drop-in Python API)
against gradients produced by the reference's own autograd (tests/golden/grad_*.npz) and the CPU oracle.

Gate: gradients are fp32 sums of up to 10^5 terms accumulated with atomics, so the bound is relative to the largest
entry of each gradient tensor: 1e-4, for the nn.Linear products on the fp32 CUDA cores and on tcgen05 alike (three
bf16 parts per operand = 24 significant bits; a two-part split was measured at 2.6e-3: its forward error flips ReLU
masks, and a flipped mask moves a gradient sum by a whole term)."""
import ctypes as C
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from helpers import GRAD_CASES, load_grad_case, rel_err
from oracle import oracle as O
from oracle.train_ops import OracleOps

pytestmark = pytest.mark.gpu

DEV = "cuda"
TOL = 1e-4
GEMM_MODES = ((1, 1e-4), (0, 1e-4))        # (ucn_train_gemm_mode, gradient gate)


@pytest.fixture(params=GEMM_MODES, ids=("simt_fp32", "tcgen05_bf16x3parts"))
def gemm(request):
    from ucnerf_b200 import lib as L
    mode, tol = request.param
    L.check(L.load().ucn_train_gemm_mode(mode))
    yield tol
    L.check(L.load().ucn_train_gemm_mode(0))


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _oracle_inputs(c):
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
    py, px = O.chunk_pixels(c["H"], c["W"])
    R = c["R"]
    r = O.build_rays_test(osc, py[:R], px[:R], c["t_rand"][:R])
    return osc, r, O.gen_pts_feats(osc, r["pts"], r["ndc_all"]), O.dir_feature(osc, r["rays_d"])


@pytest.mark.parametrize("gemm_mode,tol", ((1, 1e-5), (0, 1e-5)))
def test_linear_primitives_match_torch(gemm_mode, tol):
    """ucn_linear_fwd / _bwd_input / _bwd_weight on strided views and ragged sizes vs torch fp64 matmul, on the fp32
    CUDA-core kernel (mode 1) and on the tcgen05 kernel (mode 0: three bf16 parts per operand, six MMAs per K step)."""
    from ucnerf_b200 import lib as L
    from ucnerf_b200.train_path import CudaOps
    ops = CudaOps()
    L.check(L.load().ucn_train_gemm_mode(gemm_mode))
    try:
        g = torch.Generator(device="cpu").manual_seed(0)
        for M, K, N in ((1000, 191, 128), (257, 63, 128), (4099, 128, 3), (130, 37, 128), (64, 155, 64), (5, 128, 1),
                        (20000, 128, 128), (333, 16, 24)):
            Xb = torch.randn((M, K + 5), generator=g).to(DEV)
            W = torch.randn((N, K + 3), generator=g).to(DEV)
            b = torch.randn((N,), generator=g).to(DEV)
            X, Wv = Xb[:, 2:2 + K], W[:, 1:1 + K]
            Yb = torch.zeros((M, N + 2), device=DEV)
            Y = Yb[:, 1:1 + N]
            ops.linear_fwd(X, Wv, b, Y)
            ref = X.double() @ Wv.double().t() + b.double()
            assert rel_err(Y.cpu().numpy(), ref.cpu().numpy()) <= tol, (M, K, N)
            assert float(Yb[:, 0].abs().max()) == 0.0 and float(Yb[:, N + 1].abs().max()) == 0.0
            ops.linear_fwd(X, Wv, None, Y, accumulate=True, relu=True)
            ref2 = torch.relu(ref + X.double() @ Wv.double().t())
            assert rel_err(Y.cpu().numpy(), ref2.cpu().numpy()) <= tol, (M, K, N)
            dY = torch.randn((M, N), generator=g).to(DEV)
            dX = torch.full((M, K), 7.0, device=DEV)
            ops.linear_bwd_input(dY, Wv, dX)
            assert rel_err(dX.cpu().numpy(), (dY.double() @ Wv.double()).cpu().numpy()) <= tol, (M, K, N)
            ops.linear_bwd_input(dY, Wv, dX, accumulate=True)
            assert rel_err(dX.cpu().numpy(), (2 * dY.double() @ Wv.double()).cpu().numpy()) <= tol, (M, K, N)
            dWb, db = torch.zeros((N, K + 3), device=DEV), torch.zeros((N,), device=DEV)
            ops.linear_bwd_weight(dY, X, dWb[:, 1:1 + K], db)
            assert rel_err(dWb[:, 1:1 + K].cpu().numpy(), (dY.double().t() @ X.double()).cpu().numpy()) <= tol, (M, K, N)
            assert rel_err(db.cpu().numpy(), dY.double().sum(0).cpu().numpy()) <= 1e-5
            assert float(dWb[:, 0].abs().max()) == 0.0 and float(dWb[:, 1 + K:].abs().max()) == 0.0
    finally:
        L.check(L.load().ucn_train_gemm_mode(0))


def test_warp_specialised_gemm_matches_the_single_role_kernel():
    """gemm_ws_kernel (forward / grad-input with 16-byte-aligned activation rows) against gemm_tc_kernel (debug flag 64
    routes the same call to it) and torch fp64: ragged row counts, partial 32-wide slabs, partial column tiles, the
    vector (ldc % 4 == 0) and scalar output forms, bias / accumulate / ReLU.  Both kernels issue the same six MMAs per
    K16 step on the same three-part operands, so the outputs must be bit-identical."""
    from ucnerf_b200 import lib as L
    lib = L.load()
    lib.ucn_debug_gemm_flags.argtypes = [C.c_int]
    g = torch.Generator(device="cpu").manual_seed(3)
    try:
        for M, K, N, ldy in ((1000, 40, 128, 128), (1153, 128, 64, 64), (2500, 192, 200, 200), (1100, 128, 130, 131),
                             (4096, 100, 128, 132), (20000, 128, 256, 256)):
            X = torch.randn((M, K), generator=g).to(DEV)
            W = torch.randn((N, K), generator=g).to(DEV)
            b = torch.randn((N + 3,), generator=g).to(DEV)[:N].contiguous()
            seed = torch.randn((M, ldy), generator=g).to(DEV)
            ref = X.double() @ W.double().t() + b.double()
            for flags in (0, 1, 3):                      # plain, accumulate, accumulate + ReLU
                outs = []
                for dbg in (64, 0):
                    lib.ucn_debug_gemm_flags(dbg)
                    Y = seed.clone()
                    L.check(lib.ucn_linear_fwd(L.ptr(X), K, L.ptr(W), K, L.ptr(b), M, K, N, L.ptr(Y), ldy, flags, L.stream_ptr()),
                            "ucn_linear_fwd")
                    outs.append(Y)
                assert torch.equal(outs[0], outs[1]), (M, K, N, ldy, flags)
                want = ref + (seed[:, :N].double() if flags & 1 else 0.0)
                if flags & 2:
                    want = torch.relu(want)
                assert rel_err(outs[1][:, :N].cpu().numpy(), want.cpu().numpy()) <= 1e-5, (M, K, N, flags)
                if ldy > N:
                    assert torch.equal(outs[1][:, N:], seed[:, N:]), "columns beyond N must stay untouched"
    finally:
        lib.ucn_debug_gemm_flags(0)


def test_mn_major_grad_weight_pipeline_matches_fp64():
    """gemm_wsw_kernel (grad-weight over >= 4096 samples, both operands staged in their natural layout as MN-major UMMA
    operands) against torch fp64 and against gemm_tc_kernel's split reduction (debug flag 128): ragged sample counts (a
    partial last 32-sample slab), partial output tiles in both directions, two column tiles, X views with a wider pitch,
    accumulation into a non-zero dW, and the forced 4-byte staging of an unaligned X (flag 256)."""
    from ucnerf_b200 import lib as L
    lib = L.load()
    lib.ucn_debug_gemm_flags.argtypes = [C.c_int]
    g = torch.Generator(device="cpu").manual_seed(5)
    try:
        for M, K, N, ldx, dbg in ((5000, 128, 128, 128, 0), (4100, 191, 128, 192, 0), (9001, 63, 256, 192, 0),
                                  (4097, 40, 64, 40, 0), (20000, 155, 128, 160, 0), (4500, 37, 128, 37, 256)):
            Xb = torch.randn((M, ldx), generator=g).to(DEV)
            X = Xb[:, :K]
            dY = torch.randn((M, N), generator=g).to(DEV)
            init = torch.randn((N, K + 2), generator=g).to(DEV)
            ref = init[:, :K].double() + dY.double().t() @ X.double()
            outs = []
            for fl in (128, dbg):
                lib.ucn_debug_gemm_flags(fl)
                dW, db = init.clone(), torch.zeros((N,), device=DEV)
                L.check(lib.ucn_linear_bwd_weight(L.ptr(dY), N, L.ptr(Xb), ldx, M, K, N, L.ptr(dW), K + 2, L.ptr(db), L.stream_ptr()),
                        "ucn_linear_bwd_weight")
                outs.append(dW)
                assert rel_err(dW[:, :K].cpu().numpy(), ref.cpu().numpy()) <= 1e-5, (M, K, N, fl)
                assert torch.equal(dW[:, K:], init[:, K:]), "columns beyond K must stay untouched"
                assert rel_err(db.cpu().numpy(), dY.double().sum(0).cpu().numpy()) <= 1e-5
            assert rel_err(outs[1][:, :K].cpu().numpy(), outs[0][:, :K].cpu().numpy()) <= 1e-5
    finally:
        lib.ucn_debug_gemm_flags(0)


@pytest.mark.parametrize("name", GRAD_CASES)
def test_composite_backward(name):
    from ucnerf_b200 import lib as L
    c = load_grad_case(name)
    R, S = c["R"], c["S"]
    raw, z = T(c["g"]["raw"][:R]), T(c["g"]["z"][:R])
    d_raw = torch.empty_like(raw)
    g_rgb, g_depth = T(c["g_rgb"]), T(c["g_depth"])          # named: a temporary would be freed before the launch
    L.check(L.load().ucn_composite_bwd(L.ptr(raw), L.ptr(z), R, S, 0, L.ptr(g_rgb), L.ptr(g_depth), L.ptr(d_raw),
                                       L.stream_ptr()))
    assert rel_err(d_raw.cpu().numpy(), c["gg"]["d_raw"]) <= TOL
    assert rel_err(d_raw.cpu().numpy(), O.raw2outputs_bwd(c["g"]["raw"][:R], c["g"]["z"][:R], c["g_rgb"], c["g_depth"])) <= TOL
    # white background (renderer.py:138-139): rgb += 1 - acc  =>  every sample's cotangent shifts by -sum(g_rgb)
    from ucnerf_b200.train_path import CompositeFn
    raw_g = raw.clone().requires_grad_(True)
    rgb, depth = CompositeFn.apply(raw_g, z, True)
    (rgb * g_rgb).sum().backward()
    sig = raw.clone().double().requires_grad_(True)
    alpha = 1 - torch.exp(-sig[..., 3])
    Tr = torch.cumprod(torch.cat([torch.ones_like(alpha[:, :1]), 1 - alpha + 1e-10], -1), -1)[:, :-1]
    w = alpha * Tr
    ref = ((w[..., None] * sig[..., :3]).sum(1) + (1 - w.sum(1))[:, None])
    (ref * g_rgb.double()).sum().backward()
    assert rel_err(rgb.detach().cpu().numpy(), ref.detach().cpu().numpy()) <= 1e-5
    assert rel_err(raw_g.grad.cpu().numpy(), sig.grad.cpu().numpy()) <= TOL


@pytest.mark.parametrize("name", GRAD_CASES)
def test_mlp_train_primitives(name, gemm):
    """K3b: forward with saved activations + backward on the CUDA primitives vs the reference's autograd gradients."""
    TOL = gemm
    from ucnerf_b200 import train_path as TP
    c = load_grad_case(name)
    g = c["gg"]
    osc, r, feat, ang = _oracle_inputs(c)
    R, S, F = feat.shape
    P = {n: T(c["weights"]["nerf." + n]) for n in TP.PARAM_NAMES}
    raw, saved = TP.mlp_forward_train(TP.CudaOps(), P, T(r["ndc"]["ndc"].reshape(-1, 3)), T(ang), T(feat.reshape(-1, F)), S)
    got, want = raw.cpu().numpy().reshape(R, S, 4), c["g"]["raw"][:R]
    assert np.abs(got[..., :3] - want[..., :3]).max() <= 1e-4          # rgb in (0, 1): absolute
    assert rel_err(got[..., 3], want[..., 3]) <= 1e-4                   # sigma is unbounded: relative to its largest value
    G = {n: torch.zeros_like(p) for n, p in P.items()}
    d_feat = TP.mlp_backward(TP.CudaOps(), P, saved, T(g["d_raw"].reshape(-1, 4)), G)
    assert rel_err(d_feat.cpu().numpy().reshape(R, S, F), g["d_feat"]) <= TOL
    for n in TP.PARAM_NAMES:
        assert rel_err(G[n].cpu().numpy(), g["p:nerf." + n]) <= TOL, n


@pytest.mark.parametrize("name", GRAD_CASES)
def test_gather_backward(name):
    from ucnerf_b200 import lib as L
    from ucnerf_b200.frame import PackedFrame
    c = load_grad_case(name)
    g = c["gg"]
    osc, r, feat, ang = _oracle_inputs(c)
    fr = PackedFrame.from_scene(c["scene"], c["S"], device=DEV)
    sc = c["scene"]
    outs = [torch.zeros(sc.volumes[k].shape, device=DEV) for k in range(3)] + \
           [torch.zeros(sc.img_feats.shape, device=DEV), torch.zeros(sc.confidence.shape, device=DEV)]
    a = [T(r["pts"])] + [T(r["ndc"][k]) for k in ("stage1", "stage2", "stage3")]
    d_feat = T(g["d_feat"])
    L.check(L.load().ucn_gather_bwd(C.byref(fr.c), *[L.ptr(t) for t in a], c["R"] * c["S"], L.ptr(d_feat),
                                    *[L.ptr(o) for o in outs], L.stream_ptr()))
    for k in range(3):
        assert rel_err(outs[k].cpu().numpy(), g[f"d_vol{k}"]) <= TOL, k
    assert rel_err(outs[3].cpu().numpy(), g["d_img_feats"]) <= TOL
    assert rel_err(outs[4].cpu().numpy(), g["d_conf"]) <= TOL
    dv, dfe, dco = O.gen_pts_feats_bwd(osc, r["pts"], r["ndc_all"], g["d_feat"])
    assert rel_err(outs[0].cpu().numpy()[0], dv[0]) <= TOL


@pytest.mark.parametrize("name", GRAD_CASES)
def test_drop_in_training_step_gradients(name, gemm):
    """The reference's training-step graph (train.py:147-190) on the drop-in API: build rays, `rendering` with
    autograd, loss, backward -- every gradient the reference produces, same tensors, same shapes."""
    TOL = gemm
    from test_gpu_parity import make_model, make_query, ref_inputs
    from ucnerf_b200 import rays, renderer
    c = load_grad_case(name)
    g = c["gg"]
    H, W, S, V, R = c["H"], c["W"], c["S"], c["V"], c["R"]
    outputs, pose_ref, imgs, conf = ref_inputs(c)
    vols = [outputs[f"stage{k + 1}"]["volume_feature_no_ref"].requires_grad_(True) for k in range(3)]
    img_feats = outputs["stage3"]["img_feats"].requires_grad_(True)
    conf.requires_grad_(True)
    args = SimpleNamespace(view_num=V + 1, feat_dim=c["scene"].feat_dim, img_downscale=1.0, use_color_volume=False,
                           net_type="v0", netchunk=1 << 20)
    model, query = make_model(c, "fp32"), make_query()
    model.train()
    near_fars = pose_ref["near_fars"]
    with torch.no_grad():
        pts, dirs, ndc, z, rays_o, ndcp = rays.build_rays_test(H, W, pose_ref["c2ws"][0], pose_ref["w2cs"][0],
                                                               pose_ref["intrinsics"][0], near_fars, near_fars[-1], S,
                                                               chunk=R, idx=0, outputs=outputs, t_rand=T(c["t_rand"][:R]))
    rgb, depth = renderer.rendering(args, pose_ref, pts, ndc, z, dirs, outputs, imgs, near_fars=near_fars[0],
                                    img_feat=img_feats, confidence=conf, ndc_parameters=ndcp, network_fn=model,
                                    network_query_fn=query)
    assert np.abs(rgb.detach().cpu().numpy() - g["rgb"]).max() <= 1e-4
    assert np.abs(depth.detach().cpu().numpy() - g["depth"]).max() <= 1e-4
    loss = (rgb * T(c["g_rgb"])).sum() + (depth * T(c["g_depth"])).sum()
    loss.backward()
    for k in range(3):
        assert rel_err(vols[k].grad.cpu().numpy(), g[f"d_vol{k}"]) <= TOL, k
    assert rel_err(img_feats.grad.cpu().numpy(), g["d_img_feats"]) <= TOL
    assert rel_err(conf.grad.cpu().numpy(), g["d_conf"]) <= TOL
    n_checked = 0
    for n, p in model.named_parameters():
        if "p:" + n in g:
            assert rel_err(p.grad.cpu().numpy(), g["p:" + n]) <= TOL, n
            n_checked += 1
        else:
            assert p.grad is None, n                           # the three unused layers (Q9)
    assert n_checked == 30
    # an optimizer step on these gradients moves the render (the inference weight packs follow the parameters)
    zeros = torch.zeros((R, S, c["scene"].feat_dim), device=DEV)
    unit = dirs / dirs.norm(dim=-1, keepdim=True)
    with torch.no_grad():
        before = model.forward_fused(ndc["ndc"], unit, zeros).clone()
    torch.optim.SGD(model.parameters(), lr=1e-3).step()
    with torch.no_grad():
        after = model.forward_fused(ndc["ndc"], unit, zeros)
    assert float((after - before).abs().max()) > 0


@pytest.mark.parametrize("name", ("train_rays_small_v3", "train_rays_s126_v3"))
def test_build_rays_matches_reference_sampler(name):
    """a13: drop-in build_rays (selection injected) vs the reference's build_rays output, bit-exact."""
    from helpers import bits, load_train_rays_case
    from test_gpu_parity import ref_inputs
    from ucnerf_b200 import rays
    c = load_train_rays_case(name)
    g = c["tg"]
    outputs, pose_ref, src_imgs, conf = ref_inputs(c)
    H, W, S, V = c["H"], c["W"], c["S"], c["V"]
    imgs = torch.cat([T(c["target"])[None, None], src_imgs.reshape(1, V, 3, H, W)], dim=1)
    args = SimpleNamespace(patch_num=c["patch_num"], patch_size=c["patch_size"])
    pts, dirs, colors, ndc, z, rays_o, rays_depths, ndcp, pix = rays.build_rays(
        args, imgs, conf, None, T(c["coords"]), pose_ref, pose_ref["w2cs"], pose_ref["c2ws"], pose_ref["intrinsics"],
        c["N_rays"], S, with_depth=True, outputs=outputs, pixel_coordinates=T(g["pix"]), t_rand=T(g["t_rand"]))
    for got, key in ((z, "z"), (dirs, "rays_d"), (pts, "pts"), (colors, "colors"), (rays_o, "rays_o")):
        assert np.array_equal(bits(got.cpu().numpy()), bits(g[key])), key
    for k in ("stage1", "stage2", "stage3", "ndc"):
        assert np.array_equal(bits(ndc[k].cpu().numpy()), bits(g[k])), k
    assert np.array_equal(pix.cpu().numpy(), g["pix"]) and rays_depths is None
    with pytest.raises(Exception):
        rays.build_rays(args, imgs, conf, None, T(c["coords"]), pose_ref, pose_ref["w2cs"], pose_ref["c2ws"],
                        pose_ref["intrinsics"], c["N_rays"], S, with_depth=False, outputs=outputs)


@pytest.mark.parametrize("fused", (True, False))
def test_training_pixel_selection_properties(fused):
    """Selection in one launch (ucn_select_training_pixels) and its PyTorch form (with a torch.Generator): counts, bounds,
    patch structure, sparse-depth rays last (utils/utils.py:450-496)."""
    from ucnerf_b200 import rays
    H, W, ps, pn, N = 64, 80, 4, 10, 256
    conf = torch.rand((1, H, W), device=DEV)
    coords = torch.stack((torch.randint(0, H, (50,)), torch.randint(0, W, (50,))), 1).to(DEV)
    gen = None if fused else torch.Generator(DEV).manual_seed(5)
    pix = rays.select_training_pixels(H, W, conf, coords, N, pn, ps, generator=gen)
    assert tuple(pix.shape) == (2, N + 50) and pix.dtype == torch.float32
    assert float(pix[0].min()) >= 0 and float(pix[0].max()) <= H - 1 and float(pix[1].min()) >= 0 and float(pix[1].max()) <= W - 1
    patches = pix[:, :pn * ps * ps].reshape(2, pn, ps, ps)
    assert torch.equal(patches[0, :, 1:, :] - patches[0, :, :-1, :], torch.ones_like(patches[0, :, 1:, :]))   # rows step by 1
    assert torch.equal(patches[1, :, :, 1:] - patches[1, :, :, :-1], torch.ones_like(patches[1, :, :, 1:]))   # cols step by 1
    assert torch.equal(pix[:, N:].long(), coords.t())
    with pytest.raises(Exception):
        rays.select_training_pixels(H, W, conf, coords, pn * ps * ps - 1, pn, ps, generator=gen)        # quirk Q20
    if fused:                                                    # same seed -> same selection; another seed -> another one
        a = rays.select_training_pixels(H, W, conf, coords, N, pn, ps, seed=11)
        b = rays.select_training_pixels(H, W, conf, coords, N, pn, ps, seed=11)
        c = rays.select_training_pixels(H, W, conf, coords, N, pn, ps, seed=12)
        assert torch.equal(a, b) and not torch.equal(a, c)
        crop = rays.select_training_pixels(H, W, conf, None, 4096, pn, ps, seed=3, is_precrop_iters=True)
        assert crop.shape[1] == 4096


def test_fused_selection_follows_the_confidence():
    """Distribution check of the one-launch selection: with confidence concentrated on the left half of the image the
    confidence-drawn patches land there and the (1 - confidence)-drawn patches on the right (utils/utils.py:450-468);
    distinct pixels per draw (sampling WITHOUT replacement); uniform pixels cover the image; statistics agree with the
    torch.multinomial form on the same map."""
    from ucnerf_b200 import rays
    H, W, ps, pn = 64, 96, 4, 40
    conf = torch.full((1, H, W), 0.02, device=DEV)
    conf[:, :, :W // 2] = 0.98
    left_c, left_u, distinct = [], [], True
    for seed in range(40):
        pix = rays.select_training_pixels(H, W, conf, None, pn * ps * ps + 512, pn, ps, seed=seed)
        firsts = pix[:, :pn * ps * ps].reshape(2, pn, ps * ps)[:, :, 0]                      # top-left pixel of every patch
        left_c.append((firsts[1, :pn // 2] < W // 2 - ps).float().mean().item())             # drawn by confidence
        left_u.append((firsts[1, pn // 2:] < W // 2 - ps).float().mean().item())             # drawn by 1 - confidence
        rnd = pix[:, pn * ps * ps:]
        assert 0 <= float(rnd[0].min()) and float(rnd[0].max()) <= H - 1 and 0 <= float(rnd[1].min()) and float(rnd[1].max()) <= W - 1
        assert float(rnd[1].mean()) > 0.35 * W and float(rnd[1].mean()) < 0.65 * W
    # P(confidence draw on the left) = 0.98 / (0.98 + 0.02) = 0.98; the patch offset clamp / shift moves a few across
    assert np.mean(left_c) > 0.85 and np.mean(left_u) < 0.15, (np.mean(left_c), np.mean(left_u))
    gen = torch.Generator(DEV).manual_seed(0)
    ref = [rays.select_training_pixels(H, W, conf, None, pn * ps * ps + 512, pn, ps, generator=gen) for _ in range(40)]
    ref_c = np.mean([(p[:, :pn * ps * ps].reshape(2, pn, ps * ps)[1, :pn // 2, 0] < W // 2 - ps).float().mean().item() for p in ref])
    assert abs(ref_c - np.mean(left_c)) < 0.08, (ref_c, np.mean(left_c))


@pytest.mark.parametrize("R,S,V", ((37, 12, 3), (1, 3, 1), (5, 126, 5), (300, 3, 2)))
def test_mlp_train_ragged_and_tiny_batches(R, S, V, gemm):
    """Sample counts that are not multiples of the 128-row tensor-core tile (444), a single ray with the minimum S=3,
    V=1 .. 5 source views: CUDA primitives vs the CPU restatement (oracle/train_ops.py) on the same random inputs."""
    from ucnerf_b200 import synthetic, train_path as TP
    TOL = gemm
    rng = np.random.default_rng(R * 1000 + S * 10 + V)
    F = 24 + 12 * V + 1
    M = R * S
    wts = synthetic.make_weights(V, seed=5)
    feat = rng.standard_normal((M, F)).astype(np.float32)
    feat[:, -1] = rng.random(M).astype(np.float32)                      # confidence in [0, 1)
    ndc = rng.random((M, 3)).astype(np.float32)
    ang = rng.standard_normal((R, 3)).astype(np.float32)
    ang /= np.linalg.norm(ang, axis=1, keepdims=True)
    d_raw = rng.standard_normal((M, 4)).astype(np.float32)
    Tc = torch.from_numpy
    Pc = {n: Tc(np.ascontiguousarray(wts["nerf." + n])) for n in TP.PARAM_NAMES}
    raw_c, saved_c = TP.mlp_forward_train(OracleOps(), Pc, Tc(ndc.copy()), Tc(ang.copy()), Tc(feat.copy()), S)
    Gc = {n: torch.zeros_like(p) for n, p in Pc.items()}
    dfeat_c = TP.mlp_backward(OracleOps(), Pc, saved_c, Tc(d_raw.copy()), Gc)
    Pg = {n: T(wts["nerf." + n]) for n in TP.PARAM_NAMES}
    raw_g, saved_g = TP.mlp_forward_train(TP.CudaOps(), Pg, T(ndc), T(ang), T(feat), S)
    Gg = {n: torch.zeros_like(p) for n, p in Pg.items()}
    dfeat_g = TP.mlp_backward(TP.CudaOps(), Pg, saved_g, T(d_raw), Gg)
    assert np.abs(raw_g.cpu().numpy()[:, :3] - raw_c.numpy()[:, :3]).max() <= 1e-4
    assert rel_err(raw_g.cpu().numpy()[:, 3], raw_c.numpy()[:, 3]) <= 1e-4
    assert rel_err(dfeat_g.cpu().numpy(), dfeat_c.numpy()) <= TOL
    for n in TP.PARAM_NAMES:
        assert rel_err(Gg[n].cpu().numpy(), Gc[n].numpy()) <= TOL, n


def test_backward_edge_cases():
    """Empty batches are no-ops; white background and out-of-frustum samples go through K4b / K2b without NaNs."""
    from ucnerf_b200 import lib as L
    from ucnerf_b200.frame import PackedFrame
    from ucnerf_b200 import synthetic
    lib = L.load()
    one = torch.zeros((1, 12, 4), device=DEV)
    z = torch.zeros((1, 12), device=DEV)
    g3, g1, dr = torch.zeros((1, 3), device=DEV), torch.zeros((1,), device=DEV), torch.ones((1, 12, 4), device=DEV)
    L.check(lib.ucn_composite_bwd(L.ptr(one), L.ptr(z), 0, 12, 0, L.ptr(g3), L.ptr(g1), L.ptr(dr), L.stream_ptr()))
    assert float(dr.min()) == 1.0                                       # R = 0: nothing written
    sc = synthetic.make_scene(16, 20, 3, seed=9)
    fr = PackedFrame.from_scene(sc, 12, device=DEV)
    n = 64
    far = torch.full((n, 3), 1e4, device=DEV)                         # far outside every frustum / image
    ndc = torch.full((n, 3), 7.0, device=DEV)
    d_feat = torch.randn((n, 24 + 12 * 3 + 1), device=DEV)
    outs = [torch.zeros(sc.volumes[k].shape, device=DEV) for k in range(3)] + \
           [torch.zeros(sc.img_feats.shape, device=DEV), torch.zeros(sc.confidence.shape, device=DEV)]
    L.check(lib.ucn_gather_bwd(C.byref(fr.c), L.ptr(far), L.ptr(ndc), L.ptr(ndc), L.ptr(ndc), n, L.ptr(d_feat),
                               *[L.ptr(o) for o in outs], L.stream_ptr()))
    L.check(lib.ucn_gather_bwd(C.byref(fr.c), L.ptr(far), L.ptr(ndc), L.ptr(ndc), L.ptr(ndc), 0, L.ptr(d_feat),
                               *[L.ptr(o) for o in outs], L.stream_ptr()))
    for o in outs:
        assert bool(torch.isfinite(o).all())
    # border padding: a clamped sample deposits its whole cotangent on border texels (weights sum to 1 per sample)
    assert abs(float(outs[4].sum()) - float(d_feat[:, -1].sum())) <= 1e-3 * float(d_feat[:, -1].abs().sum())
    assert abs(float(outs[0].sum()) - float(d_feat[:, :8].sum())) <= 1e-3 * float(d_feat[:, :8].abs().sum())
