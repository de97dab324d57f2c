"""bench_train.py -- BASELINE.json configs[3]: one UC-NeRF training step per rank with an NCCL gradient all-reduce.

Driven by `bench.py --workload c4` (same flags, same JSON contract).  A "step" is the reference's
`training_step` body for the per-ray path (train.py:136-190) on config 4 of SURVEY.md 8(d): C1-shaped maps
(320x256, 3 source views, 126 samples/ray) resident in HBM as the encoder would leave them (volumes, FPN features and
confidence require grad), batch_size 1024 (50 confidence-drawn 4x4 patches + 224 random rays) + 1024 sparse-depth
rays = 2 048 rays:

    build_rays (selection in PyTorch, sampling on K1)  ->  rendering (K2, K3 training primitives, K4)
    ->  loss = 5 * img2mse + 0.05 * (weighted sparse-depth MSE + edge-preserving patch smoothness)
               + 0.008 * scale-invariant patch gradient loss      (train.py:164-188, ucnerf_b200/losses.py;
        cas_mvsnet_loss acts on the encoder's outputs and is not part of this path)
    ->  backward (K4b, K3b, K2b)  ->  one all-reduce(sum)/P of the flat gradient bucket  ->  Adam step

Every rank has its own target pose and sampling seed (data parallel); the only collective is the gradient bucket
(177 034 fp32 MLP parameters; the reference also trains its encoder, which is outside this path).
`value` = training rays/s of the whole job (P * 2 048 / max-over-ranks step time, CUDA events).
"""
import json
import os
import sys
import time

import numpy as np

CFG = dict(H=256, W=320, V=3, S=126, batch_size=1024, patch_num=50, patch_size=4, n_depth=1024)
NAME = ("UC-NeRF training step: 320x256 maps, 3 source views, 126 samples/ray (reference rejects 128), batch 1024 rays "
        "(50 4x4 patches + 224 random) + 1024 sparse-depth rays, img + sparse-depth loss, NCCL grad all-reduce")
FLOP_PER_SAMPLE = lambda V: 2 * (138112 + 1536 * V)


def _oracle_step(O, OracleOps, TP, sc, wts, pix, t_rand, target, sparse, S):
    """The same training step restated on the CPU (oracle/): forward, loss cotangents, backward."""
    import torch
    osc = O.OracleScene(sc, S, dir_switch=1 << 30)
    r = O.build_rays_test(osc, pix[0], pix[1], t_rand)
    feat = O.gen_pts_feats(osc, r["pts"], r["ndc_all"])
    ang = O.dir_feature(osc, r["rays_d"])
    R, _, F = feat.shape
    T = torch.from_numpy
    P = {n: T(np.ascontiguousarray(wts["nerf." + n])) for n in TP.PARAM_NAMES}
    raw, saved = TP.mlp_forward_train(OracleOps(), P, T(r["ndc"]["ndc"].reshape(-1, 3).copy()), T(ang.copy()),
                                      T(feat.reshape(-1, F).copy()), S)
    raw = raw.numpy().reshape(R, S, 4)
    rgb, depth = O.raw2outputs(raw, r["z"])
    g_rgb = (10.0 / (3 * R)) * (rgb - target)
    g_depth = (0.1 / R) * (depth - sparse)
    d_raw = O.raw2outputs_bwd(raw, r["z"], g_rgb.astype(np.float32), g_depth.astype(np.float32))
    G = {n: torch.zeros_like(p) for n, p in P.items()}
    d_feat = TP.mlp_backward(OracleOps(), P, saved, T(d_raw.reshape(-1, 4)), G)
    O.gen_pts_feats_bwd(osc, r["pts"], r["ndc_all"], d_feat.numpy().reshape(R, S, F))
    return float(((rgb - target) ** 2).mean())


def cpu_step_rate(n_rays, repeats=1):
    """rays/s of the CPU restatement of the training step on a bounded sample of the batch."""
    from oracle import oracle as O
    from oracle.train_ops import OracleOps
    from ucnerf_b200 import synthetic, train_path as TP
    H, W, V, S = CFG["H"], CFG["W"], CFG["V"], CFG["S"]
    sc, wts = synthetic.make_scene(H, W, V, seed=0), synthetic.make_weights(V, seed=0)
    O.set_num_threads(os.cpu_count())
    rng = np.random.default_rng(0)
    pix = np.stack((rng.integers(0, H, n_rays), rng.integers(0, W, n_rays))).astype(np.float32)
    t_rand = rng.random((n_rays, S), dtype=np.float32)
    target, sparse = rng.random((n_rays, 3), dtype=np.float32), rng.random((n_rays,), dtype=np.float32)
    t0 = time.perf_counter()
    for _ in range(repeats):
        _oracle_step(O, OracleOps, TP, sc, wts, pix, t_rand, target, sparse, S)
    dt = (time.perf_counter() - t0) / repeats
    return n_rays / dt, dt, O.num_threads()


def run_reference(args):
    """`--impl reference --workload c4`: the CPU restatement of the step on the host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.cpu_rays or 256
    cpu_step_rate(n)                                    # warm-up (BLAS threads, page-in)
    rate, dt, cores = cpu_step_rate(n, repeats=max(1, args.steps))
    print(json.dumps({
        "impl": "reference", "metric": "rays/sec", "value": rate, "unit": "rays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": NAME, **CFG},
        "cpu_baseline": {"value": rate, "unit": "rays/s", "cores": cores, "kind": "port",
                         "sample": f"{n} of the 2048 rays of one training step per step (forward + backward, oracle/)"},
        "e2e": {"value": rate, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def run(args, ClockSampler, peaks):
    """`bench.py --workload c4`: one JSON line for the training step."""
    import torch.distributed as dist
    import bench
    ctx = bench.init_dist()
    line = measure(args, ClockSampler, peaks, ctx, full=True)
    if line is not None:
        print(json.dumps(line))
    if ctx[1] > 1:
        dist.destroy_process_group()


def measure(args, ClockSampler, peaks, ctx, full=True):
    """The config-4 measurement on the current process group; returns the line (rank 0) or None."""
    import torch
    import torch.distributed as dist
    from types import SimpleNamespace
    from ucnerf_b200 import lib as L, losses, models, rays, renderer, synthetic, train_path as TP

    rank, world, local, dev = ctx
    L.check(L.load().ucn_check_device(), "ucn_check_device")
    H, W, V, S = CFG["H"], CFG["W"], CFG["V"], CFG["S"]
    N_rays, n_depth = CFG["batch_size"], CFG["n_depth"]
    R = N_rays + n_depth
    sc = synthetic.make_scene(H, W, V, seed=0, target_motion=0.25 * rank)     # every rank trains on its own target view
    wts = synthetic.make_weights(V, seed=0)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    outputs = {f"stage{k + 1}": {"depth_values": T(sc.depth_values[k]),
                                 "volume_feature_no_ref": T(sc.volumes[k]).requires_grad_(True)} for k in range(3)}
    outputs["stage3"]["img_feats"] = T(sc.img_feats).requires_grad_(True)
    conf = T(sc.confidence).requires_grad_(True)
    map_leaves = [outputs[f"stage{k + 1}"]["volume_feature_no_ref"] for k in range(3)] + [outputs["stage3"]["img_feats"], conf]
    src_imgs = T(sc.imgs)
    gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
    target_img = torch.rand((1, 1, 3, H, W), generator=gen).to(dev)
    imgs_full = torch.cat([target_img, src_imgs.reshape(1, V, 3, H, W)], dim=1)
    sparse_depths = (sc.near_fars[0, 0] + (sc.near_fars[0, 1] - sc.near_fars[0, 0]) * torch.rand((1, H, W), generator=gen)).to(dev)
    sparse_weights = torch.ones((1, H, W), device=dev)
    dpt = (1.0 + torch.rand((1, H, W), generator=gen)).to(dev)              # monocular depth prior (train.py:174)
    w2cs, c2ws, Ks, near_fars = T(sc.w2cs), T(sc.c2ws), T(sc.K), T(sc.near_fars)
    targs = SimpleNamespace(view_num=V + 1, feat_dim=sc.feat_dim, img_downscale=1.0, use_color_volume=False,
                            net_type="v0", netchunk=1 << 20, patch_num=CFG["patch_num"], patch_size=CFG["patch_size"],
                            batch_size=N_rays, N_samples=S, multires=10, multires_views=4, netdepth=6, netwidth=128)
    model = models.UCNeRF(D=6, W=128, input_ch_pts=63, skips=[4], input_ch_views=27, input_ch_feat=sc.feat_dim,
                          view_num=V + 1, mode="fp32")
    model.load_state_dict({k: torch.from_numpy(v) for k, v in wts.items()})
    model.to(dev).train()
    embed_fn, _ = models.get_embedder(10, 0)
    embeddirs_fn, _ = models.get_embedder(4, 0)
    query = lambda pts, dirs, feats, fn: renderer.run_network_mvs(pts, dirs, feats, fn, embed_fn=embed_fn,
                                                                  embeddirs_fn=embeddirs_fn, netchunk=1 << 20)
    opt = torch.optim.Adam(model.parameters(), lr=5e-4, betas=(0.9, 0.999))
    params = list(model.parameters())
    n_grad_params = sum(p.numel() for p in params)
    host_rng = np.random.default_rng(99 + rank)
    coords_pinned = torch.empty((n_depth, 2), dtype=torch.int64).pin_memory()
    phases = ("sample", "forward", "backward", "allreduce", "optimizer")
    box = {"bucket": None, "loss": None}

    def step(i, timers=None, read_loss=False):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(phases) + 1)] if timers is not None else None
        rays._FRAME_CACHE.clear()        # the encoder emits new maps every step: K0 packing is part of the step
        mark = (lambda k: ev[k].record()) if ev else (lambda k: None)
        # the data loader's part: this step's sparse-depth coordinates arrive from the host
        coords_pinned.copy_(torch.from_numpy(np.stack((host_rng.integers(0, H, n_depth), host_rng.integers(0, W, n_depth)), 1)))
        coords = coords_pinned.to(dev, non_blocking=True)
        mark(0)
        pose_ref = {"w2cs": w2cs, "intrinsics": Ks, "c2ws": c2ws, "near_fars": near_fars}
        with torch.no_grad():
            pts, dirs, target_s, ndc, z, rays_o, _, ndcp, pix = rays.build_rays(
                targs, imgs_full, conf, sparse_depths, coords, pose_ref, w2cs, c2ws, Ks, N_rays, S, with_depth=True,
                outputs=outputs, seed=1000 * rank + i)
        mark(1)
        rgb, depth_pred = renderer.rendering(targs, pose_ref, pts, ndc, z, dirs, outputs, src_imgs,
                                             near_fars=near_fars[0], img_feat=outputs["stage3"]["img_feats"],
                                             confidence=conf, ndc_parameters=ndcp, network_fn=model, network_query_fn=query)
        # train.py:164-188: value and gradient of every renderer-side term from one launch (ucn_render_losses)
        loss, _ = losses.render_losses_fused(rgb, depth_pred, target_s, pix, N_rays, CFG["patch_num"], CFG["patch_size"],
                                             sparse_depths, sparse_weights, dpt)
        mark(2)
        opt.zero_grad(set_to_none=True)
        for t in map_leaves:
            t.grad = None
        loss.backward()
        mark(3)
        box["bucket"] = TP.allreduce_grads(params, world, box["bucket"])
        mark(4)
        opt.step()
        mark(5)
        if timers is not None:
            timers.append(ev)
        if read_loss:
            box["loss"] = float(loss.item())              # device -> host read of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = torch.tensor([e0.elapsed_time(e1), wall], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms[0].item() / steps, ms[1].item() / steps

    import bench as _bench
    sampler = _bench.make_clock_sampler(local) if (full and rank == 0) else None   # before the warm-up: see bench.measure_render
    if sampler:
        sampler.start()
        time.sleep(0.3)
    for i in range(max(3, args.warmup)):
        step(i, read_loss=True)
    loss0 = box["loss"]
    timers = []
    L.launch_count(reset=True)
    if sampler:
        sampler.mark()
    ms, wall_ms = timed(lambda i: step(100 + i, timers), args.steps)
    launches = L.launch_count() // args.steps
    clocks = sampler.stop() if (rank == 0 and sampler) else None
    e2e_ms, e2e_wall = timed(lambda i: step(200 + i, None, read_loss=True), args.steps)
    e2e_ms = max(e2e_ms, e2e_wall)
    kern = {p: sum(ev[k].elapsed_time(ev[k + 1]) for ev in timers) / args.steps for k, p in enumerate(phases)}

    # ---- parity of the timed configuration: the gradients of one full-size batch on the tensor-core GEMMs against the same
    # batch on the fp32 CUDA-core GEMM (ucn_train_gemm_mode 1), every parameter tensor and every map gradient -----------------
    def grads_of(mode):
        L.check(L.load().ucn_train_gemm_mode(mode), "ucn_train_gemm_mode")
        try:
            rays._FRAME_CACHE.clear()
            crd = torch.from_numpy(np.stack((np.arange(n_depth) % H, (7 * np.arange(n_depth)) % W), 1)).to(dev)
            pose_ref = {"w2cs": w2cs, "intrinsics": Ks, "c2ws": c2ws, "near_fars": near_fars}
            with torch.no_grad():
                # (same pixels and same jitter in both modes: the selection is keyed by its own seed)
                sel = rays.select_training_pixels(H, W, conf, crd, N_rays, CFG["patch_num"], CFG["patch_size"], seed=4242)
                pts, dirs, target_s, ndc, z, rays_o, _, ndcp, pix = rays.build_rays(
                    targs, imgs_full, conf, sparse_depths, crd, pose_ref, w2cs, c2ws, Ks, N_rays, S, with_depth=True,
                    outputs=outputs, pixel_coordinates=sel, seed=424242)
            rgb, depth_pred = renderer.rendering(targs, pose_ref, pts, ndc, z, dirs, outputs, src_imgs,
                                                 near_fars=near_fars[0], img_feat=outputs["stage3"]["img_feats"],
                                                 confidence=conf, ndc_parameters=ndcp, network_fn=model, network_query_fn=query)
            loss, _ = losses.render_losses_fused(rgb, depth_pred, target_s, pix, N_rays, CFG["patch_num"], CFG["patch_size"],
                                                 sparse_depths, sparse_weights, dpt)
            opt.zero_grad(set_to_none=True)
            for t in map_leaves:
                t.grad = None
            loss.backward()
            return [None if t.grad is None else t.grad.detach().clone() for t in params + map_leaves], float(loss.item())
        finally:
            L.check(L.load().ucn_train_gemm_mode(0), "ucn_train_gemm_mode")

    parity = None
    if rank == 0:
        import ctypes as _C
        lib_ = L.load()
        lib_.ucn_debug_gemm_flags.argtypes = [_C.c_int]

        def worst_rel(ga, gb):
            w, n = 0.0, 0
            for a, b in zip(ga, gb):
                if a is None or b is None:
                    continue
                scale = float(b.abs().max())
                if scale > 0.0:
                    w, n = max(w, float((a - b).abs().max()) / scale), n + 1
            return w, n

        g_tc, loss_tc = grads_of(0)                      # as timed: gemm_ws / gemm_wsw / gemm_tc
        lib_.ucn_debug_gemm_flags(192)                   # every product on the single-role kernel the golden tests pin
        try:
            g_old, loss_old = grads_of(0)
        finally:
            lib_.ucn_debug_gemm_flags(0)
        g_ref, loss_ref = grads_of(1)                    # fp32 CUDA-core GEMMs
        d_pipe, n_cmp = worst_rel(g_tc, g_old)
        d_simt, _ = worst_rel(g_tc, g_ref)
        d_simt_old, _ = worst_rel(g_old, g_ref)
        parity = {"tensors": n_cmp, "loss": [loss_tc, loss_old, loss_ref],
                  "max_rel_grad_diff_vs_single_role_tc_kernel": d_pipe, "gate": 5e-5,
                  "max_rel_grad_diff_vs_fp32_cuda_core_gemm": d_simt,
                  "single_role_tc_kernel_vs_fp32_cuda_core_gemm": d_simt_old,
                  "against": "one full-size batch (same pixels, same jitter), 30 parameter gradients + 5 map gradients, each relative "
                             "to its largest entry.  Gated: the warp-specialised pipelines against gemm_tc_kernel, the kernel the "
                             "reference-autograd goldens pin at 1e-4 on small batches (same products, other summation order of dW), "
                             "and the loss of the three forms.  Reported only: tensor-core against fp32 CUDA-core GEMMs -- at 258 K "
                             "samples x 128 units x 8 ReLU layers a 1e-6 difference in a pre-activation flips a few hundred masks "
                             "and each flip moves a gradient sum by a whole term; the single-role kernel shows the same figure",
                  "ok": bool(d_pipe <= 5e-5 and abs(loss_tc - loss_ref) <= 1e-5 * abs(loss_ref) and abs(loss_tc - loss_old) <= 1e-6 * abs(loss_old))}
        if d_pipe > 1e-3 or abs(loss_tc - loss_ref) > 1e-3 * abs(loss_ref):      # gross: the timed step is wrong
            raise SystemExit(f"bench_train: parity check failed: {parity}")
    line = None
    if rank == 0:
        pk = peaks()
        n_samples = R * S
        flops = 3 * FLOP_PER_SAMPLE(V) * n_samples                       # forward + grad-input + grad-weight
        fb_ms = kern["forward"] + kern["backward"]
        achieved = flops / (fb_ms * 1e-3) / 1e12
        line = {
            "metric": "rays/sec", "value": world * R / (ms * 1e-3), "unit": "rays/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": NAME, **CFG, "rays_per_step_per_gpu": R, "optimizer": "Adam",
                       "arithmetic": "fp32 operands split into three bf16 parts, six tcgen05 MMAs per product, fp32 accumulate (fp32-class: gradients within 1e-4 of the reference's autograd); heads and elementwise in fp32",
                       "loss": "5*img2mse + 0.05*(weighted sparse-depth MSE + patch smoothness) + 0.008*scale-invariant patch gradient (train.py:164-188 minus the encoder-side cas_mvsnet_loss)",
                       "l2": "saved activations 2.7 GB per step exceed the 126 MB L2",
                       "parallelism": f"data parallel x{world}: one all-reduce of {n_grad_params} fp32 gradients "
                                      f"({n_grad_params * 4 / 1e6:.2f} MB) per step -- the renderer MLP's parameters; the CasMVSNet "
                                      "encoder the reference also trains (934 304 parameters, network/models.py:249-250, "
                                      "train.py:136) is outside this path, so the map gradients K2b produces end at leaf tensors "
                                      "and are not in the bucket"},
            "e2e": {"value": world * R / (e2e_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": n_depth * 16,
                    "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"kernel": "gemm_ws_kernel + gemm_wsw_kernel + gemm_tc_kernel (nn.Linear forward / grad-input / grad-weight on tcgen05, three bf16 parts per fp32 operand)", "bound": "tensor",
                         "achieved": achieved, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                         "frac": achieved / pk["bf16_tflops_sustained"], "traffic": None,
                         "peak_source": pk["_source"] + " (bf16_tflops_sustained)",
                         "note": "algorithmic FLOPs (3 x forward) over the forward + backward phases of the step, which also hold the gathers, elementwise and head kernels; six MMAs are issued per algorithmic product",
                         "flops_per_step": flops, "ms": fb_ms},
            "kernels_ms_per_step": kern, "wall_ms_per_step": wall_ms,
            "loss_first_last": [loss0, box["loss"]], "parity": parity,
        }
        if full and world == 1 and not args.no_cpu_baseline:
            n = args.cpu_rays or 256
            cpu_step_rate(min(n, 64))
            rate, dt, cores = cpu_step_rate(n)
            line["cpu_baseline"] = {"value": rate, "unit": "rays/s", "cores": cores, "kind": "port",
                                    "sample": f"{n} of the 2048 rays of one training step, forward + backward ({dt:.1f} s)"}
    return line
