#!/usr/bin/env python
"""bench.py -- rays/s of the UC-NeRF per-ray rendering hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode bf16|fp32]
                    [--workload c2|c1|c5|c4] [--max-rays M]

A "step" renders ONE whole target view of the workload: K0 pack of that view's maps + K1 raygen/sampling
+ K2 projection/gathers + K3 PE+MLP + K4 compositing.  `value` = rays of the view / max-over-ranks device
time with the maps already in HBM; `e2e` = same call with HOST (pinned) maps copied in and rgb/depth
copied out inside the timed region.  N>1 (torchrun): one process per GPU, the view is partitioned into N
contiguous row tiles (shard.row_tile); every rank's compositing kernel stores its tile straight into the root's
peer-mapped frame buffer over NVLink (shard.PeerFrameBuffer), one 4-byte all-reduce per view is the barrier, and
the root reads the frame out -- STRONG scaling of one frame, the split BASELINE.json's north_star names.  Each
rank packs (and, end to end, uploads) only the rows of the target-frustum maps its tile touches.  `extra` holds
the same measurement for config 5 (full-box partition of a 1280x1024x255 frame), the config-4 training step with
its gradient all-reduce, the old one-view-per-rank replica number, and (1 GPU) the PyTorch-eager restatement of
the reference's chunk loop on the same GPU.  `--impl reference` times the CPU port of the reference's algorithm
(oracle/, pinned bit-exact to the reference) on the host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # BASELINE.json configs[1] / [0] / [4]; "128"/"256" samples are run as 126/255 because the reference
    # crashes unless N_samples % 3 == 0 (utils/utils.py:666,707-712)
    "c2": dict(H=512, W=640, V=3, S=126, name="SCARED-shaped eval render 640x512, 3 source views, 126 samples/ray "
                                              "(reference rejects 128: N_samples%3), random-init weights"),
    "c1": dict(H=256, W=320, V=3, S=126, name="UC-NeRF eval render 320x256, 3 source views, 126 samples/ray"),
    "c5": dict(H=1024, W=1280, V=5, S=255, name="High-res stress 1280x1024, 5 source views, 255 samples/ray"),
    # BASELINE.json configs[3]: training step + NCCL gradient all-reduce (bench_train.py)
    "c4": dict(H=256, W=320, V=3, S=126, name="UC-NeRF training step, 2048 rays, NCCL grad all-reduce"),
}
FLOP_PER_SAMPLE = lambda V: 2 * (138112 + 1536 * V)          # SURVEY.md §8.0 (algorithmic MLP FLOPs)
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def ncu_traffic(kernel, workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed `ncu --set full` capture
    (profiles/ncu_traffic.json, written by tools/ncu_traffic.py); None if that kernel/workload was not captured."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        return json.load(open(p)).get(workload, {}).get(kernel)
    except (OSError, ValueError):
        return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        d["_source"] = "measured"
        return d
    return dict(FALLBACK_PEAKS, _source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path, self.off = gpu_index, None, f"/tmp/ucnerf_clocks_{os.getpid()}.csv", 0

    def mark(self):
        try:
            self.off = os.path.getsize(self.path)
        except OSError:
            self.off = 0

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.proc.wait()
        sm, mx, reasons = [], [], set()
        fh = open(self.path)
        fh.seek(self.off)
        for line in fh:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        busy = sorted(sm)[len(sm) // 2:] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class NvmlClockSampler:
    """The same quantities through NVML in this process (a thread polling every 50 ms): SM clock, max SM clock and the
    throttle-reason bit mask.  `nvidia-smi -lms` polling the recipe's nine fields (power.draw among them) during the timed
    region was measured to slow the timed region itself down -- 37.7 -> 49.2 ms per frame with host submission going from
    0.6 to 12 ms per frame -- so it is the fallback (UCN_CLOCKS=smi), not the default."""
    REASONS = {0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x8: "hw_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index):
        self.idx, self.thread, self.samples, self.stop_flag, self.mx, self.t_mark = gpu_index, None, [], False, None, 0.0

    def start(self):
        import threading
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.idx]) if vis and vis.split(",")[self.idx].isdigit() else self.idx
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.thread = None
            return False

        def loop():
            while not self.stop_flag:
                try:
                    self.samples.append((time.perf_counter(), float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)),
                                         int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))))
                except Exception:
                    pass
                time.sleep(0.05)
        self.thread = threading.Thread(target=loop, daemon=True)
        self.thread.start()
        return True

    def stop(self):
        if self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        self.stop_flag = True
        self.thread.join()
        sel = [x for x in self.samples if x[0] >= self.t_mark] or self.samples      # the timed region's samples
        clk = sorted(x[1] for x in sel)
        mask = 0
        for x in sel:
            mask |= x[2]
        busy = clk[len(clk) // 2:] if clk else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": self.mx,
                "reasons": sorted(n for b, n in self.REASONS.items() if mask & b), "samples": len(sel),
                "source": "NVML in-process, 50 ms, samples taken during the timed region"}

    def mark(self):
        """Call right before the timed region: only later samples are reported."""
        self.t_mark = time.perf_counter()


def make_clock_sampler(gpu_index):
    """NVML in-process by default, `nvidia-smi -lms` with UCN_CLOCKS=smi, nothing with UCN_CLOCKS=off (diagnostics only)."""
    mode = os.environ.get("UCN_CLOCKS", "nvml")
    if mode == "off":
        return None
    if mode == "nvml":
        s = NvmlClockSampler(gpu_index)
        if s.start():
            s.start = lambda: None                       # already running
            return s
    return ClockSampler(gpu_index)


def cpu_sample_rays(O, osc, ow, H, W, S, target_s, row0):
    """Rays for a CPU sample of about `target_s` seconds: probe with 2048 rays, scale, clamp to the frame
    (whole image rows starting at `row0`, so the sample is a contiguous band of the same frame)."""
    probe = 2048
    tr = np.random.default_rng(1).random((probe, S), dtype=np.float32)
    O.render_rays(osc, ow, row0, 256, tr[:256])                       # page in / spin up the thread pool
    t0 = time.perf_counter()
    O.render_rays(osc, ow, row0, probe, tr)
    rate = probe / (time.perf_counter() - t0)
    n = int(rate * target_s) // W * W
    return max(W, min(n, H * W - row0))


def run_reference(args, wl):
    """CPU arm: the oracle port of the reference's path on all host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from ucnerf_b200 import synthetic
    from oracle import oracle as O
    H, W, V, S = wl["H"], wl["W"], wl["V"], wl["S"]
    sc = synthetic.make_scene(H, W, V, seed=0)
    wts = synthetic.make_weights(V, seed=0)
    osc, ow = O.OracleScene(sc, S, dir_switch=1024), O.OracleWeights(wts, V)
    O.set_num_threads(os.cpu_count())                 # torchrun exports OMP_NUM_THREADS=1 to its ranks
    row0 = (H // 4) * W
    n = args.cpu_rays or cpu_sample_rays(O, osc, ow, H, W, S, 5.0, row0)    # ~5 s per step
    t_rand = np.random.default_rng(0).random((n, S), dtype=np.float32)
    for _ in range(args.warmup):
        O.render_rays(osc, ow, row0, min(n, 512), t_rand[:min(n, 512)])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.render_rays(osc, ow, row0, n, t_rand)
    dt = (time.perf_counter() - t0) / args.steps
    v = n / dt
    sample = (f"{n} rays ({n // W} image rows from row {row0 // W}) of the {H}x{W} frame per step ({dt:.1f} s), S={S}, "
              f"V'={V}, OpenMP on {O.num_threads()} threads")
    print(json.dumps({
        "impl": "reference", "metric": "rays/sec", "value": v, "unit": "rays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "H": H, "W": W, "source_views": V, "samples_per_ray": S},
        "cpu_baseline": {"value": v, "unit": "rays/s", "cores": O.num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def init_dist():
    """One process per GPU (torchrun env); returns (rank, world, local, device)."""
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        # NCCL prints its version banner on fd 1 when the communicator is created: stdout carries the JSON line only
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    return rank, world, local, dev


def measure_render(args, wl_key, steps, warmup, ctx, full=True, replicas=False):
    """One render workload on the current process group.  Returns the measurement dict on rank 0, None elsewhere.
    `replicas`: every rank renders its own whole view (the round-1 weak-scaling measurement) instead of a row tile."""
    import torch
    import torch.distributed as dist
    from ucnerf_b200 import lib as L, shard, synthetic
    from ucnerf_b200.frame import FrameRenderer, HostSweep, PackedFrame, PackedWeights

    rank, world, local, dev = ctx
    wl = WORKLOADS[wl_key]
    H, W, V, S = wl["H"], wl["W"], wl["V"], wl["S"]
    n_frame = H * W
    split = world > 1 and not replicas
    peer_mode = split and args.gather == "peer"
    # the view every rank works on: one shared target view when the frame is split, one view per rank otherwise
    sc = synthetic.make_scene(H, W, V, seed=0, target_motion=0.25 * rank if (replicas and world > 1) else 0.0)
    wts = synthetic.make_weights(V, seed=0)
    pw = PackedWeights(wts, V, dev)
    host = {"dv": [torch.from_numpy(a).pin_memory() for a in sc.depth_values],
            "vol": [torch.from_numpy(a).pin_memory() for a in sc.volumes],
            "feat": torch.from_numpy(sc.img_feats).pin_memory(), "imgs": torch.from_numpy(sc.imgs).pin_memory(),
            "conf": torch.from_numpy(sc.confidence).pin_memory()}
    resident = {k: ([t.to(dev) for t in v] if isinstance(v, list) else v.to(dev)) for k, v in host.items()}
    ray0, n_tile = shard.row_tile(H, W, rank, world) if split else (0, n_frame)
    rows = (ray0 // W, (ray0 + n_tile) // W)
    # rays per launch group: the whole tile unless its staged rows would exceed ~24 GB of workspace (config 5: 124 GB)
    per_ray = S * (4 * (24 + 12 * V + 1) + 32)
    max_rays = args.max_rays or min(n_tile, max(W, int(24e9 // per_ray) // W * W))

    def pack(maps):
        fr = PackedFrame(dev)
        fr.set_target(H, W, S, sc.K[0], sc.c2ws[0], sc.w2cs[0], sc.near_fars[0, 0], sc.near_fars[0, 1],
                      dir_rot0=sc.w2cs[0], dir_rot1=sc.w2cs[1], dir_switch=1024)
        fr.set_sources(sc.w2cs[1:], sc.K[1:])
        if split:
            fr.set_row_band(*rows)                      # pack only the rows of the target-frustum maps this tile touches
        fr.pack_depth_values(maps["dv"]).pack_volumes(maps["vol"]).pack_sources(maps["imgs"], maps["feat"], maps["conf"])
        return fr

    peer = [shard.PeerFrameBuffer(H, W, dev) for _ in range(2)] if peer_mode else None
    rgb = torch.empty((n_tile, 3), device=dev)
    depth = torch.empty((n_tile,), device=dev)
    flag = torch.zeros(1, device=dev)
    ren_box, bar_ev, host_submit = {}, [], {}

    # Peer mode: the per-frame barrier runs on a side stream.  Frame i's barrier (every rank's tile is in the root's buffer
    # i&1) is needed before the root reads that buffer and before any rank's K4 of frame i+2 overwrites it -- not before
    # frame i+1 starts: taking it off the main stream stops every frame from waiting for that frame's slowest rank
    # (ranks differ by 4-5 % and jitter; the frame period becomes the ranks' average, not the per-frame maximum).
    side = torch.cuda.Stream(device=dev) if (peer_mode and args.barrier == "side") else None
    done = [None, None]
    flags2 = [torch.zeros(1, device=dev) for _ in range(2)]

    def drain():
        for d in done:
            if d is not None:
                torch.cuda.current_stream().wait_event(d)

    host_delay = float(os.environ.get("UCN_BENCH_HOST_DELAY_MS", "0")) * 1e-3   # diagnostic: does host time add to the frame?

    def step_resident(timers, seed, i=0):
        if host_delay:
            time.sleep(host_delay)
        if side is not None and done[i & 1] is not None:
            torch.cuda.current_stream().wait_event(done[i & 1])      # buffer i&1: frame i-2 is complete (and consumed)
        fr = pack(resident)
        ren = ren_box.get("r")
        if ren is None:
            ren = ren_box["r"] = FrameRenderer(fr, pw, args.mode, max_rays=max_rays)
        ren.frame = fr
        out = peer[i & 1].tile(ray0, n_tile) if peer_mode else (rgb, depth)
        ren.render(ray0, n_tile, seed=seed, out_rgb=out[0], out_depth=out[1], timers=timers)
        if split:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if side is not None:
                ev = torch.cuda.Event()
                ev.record()
                with torch.cuda.stream(side):
                    side.wait_event(ev)
                    e0.record(side)
                    dist.all_reduce(flags2[i & 1])      # frame complete on the root once this returns in stream order
                    e1.record(side)
                    d = torch.cuda.Event()
                    d.record(side)
                    done[i & 1] = d
                bar_ev.append((e0, e1))
                return
            e0.record()
            if peer_mode:
                dist.all_reduce(flag)                   # frame complete on the root once this returns in stream order
            else:
                ren_box["full"] = shard.gather_image(torch.cat([rgb, depth[:, None]], 1), H, W)
            e1.record()
            bar_ev.append((e0, e1))

    sweep = HostSweep(pw, pack, n_tile, args.mode, dev, max_rays=max_rays, ray0=ray0, frame_hw=(H, W),
                      band_keys=("dv", "vol", "conf") if split else (), peer=peer)
    def step_e2e(i, n):
        # view i from pinned host maps -> pinned host image; view i+1's maps cross PCIe while view i renders, and
        # view i+1 is submitted before view i is collected, so the GPU does not idle while the host builds a frame
        if i == 0:
            sweep.submit(0, host, host if n > 1 else None, seed=0)
        if i + 1 < n:
            sweep.submit(i + 1, host, host if i + 2 < n else None, seed=i + 1)
        return sweep.result(i)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(n):
            fn(i)
        drain()                                          # the last frames' barriers (side stream) are part of the work
        e1.record()
        submit = (time.perf_counter() - t0) * 1e3        # host time to SUBMIT the n steps (no synchronisation inside)
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = torch.tensor([e0.elapsed_time(e1), wall, submit], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        host_submit["ms"] = ms[2].item() / n
        return ms[0].item() / n, ms[1].item() / n

    # nvidia-smi is started BEFORE the warm-up: its start-up (NVML initialisation, first driver queries) has been seen to
    # stall kernel launches for ~0.4 s, which is most of a 5-step timed region (two runs at 124-130 ms per frame); its
    # samples then cover warm-up + timed region, both under load
    sampler = make_clock_sampler(local) if (full and rank == 0) else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    for i in range(warmup):
        step_resident(None, 1000 + i, i)
    do_e2e = peer_mode or not split
    if do_e2e:
        for i in range(max(2, warmup // 2)):
            step_e2e(i, max(2, warmup // 2))
    # timed region: the fused call (ucn_render_rays), as a user runs it.  The per-kernel breakdown (roofline) comes from an
    # instrumented pass of the same K steps right after it, still under the clock sampler: CUDA events between the kernels
    # cost the frame 3-4 % (no back-to-back launches), which is why they are kept out of `value`.
    bar_ev.clear()
    L.launch_count(reset=True)
    if sampler:
        sampler.mark()
    ms, wall_ms = timed(lambda i: step_resident(None, i, i), steps)
    submit_ms = host_submit.get("ms")
    launches = L.launch_count() // steps
    gather_ev = list(bar_ev)
    timers = {}
    ms_instr, _ = timed(lambda i: step_resident(timers, i, i), steps)
    clocks = sampler.stop() if (rank == 0 and sampler) else None
    e2e_ms = None
    if do_e2e:
        e2e_ms, e2e_wall = timed(lambda i: step_e2e(i, steps), steps)
        e2e_ms = max(e2e_ms, e2e_wall)                   # host work is part of the end-to-end call

    # ---- in-run checks -------------------------------------------------------------------------------------
    checks = {}
    if split:
        # the root's frame (last timed step: seed = steps-1, buffer (steps-1)&1) against the same frame rendered whole
        # on the root alone: the split must not change a single bit
        barrier()
        if rank == 0:
            got = peer[(steps - 1) & 1] if peer_mode else None
            fr = PackedFrame(dev)
            fr.set_target(H, W, S, sc.K[0], sc.c2ws[0], sc.w2cs[0], sc.near_fars[0, 0], sc.near_fars[0, 1],
                          dir_rot0=sc.w2cs[0], dir_rot1=sc.w2cs[1], dir_switch=1024)
            fr.set_sources(sc.w2cs[1:], sc.K[1:])
            fr.pack_depth_values(resident["dv"]).pack_volumes(resident["vol"]).pack_sources(resident["imgs"], resident["feat"], resident["conf"])
            whole_max = args.max_rays or min(n_frame, max(W, int(24e9 // per_ray) // W * W))
            w_rgb, w_depth = FrameRenderer(fr, pw, args.mode, max_rays=whole_max).render(seed=steps - 1)
            if peer_mode:
                same = bool(torch.equal(got.rgb, w_rgb)) and bool(torch.equal(got.depth, w_depth))
            else:
                full_img = ren_box["full"]
                same = bool(torch.equal(full_img[:, :3], w_rgb)) and bool(torch.equal(full_img[:, 3], w_depth))
            checks["sharded_frame_bit_identical_to_1gpu"] = same
            if not same:
                raise SystemExit("bench: the row-tile-sharded frame differs from the single-GPU frame")
            if peer_mode and do_e2e:
                h_rgb, h_depth = HostSweep.split_frame(sweep.result(steps - 1))
                checks["e2e_host_frame_bit_identical_to_1gpu"] = bool(torch.equal(h_rgb, w_rgb.cpu())) and \
                    bool(torch.equal(h_depth, w_depth.cpu()))
            del w_rgb, w_depth, fr
        barrier()

    kern = {k: sum(a.elapsed_time(b) for a, b in v) / steps for k, v in timers.items()}   # ms per step, this rank
    # GPU timeline of the instrumented pass on this rank: what lies BETWEEN the kernels of a frame and between frames
    # (event to event, so idle time shows up): diagnostic for the multi-GPU frame period
    timeline = None
    order = [k for k in ("raygen", "dir", "gather", "mlp", "composite") if k in timers]
    if order and all(len(timers[k]) == steps for k in order) and steps > 1:
        gaps = {}
        for a, b in zip(order, order[1:]):
            gaps[f"{a}->{b}"] = sum(timers[a][i][1].elapsed_time(timers[b][i][0]) for i in range(steps)) / steps
        gaps["frame_end->next_frame_first_kernel"] = sum(timers[order[-1]][i][1].elapsed_time(timers[order[0]][i + 1][0])
                                                         for i in range(steps - 1)) / (steps - 1)
        timeline = {"instrumented_ms_per_step": ms_instr, "gaps_ms": {k: round(v, 4) for k, v in gaps.items()}}
    per_rank = None
    if world > 1:
        mine = torch.tensor([sum(kern.values())], device=dev, dtype=torch.float64)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [round(float(t.item()), 4) for t in allr]
    n_launch = {k: len(v) // steps for k, v in timers.items()}
    h2d_rank = sweep.h2d_bytes(host)
    tot = torch.tensor([float(h2d_rank)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot)
    res = None
    if rank == 0:
        pk = peaks()
        F = 24 + 12 * V + 1
        tile_samples = n_tile * S
        mlp_ms_per_launch = kern["mlp"] / n_launch["mlp"]
        flops_per_launch = FLOP_PER_SAMPLE(V) * tile_samples / n_launch["mlp"]
        achieved_tf = flops_per_launch / (mlp_ms_per_launch * 1e-3) / 1e12
        peak_tf = pk["bf16_tflops_sustained"]
        gather_gbs = (4 * F + 16) * tile_samples / (kern["gather"] * 1e-3) / 1e9
        comp_gbs = (20 * S + 16) * n_tile / (kern["composite"] * 1e-3) / 1e9
        ray_gbs = (4 * S + 36) * n_tile / (kern["raygen"] * 1e-3) / 1e9
        total_rays = n_frame * (world if (replicas and world > 1) else 1)
        x3 = "mlp_x3p_kernel" if L.load().ucn_mlp_x3_variant(0) == 2 and V <= 7 else "mlp_x3_kernel"
        kname = {"bf16": x3, "bf16_fast": "mlp_tc_kernel", "fp32": "mlp_fp32_kernel"}[args.mode]
        if split:
            par = (f"one frame as {world} contiguous row tiles (shard.row_tile), "
                   + ("K4 peer-stores each tile into the root's frame buffer over NVLink (16 B/ray), one 4-byte all-reduce per "
                      "frame as barrier" + (" on a side stream (it gates the root's read-out and frame i+2's stores into the "
                                            "same buffer, not the start of frame i+1)" if args.barrier == "side" else " in line")
                      + ", root reads the frame out" if peer_mode else "ncclAllGather of [R/P,4] tiles")
                   + "; every rank packs only its row band of the target-frustum maps")
        elif world > 1:
            par = f"replicas x{world}: one whole view per rank, no data-path collective"
        else:
            par = "single GPU, whole frame"
        res = {
            "value": total_rays / (ms * 1e-3), "ms_per_step": ms, "wall_ms_per_step": wall_ms, "host_submit_ms_per_step": submit_ms,
            "scaling": "strong" if split else "weak",
            "config": {"workload": wl["name"], "H": H, "W": W, "source_views": V, "samples_per_ray": S,
                       "mlp_mode": args.mode, "rays_per_launch_group": max_rays, "rays_per_rank": n_tile,
                       "l2": f"inputs larger than L2: packed maps {sum(t.numel() * 4 for _, _, t in HostSweep._walk(host)) / 1e6:.0f} MB + "
                             f"staged rows {tile_samples * (4 * F + 32) / 1e9:.1f} GB per step per rank",
                       "parallelism": par},
            "e2e": None if e2e_ms is None else {
                "value": total_rays / (e2e_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": int(tot.item()),
                "d2h_bytes_per_step": n_frame * 16 * (world if (replicas and world > 1) else 1), "ms_per_step": e2e_ms,
                "h2d_bytes_per_rank": h2d_rank, "h2d_GBps_per_rank": h2d_rank / (e2e_ms * 1e-3) / 1e9},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"kernel": kname, "bound": "tensor", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved_tf / peak_tf, "traffic": ncu_traffic(kname, wl_key) if world == 1 else None,
                         "peak_source": pk["_source"] + " (bf16_tflops_sustained)",
                         "flops_per_launch": flops_per_launch, "ms_per_launch": mlp_ms_per_launch,
                         "measured_on": "rank 0" + (f", its {n_tile}-ray tile" if split else "")},
            "kernels_ms_per_step": kern,
            "kernels_from": f"instrumented pass of the same {steps} steps right after the timed region (CUDA events around every "
                            f"kernel: {ms_instr:.2f} ms per step); the timed region runs the fused ucn_render_rays call",
            "hbm_kernels": {"gather_GBps": gather_gbs, "gather_frac": gather_gbs / pk["hbm_gbs"],
                            "gather_traffic": ncu_traffic("gather_kernel", wl_key) if world == 1 else None,
                            "composite_GBps": comp_gbs, "composite_frac": comp_gbs / pk["hbm_gbs"],
                            "raygen_GBps": ray_gbs, "raygen_frac": ray_gbs / pk["hbm_gbs"],
                            "peak_GBps": pk["hbm_gbs"]},
        }
        if split:
            res["gather_ms"] = sum(a.elapsed_time(b) for a, b in gather_ev) / max(1, len(gather_ev))
            res["gather"] = ("peer-store: the image gather is K4's own stores; gather_ms is the stream-ordered barrier "
                             "(4-byte all-reduce) on rank 0, which includes waiting for the slowest rank") if peer_mode \
                else "ncclAllGather of the [R/P,4] tiles (shard.gather_image)"
            res["checks"] = checks
            res["per_rank_kernel_ms"] = per_rank            # sum of the rank's kernels per frame (instrumented pass)
            res["timeline"] = timeline
        # ---- parity of the timed configuration: a band of the same frame on the GPU vs the CPU oracle, same jitter ----
        from oracle import oracle as O
        osc, ow = O.OracleScene(sc, S, dir_switch=1024), O.OracleWeights(wts, V)
        O.set_num_threads(os.cpu_count())
        row0 = (H // 4) * W
        if full and world == 1 and not args.no_cpu_baseline:
            n = args.cpu_rays or cpu_sample_rays(O, osc, ow, H, W, S, 15.0, row0)
        else:
            n = min(4 * W, n_frame - row0)
        tr = np.random.default_rng(0).random((n, S), dtype=np.float32)
        t0 = time.perf_counter()
        rgb_o, depth_o = O.render_rays(osc, ow, row0, n, tr)
        dt = time.perf_counter() - t0
        if full and world == 1 and not args.no_cpu_baseline:
            res["cpu_baseline"] = {"value": n / dt, "unit": "rays/s", "cores": O.num_threads(), "kind": "port",
                                   "sample": f"{n} rays ({n // W} image rows from row {row0 // W}) of the same {H}x{W} frame ({dt:.1f} s)"}
        fr = PackedFrame(dev)
        fr.set_target(H, W, S, sc.K[0], sc.c2ws[0], sc.w2cs[0], sc.near_fars[0, 0], sc.near_fars[0, 1],
                      dir_rot0=sc.w2cs[0], dir_rot1=sc.w2cs[1], dir_switch=1024)
        fr.set_sources(sc.w2cs[1:], sc.K[1:])
        fr.pack_depth_values(resident["dv"]).pack_volumes(resident["vol"]).pack_sources(resident["imgs"], resident["feat"], resident["conf"])
        g_rgb, g_depth = FrameRenderer(fr, pw, args.mode, max_rays=min(n, max_rays)).render(
            row0, n, t_rand=torch.from_numpy(tr).to(dev))
        gate = 1e-4 if args.mode == "fp32" else 1e-3
        res["parity"] = {"rays": n, "max_abs_rgb": float(np.abs(g_rgb.cpu().numpy() - rgb_o).max()),
                         "max_abs_depth": float(np.abs(g_depth.cpu().numpy() - depth_o).max()), "gate": gate,
                         "against": "oracle/ (CPU restatement pinned to the reference's outputs), same rays, same jitter",
                         "mode": args.mode}
        res["parity"]["ok"] = res["parity"]["max_abs_rgb"] <= gate
    if peer:
        for b in peer:
            b.close()
    barrier()
    return res


def gpu_eager_baseline(wl_key, dev, n_chunks=8):
    """PyTorch-eager restatement of the reference's chunk loop on this GPU (baseline/torch_eager.py), checked against the
    oracle on its first chunk, `n_chunks` chunks timed from the middle of the frame and extrapolated by rate."""
    import torch
    from baseline.torch_eager import EagerPath, time_chunks
    from oracle import oracle as O
    from ucnerf_b200 import synthetic
    wl = WORKLOADS[wl_key]
    H, W, V, S = wl["H"], wl["W"], wl["V"], wl["S"]
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sc, wts = synthetic.make_scene(H, W, V, seed=0), synthetic.make_weights(V, seed=0)
    path = EagerPath(sc, wts, S, dev, chunk=1024)
    tr = np.random.default_rng(0).random((1024, S), dtype=np.float32)
    with torch.no_grad():
        rgb, depth = path.render_chunk(0, torch.from_numpy(tr).to(dev), first_chunk=True)
    O.set_num_threads(os.cpu_count())
    rgb_o, depth_o = O.render_rays(O.OracleScene(sc, S, dir_switch=1024), O.OracleWeights(wts, V), 0, 1024, tr)
    err = float(np.abs(rgb.cpu().numpy() - rgb_o).max())
    mid = (H * W // 1024) // 2
    ids = list(range(mid, mid + n_chunks))
    s_sync = time_chunks(path, ids, sync_per_chunk=True)
    s_nosync = time_chunks(path, ids, sync_per_chunk=False)
    return {"value": 1024 / s_sync, "unit": "rays/s", "ms_per_chunk": s_sync * 1e3,
            "value_without_per_chunk_cpu": 1024 / s_nosync, "ms_per_chunk_without_per_chunk_cpu": s_nosync * 1e3,
            "chunk": 1024, "chunks_timed": n_chunks, "extrapolated": f"{n_chunks} chunks of 1024 rays from the middle of the "
            f"{H}x{W} frame, rate taken as the frame's", "max_abs_rgb_vs_oracle_first_chunk": err, "gate": 1e-4,
            "what": "torch-eager restatement of train.py:251-275 (build_rays_test + rendering per 1024-ray chunk: torch.sort, "
                    f"{4 + 2 * V} F.grid_sample, 13 nn.Linear fp32 with TF32 off, cumprod; rgb.cpu()/depth.cpu() per chunk), "
                    "same GPU; not in any ratio"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default=os.environ.get("UCNERF_BENCH_MODE", "bf16"), choices=["bf16", "bf16_fast", "fp32"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--max-rays", type=int, default=0, help="rays per launch group (0 = auto)")
    ap.add_argument("--cpu-rays", type=int, default=0, help="bounded CPU sample (rays); 0 = sized for ~15 s (ours) / ~5 s per step (reference)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N>1: image gather by K4 peer stores into the root's frame buffer (default) or ncclAllGather")
    ap.add_argument("--barrier", default="inline", choices=("side", "inline"),
                    help="N>1 peer mode: per-frame barrier in line on the render stream (default) or on a side stream "
                         "(1-2 %% faster at N = 8 when it works, but a resident NCCL kernel can keep one CTA of the persistent "
                         "MLP kernel off its SM: one 2-GPU run fell to 130 ms per frame -- experimental)")
    ap.add_argument("--replicas", action="store_true", help="N>1: one whole view per rank (round-1 weak-scaling measurement)")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra blocks (config 5 full box, config 4, replicas, eager)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.workload == "c4":
        import bench_train
        return bench_train.run_reference(args) if args.impl == "reference" else bench_train.run(args, ClockSampler, peaks)
    if args.impl == "reference":
        return run_reference(args, wl)

    import torch
    import torch.distributed as dist
    from ucnerf_b200 import lib as L
    ctx = init_dist()
    rank, world, local, dev = ctx
    L.check(L.load().ucn_check_device(), "ucn_check_device")
    res = measure_render(args, args.workload, args.steps, args.warmup, ctx, full=True, replicas=args.replicas)
    extra = {}
    if not args.no_extras:
        t_extra = time.perf_counter()
        if world > 1 and not args.replicas:
            r = measure_render(args, args.workload, 3, 3, ctx, full=False, replicas=True)
            if r:
                extra["replicas"] = {k: r[k] for k in ("value", "ms_per_step", "scaling", "e2e", "gpu_launches")}
                extra["replicas"]["parallelism"] = r["config"]["parallelism"]
        if args.workload != "c5":
            r = measure_render(args, "c5", 3, 3, ctx, full=False)
            if r:
                extra["c5_fullbox"] = {k: r.get(k) for k in ("value", "ms_per_step", "scaling", "config", "e2e", "gpu_launches",
                                                              "roofline", "hbm_kernels", "kernels_ms_per_step", "gather_ms",
                                                              "checks", "parity")}
        import bench_train
        targs = argparse.Namespace(**{**vars(args), "steps": 20, "warmup": 3, "no_cpu_baseline": True})
        r = bench_train.measure(targs, ClockSampler, peaks, ctx, full=False)
        if r:
            extra["train_c4"] = {k: r.get(k) for k in ("value", "ms_per_step", "scaling", "config", "e2e", "gpu_launches",
                                                        "kernels_ms_per_step", "roofline", "loss_first_last", "parity")}
        if world == 1 and rank == 0:
            try:
                extra["gpu_eager_baseline"] = gpu_eager_baseline(args.workload, dev)
            except Exception as e:                       # a baseline leg must never take the bench line down
                extra["gpu_eager_baseline"] = {"unavailable": f"{type(e).__name__}: {e}"}
        extra["extras_wall_s"] = time.perf_counter() - t_extra
    if rank == 0:
        line = {"metric": "rays/sec", "value": res["value"], "unit": "rays/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
                "scaling": res["scaling"], "vs_baseline": None, "dtype": "f32" if args.mode == "fp32" else "bf16",
                "data": "synthetic"}
        for k in ("config", "e2e", "gpu_launches", "clocks", "roofline", "kernels_ms_per_step", "kernels_from", "hbm_kernels",
                  "wall_ms_per_step", "host_submit_ms_per_step", "gather_ms", "gather", "per_rank_kernel_ms", "timeline", "checks", "parity",
                  "cpu_baseline"):
            if k in res and res[k] is not None:
                line[k] = res[k]
        if "gpu_eager_baseline" in extra:
            line["gpu_eager_baseline"] = extra.pop("gpu_eager_baseline")
        if extra:
            line["extra"] = extra
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
