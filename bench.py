#!/usr/bin/env python
"""bench.py -- rays/s of the UC-NeRF per-ray rendering hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode bf16|fp32]
                    [--workload c2|c1|c5|c4] [--max-rays M]

A "step" renders one whole target view of the workload: K0 pack of that view's maps + K1 raygen/sampling
+ K2 projection/gathers + K3 PE+MLP + K4 compositing.  `value` = rays of all ranks / max-over-ranks device
time with the maps already in HBM; `e2e` = same call with HOST (pinned) maps copied in and rgb/depth
copied out inside the timed region.  N>1 (torchrun): one process per GPU, every rank renders its own target
view of the sweep (weak scaling, no data-path collective).  `--impl reference` times the CPU port of the
reference's algorithm (oracle/, pinned bit-exact to the reference) on the host cores.
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
        self.idx, self.proc, self.path = gpu_index, None, f"/tmp/ucnerf_clocks_{os.getpid()}.csv"

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
        for line in open(self.path):
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
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "H": H, "W": W, "source_views": V, "samples_per_ray": S},
        "cpu_baseline": {"value": v, "unit": "rays/s", "cores": O.num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


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
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.workload == "c4":
        import bench_train
        return bench_train.run_reference(args) if args.impl == "reference" else bench_train.run(args, ClockSampler, peaks)
    if args.impl == "reference":
        return run_reference(args, wl)

    import torch
    import torch.distributed as dist
    from ucnerf_b200 import lib as L, synthetic
    from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
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
    L.check(L.load().ucn_check_device(), "ucn_check_device")

    H, W, V, S = wl["H"], wl["W"], wl["V"], wl["S"]
    n_rays = H * W
    # every rank renders its own target view of the sweep (weak scaling): same maps, different target pose
    sc = synthetic.make_scene(H, W, V, seed=0, target_motion=0.25 * rank)
    wts = synthetic.make_weights(V, seed=0)
    pw = PackedWeights(wts, V, dev)
    host = {"dv": [torch.from_numpy(a).pin_memory() for a in sc.depth_values],
            "vol": [torch.from_numpy(a).pin_memory() for a in sc.volumes],
            "feat": torch.from_numpy(sc.img_feats).pin_memory(), "imgs": torch.from_numpy(sc.imgs).pin_memory(),
            "conf": torch.from_numpy(sc.confidence).pin_memory()}
    h2d_bytes = sum(t.numel() * 4 for t in host["dv"] + host["vol"] + [host["feat"], host["imgs"], host["conf"]])
    resident = {k: ([t.to(dev) for t in v] if isinstance(v, list) else v.to(dev)) for k, v in host.items()}
    # rays per launch group: whole frame unless its staged rows would exceed ~24 GB of workspace (config 5: 124 GB)
    per_ray = S * (4 * (24 + 12 * V + 1) + 32)
    max_rays = args.max_rays or min(n_rays, max(W, int(24e9 // per_ray) // W * W))

    def pack(maps):
        fr = PackedFrame(dev)
        fr.set_target(H, W, S, sc.K[0], sc.c2ws[0], sc.w2cs[0], sc.near_fars[0, 0], sc.near_fars[0, 1],
                      dir_rot0=sc.w2cs[0], dir_rot1=sc.w2cs[1], dir_switch=1024)
        fr.set_sources(sc.w2cs[1:], sc.K[1:])
        fr.pack_depth_values(maps["dv"]).pack_volumes(maps["vol"]).pack_sources(maps["imgs"], maps["feat"], maps["conf"])
        return fr

    ren_box = {}
    rgb = torch.empty((n_rays, 3), device=dev)
    depth = torch.empty((n_rays,), device=dev)

    def step_resident(timers, seed):
        fr = pack(resident)
        ren = ren_box.get("r")
        if ren is None:
            ren = ren_box["r"] = FrameRenderer(fr, pw, args.mode, max_rays=max_rays)
        ren.frame = fr
        ren.render(seed=seed, out_rgb=rgb, out_depth=depth, timers=timers)

    from ucnerf_b200.frame import HostSweep
    sweep = HostSweep(pw, pack, n_rays, args.mode, dev, max_rays=max_rays)

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

    for i in range(args.warmup):
        step_resident({}, 1000 + i)
    for i in range(max(2, args.warmup // 2)):
        step_e2e(i, max(2, args.warmup // 2))
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    timers = {}
    L.launch_count(reset=True)
    ms, wall_ms = timed(lambda i: step_resident(timers, i), args.steps)
    launches = L.launch_count() // args.steps
    clocks = sampler.stop() if rank == 0 else None
    e2e_ms, e2e_wall = timed(lambda i: step_e2e(i, args.steps), args.steps)
    e2e_ms = max(e2e_ms, e2e_wall)                       # host work is part of the end-to-end call

    kern = {k: sum(a.elapsed_time(b) for a, b in v) / args.steps for k, v in timers.items()}   # ms per step
    n_launch = {k: len(v) // args.steps for k, v in timers.items()}
    pk = peaks()
    n_samples = n_rays * S
    mlp_ms_per_launch = kern["mlp"] / n_launch["mlp"]
    flops_per_launch = FLOP_PER_SAMPLE(V) * n_samples / n_launch["mlp"]
    achieved_tf = flops_per_launch / (mlp_ms_per_launch * 1e-3) / 1e12
    peak_tf = pk["bf16_tflops_sustained"]
    F = 24 + 12 * V + 1
    gather_gbs = (4 * F + 16) * n_samples / (kern["gather"] * 1e-3) / 1e9
    comp_gbs = (20 * S + 16) * n_rays / (kern["composite"] * 1e-3) / 1e9

    if rank == 0:
        line = {
            "metric": "rays/sec", "value": world * n_rays / (ms * 1e-3), "unit": "rays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if args.mode == "fp32" else "bf16",
            "data": "synthetic",
            "config": {"workload": wl["name"], "H": H, "W": W, "source_views": V, "samples_per_ray": S,
                       "mlp_mode": args.mode, "rays_per_launch_group": max_rays,
                       "l2": f"inputs larger than L2: packed maps {h2d_bytes / 1e6:.0f} MB + staged rows "
                             f"{n_samples * (4 * F + 32) / 1e9:.1f} GB per step",
                       "parallelism": f"ray/view sharding x{world}, no data-path collective"},
            "e2e": {"value": world * n_rays / (e2e_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": n_rays * 16, "ms_per_step": e2e_ms},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"kernel": {"bf16": "mlp_x3_kernel", "bf16_fast": "mlp_tc_kernel", "fp32": "mlp_fp32_kernel"}[args.mode], "bound": "tensor",
                         "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
                         "traffic": ncu_traffic({"bf16": "mlp_x3_kernel", "bf16_fast": "mlp_tc_kernel", "fp32": "mlp_fp32_kernel"}[args.mode], args.workload),
                         "peak_source": pk["_source"] + " (bf16_tflops_sustained)",
                         "flops_per_launch": flops_per_launch, "ms_per_launch": mlp_ms_per_launch},
            "kernels_ms_per_step": kern,
            "hbm_kernels": {"gather_GBps": gather_gbs, "gather_frac": gather_gbs / pk["hbm_gbs"],
                            "gather_traffic": ncu_traffic("gather_kernel", args.workload),
                            "composite_GBps": comp_gbs, "composite_frac": comp_gbs / pk["hbm_gbs"],
                            "peak_GBps": pk["hbm_gbs"]},
            "wall_ms_per_step": wall_ms,
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import oracle as O
            osc, ow = O.OracleScene(sc, S, dir_switch=1024), O.OracleWeights(wts, V)
            O.set_num_threads(os.cpu_count())
            row0 = (H // 4) * W
            n = args.cpu_rays or cpu_sample_rays(O, osc, ow, H, W, S, 15.0, row0)
            tr = np.random.default_rng(0).random((n, S), dtype=np.float32)
            t0 = time.perf_counter()
            rgb_o, _ = O.render_rays(osc, ow, row0, n, tr)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt, "unit": "rays/s", "cores": O.num_threads(), "kind": "port",
                                    "sample": f"{n} rays ({n // W} image rows from row {row0 // W}) of the same {H}x{W} frame ({dt:.1f} s)"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
