"""A/B of the whole-frame render (config 2 by default): fused ucn_render_rays vs the per-kernel staged path, both K3 variants."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from ucnerf_b200 import lib as L, synthetic
from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights
H, W, V, S = (int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (512, 640, 3, 126)))
sc = synthetic.make_scene(H, W, V, seed=0)
pw = PackedWeights(synthetic.make_weights(V, seed=0), V, "cuda")
fr = PackedFrame.from_scene(sc, S, "cuda", dir_switch=1024)
lib = L.load()
per_ray = S * (4 * (24 + 12 * V + 1) + 32)
max_rays = min(H * W, max(W, int(24e9 // per_ray) // W * W))
for variant in (2, 1, 2, 1):
    lib.ucn_mlp_x3_variant(variant)
    ren = FrameRenderer(fr, pw, "bf16", max_rays=max_rays)
    for timers in (None, {}):
        for _ in range(3):
            ren.render(seed=1, timers={} if timers is not None else None)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t = {} if timers is not None else None
        e0.record()
        for i in range(5):
            ren.render(seed=i, timers=t)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        k = {} if t is None else {n: sum(a.elapsed_time(b) for a, b in v) / 5 for n, v in t.items()}
        print(f"variant {variant} {'staged' if t is not None else 'fused '}: {ms:.2f} ms/frame  {H * W / ms / 1e3:.2f} Mrays/s  " +
              " ".join(f"{n}={v:.2f}" for n, v in k.items()), flush=True)
    del ren
