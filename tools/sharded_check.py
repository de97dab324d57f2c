"""N-GPU check (run under torchrun): ray-tile sharded render of one 320x256 target view over NCCL == the single-GPU
render, bit for bit, and within the bf16 gate of the CPU oracle on a band of rows.  Also times the sharded frame
(strong scaling: fixed frame split over N ranks, one all-gather of 16 B/ray at the end)."""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch, torch.distributed as dist
from ucnerf_b200 import shard, synthetic
from ucnerf_b200.frame import FrameRenderer, PackedFrame, PackedWeights

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
H, W, V, S = 256, 320, 3, 126
sc = synthetic.make_scene(H, W, V, seed=5)
pw = PackedWeights(synthetic.make_weights(V, seed=5), V, dev)
fr = PackedFrame.from_scene(sc, S, dev, dir_switch=1024)
t_rand = torch.from_numpy(np.random.default_rng(5).random((H * W, S), dtype=np.float32)).to(dev)
ren = FrameRenderer(fr, pw, "bf16")

def render_rays(ray0, n):
    return ren.render(ray0, n, t_rand=t_rand[ray0:ray0 + n])

rgb, depth = shard.render_sharded(render_rays, H, W, row_align=8)
full_rgb, full_depth = ren.render(0, H * W, t_rand=t_rand)
torch.cuda.synchronize()
same = bool(torch.equal(rgb, full_rgb) and torch.equal(depth, full_depth))
for _ in range(3):
    shard.render_sharded(render_rays, H, W, row_align=8)
dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    shard.render_sharded(render_rays, H, W, row_align=8)
e1.record(); dist.barrier(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / 10], device=dev)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
ok = torch.tensor([int(same)], device=dev)
dist.all_reduce(ok, op=dist.ReduceOp.MIN)

# the same frame with K4 storing every rank's tile straight into rank 0's frame buffer over NVLink (no all-gather)
buf = shard.PeerFrameBuffer(H, W, dev, root=0)

def render_into(ray0, n, out_rgb, out_depth):
    ren.render(ray0, n, t_rand=t_rand[ray0:ray0 + n], out_rgb=out_rgb, out_depth=out_depth)

res = shard.render_sharded_p2p(render_into, buf, row_align=8)
p2p_same = True if res is None else bool(torch.equal(res[0], full_rgb) and torch.equal(res[1], full_depth))
for _ in range(3):
    shard.render_sharded_p2p(render_into, buf, row_align=8)
dist.barrier(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    shard.render_sharded_p2p(render_into, buf, row_align=8)
p2p_ms = torch.tensor([(time.perf_counter() - t0) * 100.0], device=dev)       # wall ms per frame (ends with a barrier)
dist.all_reduce(p2p_ms, op=dist.ReduceOp.MAX)
t0 = time.perf_counter()
for _ in range(10):
    shard.render_sharded(render_rays, H, W, row_align=8)
    torch.cuda.synchronize()
ag_ms = torch.tensor([(time.perf_counter() - t0) * 100.0], device=dev)
dist.all_reduce(ag_ms, op=dist.ReduceOp.MAX)
if rank == 0:
    from oracle import oracle as O
    O.set_num_threads(os.cpu_count())
    osc, ow = O.OracleScene(sc, S, dir_switch=1024), O.OracleWeights(synthetic.make_weights(V, seed=5), V)
    r0, n = 100 * W, 8 * W
    rgb_o, _ = O.render_rays(osc, ow, r0, n, t_rand[r0:r0 + n].cpu().numpy())
    err = float(np.abs(rgb[r0:r0 + n].cpu().numpy() - rgb_o).max())
    print(json.dumps({"check": "sharded render == single-GPU render (bit-identical on every rank)", "n_gpus": world,
                      "bit_identical": bool(ok.item()), "max_abs_rgb_err_vs_oracle": err, "gate": 1e-3,
                      "frame": f"{W}x{H}, S={S}, V'={V}", "sharded_frame_ms": ms.item(),
                      "rays_per_s": H * W / (ms.item() * 1e-3),
                      "p2p_epilogue": {"what": "K4 stores each tile into rank 0's frame buffer over NVLink (CUDA IPC), one barrier",
                                       "bit_identical_on_root": p2p_same, "wall_ms_per_frame": p2p_ms.item(),
                                       "allgather_wall_ms_per_frame": ag_ms.item()}}))
    assert ok.item() == 1 and err <= 1e-3 and p2p_same
buf.close()
dist.destroy_process_group()
