"""CPU, world_size 2 over gloo: ray-tile sharding + final image gather reproduce the single-rank frame
bit-identically (rays are independent).  The per-rank "renderer" here is the CPU oracle -- the product
kernels need a GPU -- so this covers the host-side N>1 logic: partition, uneven tiles, gather order."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ucnerf_b200 import shard


def test_row_tile_partition():
    for H, W, world, align in ((256, 320, 8, 1), (30, 7, 4, 1), (512, 640, 8, 8), (17, 5, 3, 4), (4, 4, 8, 1)):
        spans = [shard.row_tile(H, W, p, world, align) for p in range(world)]
        assert spans[0][0] == 0
        for (a0, an), (b0, _) in zip(spans, spans[1:]):
            assert a0 + an == b0
        assert spans[-1][0] + spans[-1][1] == H * W
        assert all(n % W == 0 for _, n in spans)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, H, W, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from helpers import load_case
    from oracle import oracle as O
    c = load_case("small_v3")
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
    ow = O.OracleWeights(c["weights"], c["V"])

    def render_rays(ray0, n):
        rgb, depth = O.render_rays(osc, ow, ray0, n, c["t_rand"][ray0:ray0 + n])
        return torch.from_numpy(rgb), torch.from_numpy(depth)
    rgb, depth = shard.render_sharded(render_rays, H, W, row_align=5)      # uneven tiles on purpose
    if rank == 0:
        np.savez(os.path.join(out_dir, "out.npz"), rgb=rgb.numpy(), depth=depth.numpy())
    dist.destroy_process_group()


def test_sharded_render_matches_single_rank(tmp_path):
    from helpers import load_case
    from oracle import oracle as O
    c = load_case("small_v3")
    H, W = c["H"], c["W"]
    mp.spawn(_worker, args=(2, _free_port(), H, W, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "out.npz")
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
    ow = O.OracleWeights(c["weights"], c["V"])
    rgb, depth = O.render_rays(osc, ow, 0, H * W, c["t_rand"])
    assert np.array_equal(got["rgb"], rgb) and np.array_equal(got["depth"], depth)
    assert np.abs(rgb - c["g"]["rgb"]).max() <= 1e-4


def test_stage_band_covers_every_row_a_tile_touches():
    """frame.stage_band: rows of a stage map that the rays of image rows [r0, r1) read through
    grid_sample(align_corners=False, border) when the volume is pixel-aligned with the target view
    (utils/utils.py:852-857; ATen GridSampler.h unnormalize: ((g+1)*h - 1)/2 with g = 2*y/(H-1) - 1)."""
    import numpy as np
    from ucnerf_b200.frame import stage_band
    from ucnerf_b200.shard import row_tile
    for H, W in ((512, 640), (1024, 1280), (256, 320), (16, 20), (8, 12)):
        for world in (1, 2, 3, 4, 8):
            for rank in range(world):
                ray0, n = row_tile(H, W, rank, world)
                if n == 0:
                    continue
                r0, r1 = ray0 // W, (ray0 + n) // W
                for ds in (4, 2, 1):
                    h = H // ds
                    a, b = stage_band(H, r0, r1, ds)
                    assert 0 <= a < b <= h
                    y = np.arange(r0, r1, dtype=np.float64)
                    for jitter in (-1e-3, 0.0, 1e-3):                 # projection round-off moves the row by << 1e-3
                        x = np.clip((y + jitter) / (H - 1) * h - 0.5, 0, h - 1)
                        lo, hi = np.floor(x).astype(int), np.minimum(np.floor(x).astype(int) + 1, h - 1)
                        assert lo.min() >= a and hi.max() < b, (H, world, rank, ds, a, b, lo.min(), hi.max())
                    # near/far lookup of K1: pixel row // ds (utils/utils.py:660-683)
                    assert r0 // ds >= a and (r1 - 1) // ds < b
