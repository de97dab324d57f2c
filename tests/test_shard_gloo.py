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
