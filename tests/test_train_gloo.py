"""CPU, world_size 2 over gloo: the data-parallel gradient exchange of the training step.
Each rank back-propagates its own ray batch through the CPU restatement of the path (oracle/), the flat gradient
bucket is all-reduced (sum / P), and every rank must end up with the mean of the two single-rank gradients -- the
three parameters forward() never touches (Q9) stay without a gradient and contribute zeros to the bucket."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ucnerf_b200 import train_path as TP


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_grads(rank):
    """Parameter gradients of one rank's batch (rows rank*64 .. rank*64+63 of the golden case), CPU oracle ops."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from helpers import load_grad_case
    from oracle import oracle as O
    from oracle.train_ops import OracleOps
    c = load_grad_case("grad_small_v3")
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
    py, px = O.chunk_pixels(c["H"], c["W"])
    sl = slice(rank * 64, rank * 64 + 64)
    r = O.build_rays_test(osc, py[sl], px[sl], c["t_rand"][sl])
    feat = O.gen_pts_feats(osc, r["pts"], r["ndc_all"])
    ang = O.dir_feature(osc, r["rays_d"])
    R, S, F = feat.shape
    T = torch.from_numpy
    P = {n: T(np.ascontiguousarray(c["weights"]["nerf." + n])) for n in TP.PARAM_NAMES}
    raw, saved = TP.mlp_forward_train(OracleOps(), P, T(r["ndc"]["ndc"].reshape(-1, 3).copy()), T(ang.copy()),
                                      T(feat.reshape(-1, F).copy()), S)
    d_raw = O.raw2outputs_bwd(raw.numpy().reshape(R, S, 4), r["z"], c["g_rgb"][sl], c["g_depth"][sl])
    G = {n: torch.zeros_like(p) for n, p in P.items()}
    TP.mlp_backward(OracleOps(), P, saved, T(d_raw.reshape(-1, 4)), G)
    return G


def _params_with(G):
    """A parameter list shaped like UCNeRF's: the 30 live tensors plus one never-used layer without a gradient."""
    params = []
    for n in TP.PARAM_NAMES:
        p = torch.nn.Parameter(torch.zeros_like(G[n]))
        p.grad = G[n].clone()
        params.append(p)
    params.insert(5, torch.nn.Parameter(torch.zeros(7, 3)))          # grad stays None
    return params


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    params = _params_with(_rank_grads(rank))
    bucket = TP.allreduce_grads(params, world)
    assert bucket.numel() == sum(p.numel() for p in params)
    assert params[5].grad is None
    np.savez(os.path.join(out_dir, f"g{rank}.npz"), **{str(i): p.grad.numpy() for i, p in enumerate(params) if p.grad is not None})
    dist.destroy_process_group()


def test_gradient_bucket_allreduce(tmp_path):
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    g0, g1 = _rank_grads(0), _rank_grads(1)
    got0, got1 = np.load(tmp_path / "g0.npz"), np.load(tmp_path / "g1.npz")
    names = list(TP.PARAM_NAMES)
    names.insert(5, None)
    for i, n in enumerate(names):
        if n is None:
            assert str(i) not in got0.files
            continue
        want = (g0[n].numpy().astype(np.float64) + g1[n].numpy()) / 2
        assert np.array_equal(got0[str(i)], got1[str(i)]), n                      # both ranks hold the same mean
        assert np.abs(got0[str(i)] - want).max() <= 1e-6 * max(1.0, np.abs(want).max()), n


def test_single_rank_is_identity():
    G = {n: torch.randn(3, 4) for n in TP.PARAM_NAMES}
    params = _params_with(G)
    TP.allreduce_grads(params, 1)
    for p, n in zip([p for i, p in enumerate(params) if i != 5], TP.PARAM_NAMES):
        assert torch.equal(p.grad, G[n])


def test_bucket_packs_views_of_a_flat_gradient_buffer():
    """MlpFn.backward hands out parameter gradients as views of ONE padded buffer (offsets rounded up to four floats);
    the bucket must hold their plain concatenation (no padding), zeros for a parameter without a gradient."""
    shapes = [(3, 5), (7,), (2, 2, 3), (1,)]
    offs, total = [], 0
    for sh in shapes:
        offs.append(total)
        total += (int(np.prod(sh)) + 3) & ~3
    flat = torch.arange(total, dtype=torch.float32) + 1.0
    params = []
    for sh, o in zip(shapes, offs):
        p = torch.nn.Parameter(torch.zeros(sh))
        p.grad = flat[o:o + int(np.prod(sh))].view(sh)
        params.append(p)
    params.insert(2, torch.nn.Parameter(torch.zeros(4)))               # never used: no gradient
    bucket = TP.allreduce_grads(params, 1)
    want = torch.cat([p.grad.reshape(-1) if p.grad is not None else torch.zeros(p.numel()) for p in params])
    assert torch.equal(bucket, want) and bucket.numel() == sum(p.numel() for p in params)
    again = TP.allreduce_grads(params, 1, bucket)
    assert again is bucket and torch.equal(bucket, want)               # the bucket is reused across steps
