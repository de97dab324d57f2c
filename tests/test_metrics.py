"""Evaluation metrics.  CPU: the numpy restatement against values produced by the reference's
own functions (tests/golden/metrics.npz).  GPU (-m gpu): the CUDA path through the C-ABI against the same goldens and
the restatement; a1..a3 are pixel counts and must match exactly, the fp32 means within 1e-5 relative (the reference
averages in fp32, the kernels in fp64), medians bit-exact."""
import os
import sys

import numpy as np
import pytest

from oracle import metrics as OM

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden_metrics import metric_inputs  # noqa: E402

RTOL = 1e-5


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "metrics.npz"))


def test_oracle_metrics_match_reference(gold):
    gt_rgb, pred_rgb, gt_d, pred_d, masks = metric_inputs()
    assert np.array_equal(OM.get_psnr(pred_rgb, gt_rgb), gold["psnr"])
    assert np.array_equal(OM.depth_evaluation(gt_d, pred_d), gold["depth"])
    assert np.array_equal(OM.depth_evaluation(gt_d, pred_d, masks), gold["depth_masked"])
    assert np.array_equal(OM.depth_evaluation(gt_d, pred_d, None, 1.0, 30.0), gold["depth_range"])
    gt2 = gt_d.copy()
    gt2[1] = 0.0
    assert np.array_equal(OM.depth_evaluation(gt2, pred_d), gold["depth_skip"])
    v = gt_d[2] > 1e-4
    assert np.array_equal(np.array(OM.compute_errors(gt_d[2][v], pred_d[2][v]), dtype=np.float64), gold["errors_one"])


def _close(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert np.all(np.abs(a - b) <= RTOL * np.abs(b)), (a, b)


@pytest.mark.gpu
def test_cuda_metrics_match_reference(gold):
    import torch
    from ucnerf_b200 import metrics as M
    gt_rgb, pred_rgb, gt_d, pred_d, masks = metric_inputs()
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    _close(M.get_psnr([T(p) for p in pred_rgb], [T(g) for g in gt_rgb]), gold["psnr"])
    for key, kw in (("depth", {}), ("depth_masked", {"pred_masks": T(masks)}),
                    ("depth_range", {"min_depth": 1.0, "max_depth": 30.0})):
        got = M.depth_evaluation(T(gt_d), T(pred_d), verbose=False, **kw)
        _close(got, gold[key])
    gt2 = gt_d.copy()
    gt2[1] = 0.0
    _close(M.depth_evaluation(T(gt2), T(pred_d), verbose=False), gold["depth_skip"])
    v = gt_d[2] > 1e-4
    got = M.compute_errors(T(gt_d[2][v]), T(pred_d[2][v]))
    _close(got, gold["errors_one"])
    # threshold fractions are pixel counts over the same fp32 quotients: exact
    n = int(v.sum())
    assert [round(x * n) for x in got[4:]] == [round(x * n) for x in gold["errors_one"][4:]]


@pytest.mark.gpu
@pytest.mark.parametrize("n", (1, 2, 7, 4096, 100001))
def test_cuda_median_is_exact(n):
    """Radix select == numpy's median for odd / even counts, negative and repeated values, with a mask."""
    import ctypes as C
    import torch
    from ucnerf_b200 import lib as L
    lib = L.load()
    rng = np.random.default_rng(n)
    gt = (rng.random(n) * 50 + 0.01).astype(np.float32)
    pred = (rng.standard_normal(n) * 10).astype(np.float32)
    pred[::3] = pred[0]                                           # ties
    mask = rng.random(n) > 0.25
    mask[0] = True
    gt[n // 2] = 0.0 if n > 2 else gt[n // 2]                     # an invalid pixel
    valid = (gt > 1e-4) & (gt < 100) & mask
    g, p, m = (torch.from_numpy(a).cuda() for a in (gt, pred, mask.astype(np.uint8)))
    ws = torch.empty(int(lib.ucn_metrics_workspace_bytes()), dtype=torch.uint8, device="cuda")
    mid = torch.zeros(4, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    L.check(lib.ucn_depth_median(L.ptr(g), L.ptr(p), C.c_void_p(m.data_ptr()), n, 1e-4, 100.0, C.c_void_p(ws.data_ptr()),
                                 ws.numel(), L.ptr(mid), C.c_void_p(cnt.data_ptr()), L.stream_ptr()))
    assert int(cnt.item()) == int(valid.sum())
    h = mid.cpu().numpy()
    assert (h[0] + h[1]) * np.float32(0.5) == np.median(gt[valid])
    assert (h[2] + h[3]) * np.float32(0.5) == np.median(pred[valid])


@pytest.mark.gpu
def test_cuda_metrics_full_frame_properties():
    """640x512 frames (config 2): depth errors are invariant to a global scale of the prediction (median alignment),
    a perfect prediction scores (0, 0, 0, 0, 1, 1, 1), and PSNR of a constant offset d is -20 log10 d."""
    import torch
    from ucnerf_b200 import metrics as M
    g = torch.Generator("cuda").manual_seed(3)
    gt = torch.rand((2, 512, 640), device="cuda", generator=g) * 60 + 0.5
    pred = gt * (1 + 0.1 * torch.randn((2, 512, 640), device="cuda", generator=g)).abs() + 0.01
    e1 = M.depth_evaluation(gt, pred, verbose=False)
    e2 = M.depth_evaluation(gt, pred * 4.0, verbose=False)              # power of two: exact in fp32
    assert np.allclose(e1, e2, rtol=1e-6, atol=0)
    e0 = M.depth_evaluation(gt, gt.clone(), verbose=False)
    assert np.allclose(e0, [0, 0, 0, 0, 1, 1, 1], atol=1e-12)
    rgb = torch.rand((512 * 640, 3), device="cuda", generator=g) * 0.5
    ps = M.get_psnr([rgb + 0.125], [rgb])
    assert abs(ps[0] - (-20 * np.log10(0.125))) < 1e-4
    with pytest.raises(Exception):
        M.get_psnr([rgb.cpu()], [rgb.cpu()])                            # no CPU path
