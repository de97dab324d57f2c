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


def _ssim_brute(a, b, data_range=1.0):
    """The SSIM definition evaluated window by window (loops, float64): the pin of oracle.metrics.ssim."""
    c1, c2 = (0.01 * data_range) ** 2, (0.03 * data_range) ** 2
    vals = []
    for p, q in zip(np.asarray(a, np.float64), np.asarray(b, np.float64)):
        H, W = p.shape
        s = []
        for y in range(3, H - 3):
            for x in range(3, W - 3):
                wp, wq = p[y - 3:y + 4, x - 3:x + 4].ravel(), q[y - 3:y + 4, x - 3:x + 4].ravel()
                up, uq = wp.mean(), wq.mean()
                vp, vq = wp.var(ddof=1), wq.var(ddof=1)
                vpq = ((wp - up) * (wq - uq)).sum() / 48.0
                s.append(((2 * up * uq + c1) * (2 * vpq + c2)) / ((up * up + uq * uq + c1) * (vp + vq + c2)))
        vals.append(np.mean(s))
    return float(np.mean(vals))


def _ssim_images(seed, C, H, W):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W]
    base = 0.5 + 0.4 * np.sin(0.21 * xx + 0.13 * yy)[None] * np.cos(0.05 * xx * np.arange(1, C + 1)[:, None, None])
    gt = np.clip(base + 0.02 * rng.standard_normal((C, H, W)), 0, 1).astype(np.float32)
    pred = np.clip(gt + 0.08 * rng.standard_normal((C, H, W)), 0, 1).astype(np.float32)
    return gt, pred


def test_oracle_ssim_matches_the_definition():
    """skimage is neither vendored in the reference nor installed here: the restatement is pinned to the published
    formula evaluated window by window, and to the identities SSIM(x, x) = 1 and symmetry."""
    for seed, (C, H, W) in enumerate(((3, 12, 17), (1, 7, 7), (3, 20, 9))):
        gt, pred = _ssim_images(seed, C, H, W)
        assert abs(OM.ssim(gt, pred) - _ssim_brute(gt, pred)) <= 1e-12
        assert abs(OM.ssim(gt, gt) - 1.0) <= 1e-12
        assert abs(OM.ssim(gt, pred) - OM.ssim(pred, gt)) <= 1e-12
    gt, pred = _ssim_images(9, 3, 16, 16)
    assert 0.0 < OM.ssim(gt, pred) < 0.999


@pytest.mark.gpu
def test_cuda_ssim_matches_oracle():
    import torch
    from ucnerf_b200 import metrics as M
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    for seed, (C, H, W) in enumerate(((3, 12, 17), (1, 7, 7), (3, 64, 80), (3, 41, 39), (3, 512, 640))):
        gt, pred = _ssim_images(seed, C, H, W)
        got = M.ssim(T(gt), T(pred))
        assert got.shape == (1,) and abs(got[0] - OM.ssim(gt, pred)) <= 1e-9, (C, H, W, got)
    gts = np.stack([_ssim_images(20 + i, 3, 32, 48)[0] for i in range(3)])
    preds = np.stack([_ssim_images(20 + i, 3, 32, 48)[1] for i in range(3)])
    got = M.ssim(T(gts), T(preds))
    want = [OM.ssim(g, p) for g, p in zip(gts, preds)]
    assert np.abs(got - np.array(want)).max() <= 1e-9
    assert abs(M.ssim(T(gts), T(gts)).max() - 1.0) <= 1e-12
