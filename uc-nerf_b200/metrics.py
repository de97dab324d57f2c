"""Device-resident evaluation metrics: host-side mirror of the reference's evaluation helpers over libucnerf
(ucn_sq_err_sum / ucn_depth_median / ucn_depth_errors, include/ucnerf.h).

    get_psnr          utils/utils.py:18-22
    compute_errors    utils/evaluation.py:8-27
    depth_evaluation  utils/evaluation.py:30-74

Same names, argument meaning and return values; inputs are CUDA tensors (the rendered frames never leave HBM: the
reference evaluates on the host after a per-chunk `.cpu()`), only the handful of fp64 sums comes back.  SSIM and LPIPS
(utils/evaluation.py:76-101: skimage's gaussian SSIM and an AlexNet with downloaded weights) are not part of the
render path's caller loop and are not provided.  There is no CPU fallback: a missing extension raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import lib as L


def _f32(t: torch.Tensor) -> torch.Tensor:
    if not t.is_cuda:
        raise L.UcnError("metrics: tensors must live on a CUDA device (no CPU path)")
    return t.detach().to(torch.float32).contiguous()


def get_psnr(imgs_pred: Sequence[torch.Tensor], imgs_gt: Sequence[torch.Tensor]) -> np.ndarray:
    """PSNR of every (prediction, target) pair -> float64 array, one device->host copy for the whole list."""
    lib = L.load()
    n_img = len(imgs_pred)
    if n_img == 0:
        return np.array([])
    sums = torch.zeros(n_img, dtype=torch.float64, device=imgs_pred[0].device)
    counts = []
    for i, (img, tar) in enumerate(zip(imgs_pred, imgs_gt)):
        a, b = _f32(img), _f32(tar)
        if a.numel() != b.numel():
            raise ValueError("get_psnr: prediction and target differ in size")
        L.check(lib.ucn_sq_err_sum(L.ptr(a), L.ptr(b), a.numel(), C.c_void_p(sums[i:].data_ptr()), L.stream_ptr()),
                "ucn_sq_err_sum")
        counts.append(a.numel())
    mse = sums.cpu().numpy() / np.array(counts, dtype=np.float64)
    return -10.0 * np.log(mse) / np.log(10.0)


def ssim(gts: torch.Tensor, predicts: torch.Tensor, data_range: float = 1.0) -> np.ndarray:
    """Per-image SSIM of [N,C,H,W] (or [C,H,W]) images in [0, data_range] -> float64 array [N]: what
    utils/evaluation.py:88-95 computes with skimage.metrics.structural_similarity(gt, pred, data_range=1,
    channel_axis=2) (7 x 7 uniform window, sample covariance, mean of the map cropped by 3 pixels, mean over channels),
    on the GPU; one device->host copy for the whole batch."""
    lib = L.load()
    a, b = _f32(gts), _f32(predicts)
    if a.dim() == 3:
        a, b = a[None], b[None]
    if a.shape != b.shape or a.dim() != 4:
        raise ValueError("ssim: expected two [N,C,H,W] (or [C,H,W]) tensors of the same shape")
    N, Cn, H, W = a.shape
    sums = torch.zeros(N, dtype=torch.float64, device=a.device)
    for i in range(N):
        L.check(lib.ucn_ssim_sum(L.ptr(a[i]), L.ptr(b[i]), Cn, H, W, float(data_range), C.c_void_p(sums[i:].data_ptr()),
                                 L.stream_ptr()), "ucn_ssim_sum")
    return sums.cpu().numpy() / float(Cn * (H - 6) * (W - 6))


def _errors_from_sums(s: np.ndarray):
    c = s[7]
    return (s[0] / c, s[1] / c, float(np.sqrt(s[2] / c)), float(np.sqrt(s[3] / c)), s[4] / c, s[5] / c, s[6] / c)


def compute_errors(gt: torch.Tensor, pred: torch.Tensor):
    """(abs_rel, sq_rel, rmse, rmse_log, a1, a2, a3) over every element of two equally-shaped CUDA tensors."""
    lib = L.load()
    g, p = _f32(gt), _f32(pred)
    out = torch.zeros(8, dtype=torch.float64, device=g.device)
    inf = float(np.finfo(np.float32).max)
    L.check(lib.ucn_depth_errors(L.ptr(g), L.ptr(p), None, g.numel(), -inf, inf, 1.0, C.c_void_p(out.data_ptr()),
                                 L.stream_ptr()), "ucn_depth_errors")
    return _errors_from_sums(out.cpu().numpy())


def depth_evaluation(gt_depths: torch.Tensor, pred_depths: torch.Tensor, savedir=None,
                     pred_masks: Optional[torch.Tensor] = None, min_depth: float = 0.0001, max_depth: float = 100,
                     verbose: bool = True) -> np.ndarray:
    """Mean depth errors over the frames [num, H, W] after the reference's global median scaling.  `pred_masks`
    (bool / uint8) must have the ground-truth resolution (the reference's cv2 resize is an identity then)."""
    lib = L.load()
    assert gt_depths.shape[0] == pred_depths.shape[0]
    g, p = _f32(gt_depths), _f32(pred_depths)
    m = None
    if pred_masks is not None:
        if pred_masks.shape != gt_depths.shape:
            raise ValueError("depth_evaluation: masks must have the ground-truth resolution")
        m = (pred_masks.detach() != 0).to(torch.uint8).contiguous()
    num, per = g.shape[0], g[0].numel()
    dev = g.device
    ws = torch.empty(int(lib.ucn_metrics_workspace_bytes()), dtype=torch.uint8, device=dev)
    mid = torch.zeros(4, dtype=torch.float32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    mp = None if m is None else C.c_void_p(m.data_ptr())
    L.check(lib.ucn_depth_median(L.ptr(g), L.ptr(p), mp, g.numel(), min_depth, max_depth, C.c_void_p(ws.data_ptr()),
                                 ws.numel(), L.ptr(mid), C.c_void_p(cnt.data_ptr()), L.stream_ptr()), "ucn_depth_median")
    mid_h = mid.cpu().numpy()
    if int(cnt.item()) == 0:
        raise ValueError("depth_evaluation: no valid pixel")          # the reference fails in np.concatenate
    med_gt = (mid_h[0] + mid_h[1]) * np.float32(0.5)                   # numpy's median of an even count
    med_pred = (mid_h[2] + mid_h[3]) * np.float32(0.5)
    ratio = np.float32(med_gt / med_pred)
    sums = torch.zeros((num, 8), dtype=torch.float64, device=dev)
    for i in range(num):
        mi = None if m is None else C.c_void_p(m[i].data_ptr())
        L.check(lib.ucn_depth_errors(L.ptr(g[i]), L.ptr(p[i]), mi, per, min_depth, max_depth, float(ratio),
                                     C.c_void_p(sums[i].data_ptr()), L.stream_ptr()), "ucn_depth_errors")
    s = sums.cpu().numpy()
    errors = [_errors_from_sums(s[i]) for i in range(num) if s[i, 7] > 0]   # frames without a valid pixel are skipped
    mean_errors = np.array(errors).mean(0)
    if verbose:
        print("\n  " + ("{:>8} | " * 7).format("abs_rel", "sq_rel", "rmse", "rmse_log", "a1", "a2", "a3"))
        print(("&{: 8.3f}  " * 7).format(*mean_errors.tolist()) + "\\\\")
    return mean_errors
