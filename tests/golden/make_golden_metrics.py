"""Golden values of the evaluation metrics, from the reference's own modules (build container only):
utils/utils.py:18-22 get_psnr and utils/evaluation.py:8-74 compute_errors / depth_evaluation on seeded frames
(with and without prediction masks, with out-of-range ground truth, an all-invalid frame, even and odd valid counts).
Usage: python tests/golden/make_golden_metrics.py"""
import contextlib
import io
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)


def metric_inputs(seed=77, n=3, H=48, W=64):
    rng = np.random.default_rng(seed)
    gt_rgb = rng.random((n, H, W, 3)).astype(np.float32)
    pred_rgb = np.clip(gt_rgb + 0.05 * rng.standard_normal((n, H, W, 3)), 0, 1).astype(np.float32)
    gt_d = (0.5 + 40.0 * rng.random((n, H, W))).astype(np.float32)
    pred_d = (gt_d * (0.6 + 0.1 * rng.standard_normal((n, H, W))) + 0.3).astype(np.float32)
    pred_d = np.abs(pred_d) + np.float32(1e-3)
    gt_d[0, :5] = 0.0                     # invalid rows (below min_depth)
    gt_d[1, :, :7] = 150.0                # above max_depth
    gt_d[2, 3, 3] = 0.0                   # makes the total valid count odd
    masks = (rng.random((n, H, W)) > 0.3)
    return gt_rgb, pred_rgb, gt_d, pred_d, masks


def main():
    from make_golden import import_reference
    import_reference()
    for name in ("lpips", "skimage"):
        sys.modules.setdefault(name, types.ModuleType(name))
    from utils.evaluation import compute_errors, depth_evaluation
    from utils.utils import get_psnr
    gt_rgb, pred_rgb, gt_d, pred_d, masks = metric_inputs()
    out = {}
    out["psnr"] = get_psnr(list(pred_rgb), [torch.from_numpy(g) for g in gt_rgb])
    with contextlib.redirect_stdout(io.StringIO()):
        out["depth"] = depth_evaluation(gt_d.copy(), pred_d.copy())
        out["depth_masked"] = depth_evaluation(gt_d.copy(), pred_d.copy(), pred_masks=masks)
        out["depth_range"] = depth_evaluation(gt_d.copy(), pred_d.copy(), min_depth=1.0, max_depth=30.0)
        gt2 = gt_d.copy()
        gt2[1] = 0.0                      # a frame with no valid pixel is skipped (evaluation.py:49-50)
        out["depth_skip"] = depth_evaluation(gt2, pred_d.copy())
    g, p = gt_d[2][gt_d[2] > 1e-4], pred_d[2][gt_d[2] > 1e-4]
    out["errors_one"] = np.array(compute_errors(g, p), dtype=np.float64)
    np.savez(os.path.join(HERE, "metrics.npz"), **out)
    for k, v in out.items():
        print(k, np.asarray(v))


if __name__ == "__main__":
    main()
