"""CPU restatement of the renderer-side loss terms WITH their analytic gradients (TEST INFRASTRUCTURE ONLY -- never
imported by the product).  numpy fp64; the fused CUDA kernel (ucn_render_losses) computes the same closed forms.

    network/mvs_models.py:491-509  EdgePreservingSmoothnessLoss
    utils/loss.py:54-112           compute_scale_and_shift, gradient_loss, GradientLoss
    train.py:164-188               img2mse, the sparse-depth term and the weighted sum

Pinned by tests/golden/losses.npz (values and autograd gradients from the reference's own modules).
"""
import numpy as np


def smooth_term(depth, guide, gamma=0.1):
    """depth [P,h,w], guide [P,h,w] -> (value, d value / d depth).  Mean over patches and neighbour pairs of
    |w * (d0 - d1)|, w = exp(-|g0 - g1| / gamma), averaged over the four directions."""
    depth, guide = depth.astype(np.float64), guide.astype(np.float64)
    grad = np.zeros_like(depth)
    total = 0.0
    sl = ((np.s_[:, :, :-1], np.s_[:, :, 1:]), (np.s_[:, :-1, :], np.s_[:, 1:, :]),
          (np.s_[:, :-1, :-1], np.s_[:, 1:, 1:]), (np.s_[:, 1:, :-1], np.s_[:, :-1, 1:]))
    for s0, s1 in sl:
        w = np.exp(-np.abs(guide[s0] - guide[s1]) / gamma)
        v = w * (depth[s0] - depth[s1])
        total += np.abs(v).mean() / 4
        gq = w * np.sign(v) / v.size / 4
        grad[s0] += gq
        grad[s1] -= gq
    return total, grad


def scale_invariant_term(pred, target):
    """pred, target [P,h,w] -> (value, d value / d pred): least-squares (scale, shift) per patch, L1 of the residual's
    horizontal + vertical differences, summed over patches / P.  The shift cancels in the differences; the scale
    carries d scale / d pred."""
    pred, target = pred.astype(np.float64), target.astype(np.float64)
    P = pred.shape[0]
    grad = np.zeros_like(pred)
    total = 0.0
    for p in range(P):
        x, t = pred[p], target[p]
        n = x.size
        a00, a01, b0, b1 = (x * x).sum(), x.sum(), (x * t).sum(), t.sum()
        det = a00 * n - a01 * a01
        if det != 0:
            num = n * b0 - a01 * b1
            s = num / det
            ds = ((n * t - b1) - s * (2 * n * x - 2 * a01)) / det
        else:
            s, ds = 0.0, np.zeros_like(x)
        G, gk = 0.0, np.zeros_like(x)
        for s0, s1 in ((np.s_[:, 1:], np.s_[:, :-1]), (np.s_[1:, :], np.s_[:-1, :])):
            r = s * (x[s0] - x[s1]) - (t[s0] - t[s1])
            sg = np.sign(r)
            total += np.abs(r).sum() / P
            G += (sg * (x[s0] - x[s1])).sum()
            gk[s0] += sg
            gk[s1] -= sg
        grad[p] = (G * ds + s * gk) / P
    return total, grad


def render_losses(rgb, depth_pred, target_s, pix, N_rays, patch_num, patch_size, sparse_depths, sparse_weights, dpt):
    """train.py:164-188 (renderer side) -> (terms [loss, img, nerf_depth, smooth, scale_inv], d loss/d rgb, d loss/d depth)."""
    R = rgb.shape[0]
    ps, half = patch_size, patch_num // 2
    pp = patch_num * ps * ps
    rgb64, tgt64, d64 = rgb.astype(np.float64), target_s.astype(np.float64), depth_pred.astype(np.float64)
    img = ((rgb64 - tgt64) ** 2).mean()
    d_rgb = 5.0 * 2 * (rgb64 - tgt64) / rgb64.size
    d_depth = np.zeros(R)
    td = sparse_depths[0][pix[0, N_rays:], pix[1, N_rays:]].astype(np.float64)
    tw = sparse_weights[0][pix[0, N_rays:], pix[1, N_rays:]].astype(np.float64)
    nd = (((d64[N_rays:] - td) ** 2) * tw).mean()
    d_depth[N_rays:] = 0.05 * 2 * (d64[N_rays:] - td) * tw / td.size
    pd = d64[:pp].reshape(-1, ps, ps)
    gd = dpt[0][pix[0, :pp], pix[1, :pp]].reshape(-1, ps, ps)
    sm, g_sm = smooth_term(pd[:half], gd[:half])
    si, g_si = scale_invariant_term(pd[half:], gd[half:])
    d_depth[:half * ps * ps] += 0.05 * g_sm.reshape(-1)
    d_depth[half * ps * ps:pp] += 0.008 * g_si.reshape(-1)
    loss = 0.05 * nd + 0.05 * sm + 0.008 * si + 5.0 * img
    return np.array([loss, img, nd, sm, si]), d_rgb, d_depth
