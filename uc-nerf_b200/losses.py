"""Renderer-side loss terms of the training step (train.py:164-190), SURVEY.md 8f rank 3.

These are the terms of the reference's loss that act on the renderer's outputs (rgb, depth of the selected rays), on
~2 048 rays and 50 4x4 patches.  `render_losses_fused` is the product path: ONE launch of `ucn_render_losses`
(csrc/ucn_losses.cu) returns every term and the closed-form gradient d loss / d (rgb, depth), CUDA tensors only.
`render_losses` states the same contract as elementwise torch code (the mirror the tests pin to the reference's loss
modules).  `cas_mvsnet_loss` (network/mvs_models.py:504-530) acts on the encoder's outputs and is not part of this path.

    utils/utils.py:11-12         img2mse, mse2psnr
    network/mvs_models.py:491-509 EdgePreservingSmoothnessLoss  -> edge_preserving_smoothness
    utils/loss.py:54-112          compute_scale_and_shift + gradient_loss + GradientLoss -> scale_invariant_gradient
    train.py:164-188              the weighted sum -> render_losses
"""
import torch


def img2mse(x, y):
    return torch.mean((x - y) ** 2)


def mse2psnr(x):
    return -10.0 * torch.log10(x)


def edge_preserving_smoothness(depth, guide, gamma: float = 0.1):
    """depth [P,h,w], guide [P,h,w,C]: mean |w * d depth| over the four neighbour directions, with bilateral weights
    w = exp(-|d guide|_1 / gamma) from the guide (the monocular depth prior in train.py:174-178)."""
    def pairs(t):
        return ((t[:, :, :-1], t[:, :, 1:]), (t[:, :-1, :], t[:, 1:, :]), (t[:, :-1, :-1], t[:, 1:, 1:]),
                (t[:, 1:, :-1], t[:, :-1, 1:]))
    total = 0.0
    for (g0, g1), (d0, d1) in zip(pairs(guide), pairs(depth)):
        w = torch.exp(-(g0 - g1).abs().sum(-1) / gamma)
        total = total + (w * (d0 - d1)).abs().mean()
    return total / 4


def scale_invariant_gradient(pred, target, mask=None):
    """pred, target [P,h,w]: align pred to target per patch by the least-squares scale and shift, then sum the L1 norms of
    the horizontal and vertical gradients of the residual, divided by P (utils/loss.py:54-112; the masks are only used
    by the alignment there, as here)."""
    mask = torch.ones_like(pred) if mask is None else mask
    a00, a01, a11 = (mask * pred * pred).sum((1, 2)), (mask * pred).sum((1, 2)), mask.sum((1, 2))
    b0, b1 = (mask * pred * target).sum((1, 2)), (mask * target).sum((1, 2))
    det = a00 * a11 - a01 * a01
    ok = det != 0
    safe = torch.where(ok, det, torch.ones_like(det))
    scale = torch.where(ok, (a11 * b0 - a01 * b1) / safe, torch.zeros_like(det))
    shift = torch.where(ok, (-a01 * b0 + a00 * b1) / safe, torch.zeros_like(det))
    diff = scale.view(-1, 1, 1) * pred + shift.view(-1, 1, 1) - target
    gx = (diff[:, :, 1:] - diff[:, :, :-1]).abs().sum((1, 2))
    gy = (diff[:, 1:, :] - diff[:, :-1, :]).abs().sum((1, 2))
    return (gx + gy).sum() / pred.shape[0]


def render_losses(rgb, depth_pred, target_s, pix, N_rays, patch_num, patch_size, sparse_depths, sparse_weights, dpt):
    """The renderer-side part of train.py:164-188: -> (loss, dict of terms).
    rgb [R,3], depth_pred [R], target_s [R,3], pix [2,R] int64 (rows, cols); the first patch_num*patch_size^2 rays are
    the patches, rays N_rays.. are the sparse-depth rays; sparse_depths / sparse_weights / dpt [1,H,W]."""
    ps, half = patch_size, patch_num // 2
    patch_pts = patch_num * ps * ps
    target_depths = sparse_depths[:, pix[0, N_rays:], pix[1, N_rays:]]
    target_weights = sparse_weights[:, pix[0, N_rays:], pix[1, N_rays:]]
    patch_depth = depth_pred[:patch_pts].reshape(-1, ps, ps)
    patch_dpt = dpt[:, pix[0, :patch_pts], pix[1, :patch_pts]].reshape(-1, ps, ps, 1)
    smooth = edge_preserving_smoothness(patch_depth[:half], patch_dpt[:half])
    nerf_depth = torch.mean(((depth_pred[N_rays:] - target_depths) ** 2) * target_weights)
    scale_inv = scale_invariant_gradient(patch_depth[half:], patch_dpt[half:].squeeze(-1))
    img = img2mse(rgb, target_s)
    loss = nerf_depth * 0.05 + smooth * 0.05 + scale_inv * 0.008 + img * 5.0
    return loss, {"img": img, "nerf_depth": nerf_depth, "smooth": smooth, "scale_invariant": scale_inv}


class _FusedLossFn(torch.autograd.Function):
    """loss and d loss / d (rgb, depth) from ONE launch (ucn_render_losses); backward only scales the saved gradients."""

    @staticmethod
    def forward(ctx, rgb, depth_pred, target_s, pix, N_rays, patch_num, patch_size, sparse_depths, sparse_weights, dpt):
        import ctypes as C
        from . import lib as L
        lib = L.load()
        R = rgb.shape[0]
        H, W = dpt.shape[-2:]
        f = lambda t: t.detach().to(torch.float32).contiguous()
        rgb_c, dep_c, tgt_c = f(rgb), f(depth_pred), f(target_s)
        rows, cols = pix[0].to(torch.int64).contiguous(), pix[1].to(torch.int64).contiguous()
        terms = torch.empty(5, device=rgb.device, dtype=torch.float32)
        d_rgb, d_depth = torch.empty_like(rgb_c), torch.empty_like(dep_c)
        L.check(lib.ucn_render_losses(L.ptr(rgb_c), L.ptr(dep_c), L.ptr(tgt_c), C.c_void_p(rows.data_ptr()),
                                      C.c_void_p(cols.data_ptr()), R, int(N_rays), int(patch_num), int(patch_size),
                                      L.ptr(f(sparse_depths)), L.ptr(f(sparse_weights)), L.ptr(f(dpt)), int(H), int(W),
                                      L.ptr(terms), L.ptr(d_rgb), L.ptr(d_depth), L.stream_ptr()), "ucn_render_losses")
        ctx.save_for_backward(d_rgb, d_depth)
        ctx.mark_non_differentiable(terms)
        return terms[0].clone(), terms

    @staticmethod
    def backward(ctx, g_loss, _g_terms):
        d_rgb, d_depth = ctx.saved_tensors
        return g_loss * d_rgb, g_loss * d_depth, None, None, None, None, None, None, None, None


def render_losses_fused(rgb, depth_pred, target_s, pix, N_rays, patch_num, patch_size, sparse_depths, sparse_weights, dpt):
    """Same contract as `render_losses`, one CUDA launch for the value and the gradient (CUDA tensors only; no CPU
    path).  -> (loss, dict of terms)."""
    if not rgb.is_cuda:
        from .lib import UcnError
        raise UcnError("render_losses_fused: CUDA tensors only (use render_losses for the torch glue)")
    loss, t = _FusedLossFn.apply(rgb, depth_pred, target_s, pix, N_rays, patch_num, patch_size, sparse_depths,
                                 sparse_weights, dpt)
    return loss, {"img": t[1], "nerf_depth": t[2], "smooth": t[3], "scale_invariant": t[4]}
