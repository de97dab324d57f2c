"""Training-step path: the per-ray renderer with gradients.

The reference trains through autograd (train.py:147-190): `rendering` (network/renderer.py:215-255) is called on the
rays `build_rays` selected, the loss is back-propagated into the UC-NeRF parameters (network/models.py:87-104) and,
through the two `grid_sample` groups of `gen_pts_feats` (utils/utils.py:776-788,859-890), into the three feature
volumes, the source-view FPN features and the confidence map.  Sample positions carry no gradient (the depth
hypotheses are detached, network/mvs_models.py:716-717).

This module is the host side of that graph on libucnerf.so:

    CompositeFn   K4  ucn_composite      /  K4b ucn_composite_bwd
    MlpFn         K3  fp32 primitives with saved activations  /  K3b (ucn_linear_*, ucn_film_relu_*, ucn_mul_*,
                      ucn_blend_*; layer sequence = `mlp_forward_train` / `mlp_backward` below)
    GatherFn      K2  ucn_gather         /  K2b ucn_gather_bwd

The layer sequence is written once against an `ops` object.  The product always passes `CudaOps` (ctypes calls on
CUDA tensors, no fallback); tests/ passes oracle.train_ops.OracleOps to check the host logic on CPU against the
reference's own autograd gradients (tests/golden/grad_*.npz).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

import torch

from . import lib as L

# parameters the live forward touches, in the order MlpFn receives them (models.py:87-104; Q9: the three unused
# layers get no gradient, exactly as in the reference)
PARAM_NAMES: List[str] = (
    [f"pts_linears.{i}.{s}" for i in range(6) for s in ("weight", "bias")]
    + [f"{n}.{s}" for n in ("pts_bias_depth_fine", "pts_bias_confidence", "feature_linear", "views_linears.0",
                            "view_confi_linears.0", "rgb_linear", "alpha_linear", "confi_rgb_linear", "alpha_linear_1")
       for s in ("weight", "bias")])


def _v(t):
    """(device pointer, leading dimension) of a row-major 2-D fp32 view."""
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1) or t.dtype != torch.float32:
        raise L.UcnError(f"expected a row-major 2-D fp32 view, got shape {tuple(t.shape)} stride {t.stride()}")
    if not t.is_cuda:
        raise L.UcnError("libucnerf has no CPU path: tensor must live on a CUDA device")
    return C.c_void_p(t.data_ptr()), int(t.stride(0))


class CudaOps:
    """The K3b primitives of include/ucnerf.h on CUDA tensors (current stream)."""

    def __init__(self):
        self.lib = L.load()

    def linear_fwd(self, X, W, b, Y, accumulate=False, relu=False):
        (x, ldx), (w, ldw), (y, ldy) = _v(X), _v(W), _v(Y)
        L.check(self.lib.ucn_linear_fwd(x, ldx, w, ldw, L.ptr(b), X.shape[0], X.shape[1], W.shape[0], y, ldy,
                                        int(accumulate) | (2 if relu else 0), L.stream_ptr()), "ucn_linear_fwd")

    def linear_bwd_input(self, dY, W, dX, accumulate=False):
        (g, ldy), (w, ldw), (x, ldx) = _v(dY), _v(W), _v(dX)
        L.check(self.lib.ucn_linear_bwd_input(g, ldy, w, ldw, dY.shape[0], W.shape[1], W.shape[0], x, ldx,
                                              int(accumulate), L.stream_ptr()), "ucn_linear_bwd_input")

    def linear_bwd_weight(self, dY, X, dW, db):
        (g, ldy), (x, ldx), (w, ldw) = _v(dY), _v(X), _v(dW)
        L.check(self.lib.ucn_linear_bwd_weight(g, ldy, x, ldx, dY.shape[0], X.shape[1], dY.shape[1], w, ldw, L.ptr(db),
                                               L.stream_ptr()), "ucn_linear_bwd_weight")

    def prep(self, ndc, angle, S, PH, FTV):
        (ph, ldph), (ftv, ldftv) = _v(PH), _v(FTV)
        L.check(self.lib.ucn_train_prep(L.ptr(ndc), L.ptr(angle), S, ndc.shape[0], ph, ldph, ftv, ldftv, L.stream_ptr()),
                "ucn_train_prep")

    def film_relu_fwd(self, PRE, MOD, H):
        h, ldh = _v(H)
        L.check(self.lib.ucn_film_relu_fwd(L.ptr(PRE), L.ptr(MOD), h, ldh, PRE.shape[0], L.stream_ptr()), "ucn_film_relu_fwd")

    def film_relu_bwd(self, dH, PRE, MOD, dPRE, dMOD, accumulate, db):
        g, ldg = _v(dH)
        L.check(self.lib.ucn_film_relu_bwd(g, ldg, L.ptr(PRE), L.ptr(MOD), L.ptr(dPRE), L.ptr(dMOD), int(accumulate),
                                           L.ptr(db), PRE.shape[0], L.stream_ptr()), "ucn_film_relu_bwd")

    def mul_fwd(self, A, B, OUT):
        a, lda = _v(A)
        L.check(self.lib.ucn_mul_fwd(a, lda, L.ptr(B), L.ptr(OUT), B.shape[0], L.stream_ptr()), "ucn_mul_fwd")

    def mul_bwd(self, dOUT, A, B, dA, accumulate, dB):
        a, lda = _v(A)
        L.check(self.lib.ucn_mul_bwd(L.ptr(dOUT), a, lda, L.ptr(B), L.ptr(dA), int(accumulate), L.ptr(dB), B.shape[0],
                                     L.stream_ptr()), "ucn_mul_bwd")

    def relu_bwd(self, dY, Y, dX):
        L.check(self.lib.ucn_relu_bwd(L.ptr(dY), L.ptr(Y), L.ptr(dX), Y.numel(), L.stream_ptr()), "ucn_relu_bwd")

    def blend_fwd(self, BASE4, ADAPT4, conf, raw):
        c, ldc = _v(conf)
        L.check(self.lib.ucn_blend_fwd(L.ptr(BASE4), L.ptr(ADAPT4), c, ldc, L.ptr(raw), raw.shape[0], L.stream_ptr()),
                "ucn_blend_fwd")

    def blend_bwd(self, d_raw, raw, BASE4, ADAPT4, conf, dBASE4, dADAPT4, d_conf):
        (c, ldc), (dc, lddc) = _v(conf), _v(d_conf)
        L.check(self.lib.ucn_blend_bwd(L.ptr(d_raw), L.ptr(raw), L.ptr(BASE4), L.ptr(ADAPT4), c, ldc, L.ptr(dBASE4),
                                       L.ptr(dADAPT4), dc, lddc, raw.shape[0], L.stream_ptr()), "ucn_blend_bwd")


def mlp_forward_train(ops, P: Dict[str, torch.Tensor], ndc, angle, feat, S: int):
    """network/models.py:138-184 with every activation the backward needs kept.
    ndc [M,3], angle [R,3] (per ray, M = R*S), feat [M,F] -> raw [M,4], saved dict."""
    M, F = feat.shape
    V = (F - 25) // 12
    Kf, Kb = 24 + 4 * V, 8 * V
    new = lambda *shape: feat.new_empty(shape)
    # every column that is ever read (0..190 / 0..154) is written below; the pad columns only round the pitch up to a
    # multiple of four floats and are masked out by the kernels, so the buffers need no zero fill
    PH = new(M, 192)                       # [PE_10(ndc) 0..62 | h4 63..190 | pad]   (skip concat, models.py:156-157)
    FTV = new(M, 160)                      # [feature 0..127 | PE_4(dir) 128..154 | pad]  (models.py:166)
    ops.prep(ndc, angle, S, PH, FTV)
    MOD, FB = new(M, 128), new(M, 128)
    ops.linear_fwd(feat[:, :Kf], P["pts_bias_depth_fine.weight"], P["pts_bias_depth_fine.bias"], MOD)       # :150
    ops.linear_fwd(feat[:, Kf:Kf + Kb], P["pts_bias_confidence.weight"], P["pts_bias_confidence.bias"], FB)  # :151
    PRE, H, H5 = new(6, M, 128), new(4, M, 128), new(M, 128)
    x = PH[:, :63]
    for i in range(6):                                                                                      # :153-157
        ops.linear_fwd(x, P[f"pts_linears.{i}.weight"], P[f"pts_linears.{i}.bias"], PRE[i])
        out = H[i] if i < 4 else (PH[:, 63:191] if i == 4 else H5)
        ops.film_relu_fwd(PRE[i], MOD, out)
        x = PH[:, :191] if i == 4 else out
    BASE4, ADAPT4, HFB, VH, raw = new(M, 4), new(M, 4), new(M, 128), new(M, 128), new(M, 4)
    # base heads (:161-162) stacked into one 4-wide product: h is streamed once instead of twice
    WB = torch.cat((P["confi_rgb_linear.weight"], P["alpha_linear_1.weight"]), 0)
    ops.linear_fwd(H5, WB, torch.cat((P["confi_rgb_linear.bias"], P["alpha_linear_1.bias"]), 0), BASE4)
    ops.mul_fwd(H5, FB, HFB)
    ops.linear_fwd(HFB, P["feature_linear.weight"], P["feature_linear.bias"], FTV[:, :128])                  # :164
    # the two 155 -> 64 view branches (:168-169, :172-173) read the same input: one 128-wide product on the stacked
    # weights fills the 128 x 128 tensor-core tile that each 64-wide product would leave half empty
    WV = torch.cat((P["views_linears.0.weight"], P["view_confi_linears.0.weight"]), 0)
    BV = torch.cat((P["views_linears.0.bias"], P["view_confi_linears.0.bias"]), 0)
    ops.linear_fwd(FTV[:, :155], WV, BV, VH, relu=True)
    # adapt heads (:170, :174): rgb reads columns 0..63 of VH, alpha columns 64..127 -- one 4-wide product on the
    # block matrix [[rgb_w, 0], [0, alpha_w]] (the zero blocks add exact zeros)
    WA = WB.new_zeros((4, 128))
    WA[:3, :64] = P["rgb_linear.weight"]
    WA[3:, 64:] = P["alpha_linear.weight"]
    ops.linear_fwd(VH, WA, torch.cat((P["rgb_linear.bias"], P["alpha_linear.bias"]), 0), ADAPT4)
    ops.blend_fwd(BASE4, ADAPT4, feat[:, F - 1:F], raw)                                                      # :149,177-180
    saved = dict(PH=PH, FTV=FTV, MOD=MOD, FB=FB, PRE=PRE, H=H, H5=H5, BASE4=BASE4, ADAPT4=ADAPT4, HFB=HFB, VH=VH,
                 raw=raw, feat=feat, WV=WV, WB=WB, WA=WA)
    return raw, saved


def mlp_backward(ops, P: Dict[str, torch.Tensor], saved, d_raw, G: Dict[str, torch.Tensor]):
    """Backward of `mlp_forward_train`: accumulates parameter gradients into the zero-initialised G[name]
    (reference layouts) and returns d_feat [M,F]."""
    s = saved
    feat, PH, FTV, MOD, FB, PRE, H, H5, VH, HFB = (s[k] for k in ("feat", "PH", "FTV", "MOD", "FB", "PRE", "H", "H5", "VH", "HFB"))
    M, F = feat.shape
    V = (F - 25) // 12
    Kf, Kb = 24 + 4 * V, 8 * V
    new = lambda *shape: feat.new_empty(shape)
    d_feat = new(M, F)                     # columns [0,Kf) | [Kf,Kf+Kb) | F-1 are each written (not accumulated) below
    dBASE4, dADAPT4 = new(M, 4), new(M, 4)
    ops.blend_bwd(d_raw, s["raw"], s["BASE4"], s["ADAPT4"], feat[:, F - 1:F], dBASE4, dADAPT4, d_feat[:, F - 1:F])
    # adapt heads (models.py:170,174) and the two 155 -> 64 view branches (:168-173)
    dVH, dVPRE, dFT = new(M, 128), new(M, 128), new(M, 128)
    WA, WB = s["WA"], s["WB"]
    GA, GAb = torch.zeros_like(WA), WA.new_zeros(4)
    ops.linear_bwd_input(dADAPT4, WA, dVH)
    ops.linear_bwd_weight(dADAPT4, VH, GA, GAb)
    G["rgb_linear.weight"].copy_(GA[:3, :64]); G["alpha_linear.weight"].copy_(GA[3:, 64:])
    G["rgb_linear.bias"].copy_(GAb[:3]); G["alpha_linear.bias"].copy_(GAb[3:])
    ops.relu_bwd(dVH, VH, dVPRE)
    WV = s["WV"]                                                       # stacked [views_linears.0 ; view_confi_linears.0]
    ops.linear_bwd_input(dVPRE, WV[:, :128], dFT)
    GWV, GBV = torch.zeros_like(WV), WV.new_zeros(128)
    ops.linear_bwd_weight(dVPRE, FTV[:, :155], GWV, GBV)
    G["views_linears.0.weight"].copy_(GWV[:64]); G["view_confi_linears.0.weight"].copy_(GWV[64:])
    G["views_linears.0.bias"].copy_(GBV[:64]); G["view_confi_linears.0.bias"].copy_(GBV[64:])
    # feature_linear on h * feats_bias (:164), base heads (:161-162)
    dHFB, dH, dFB = new(M, 128), new(M, 128), new(M, 128)
    ops.linear_bwd_input(dFT, P["feature_linear.weight"], dHFB)
    ops.linear_bwd_weight(dFT, HFB, G["feature_linear.weight"], G["feature_linear.bias"])
    GB, GBb = torch.zeros_like(WB), WB.new_zeros(4)
    ops.linear_bwd_input(dBASE4, WB, dH)
    ops.linear_bwd_weight(dBASE4, H5, GB, GBb)
    G["confi_rgb_linear.weight"].copy_(GB[:3]); G["alpha_linear_1.weight"].copy_(GB[3:])
    G["confi_rgb_linear.bias"].copy_(GBb[:3]); G["alpha_linear_1.bias"].copy_(GBb[3:])
    ops.mul_bwd(dHFB, H5, FB, dH, True, dFB)
    # the six FiLM-modulated layers, last to first (:153-157)
    dPRE, dMOD = new(M, 128), new(M, 128)
    for i in range(5, -1, -1):
        ops.film_relu_bwd(dH, PRE[i], MOD, dPRE, dMOD, i != 5, G[f"pts_linears.{i}.bias"])     # bias grad = colsum(dPRE)
        x_i = PH[:, :63] if i == 0 else (PH[:, :191] if i == 5 else H[i - 1])
        ops.linear_bwd_weight(dPRE, x_i, G[f"pts_linears.{i}.weight"], None)
        if i == 0:
            break
        W = P[f"pts_linears.{i}.weight"]
        ops.linear_bwd_input(dPRE, W[:, 63:] if i == 5 else W, dH)       # PE(ndc) carries no gradient
    # FiLM / feature-bias generators (:150-151)
    ops.linear_bwd_input(dMOD, P["pts_bias_depth_fine.weight"], d_feat[:, :Kf])
    ops.linear_bwd_weight(dMOD, feat[:, :Kf], G["pts_bias_depth_fine.weight"], G["pts_bias_depth_fine.bias"])
    ops.linear_bwd_input(dFB, P["pts_bias_confidence.weight"], d_feat[:, Kf:Kf + Kb])
    ops.linear_bwd_weight(dFB, feat[:, Kf:Kf + Kb], G["pts_bias_confidence.weight"], G["pts_bias_confidence.bias"])
    return d_feat


# ---- torch.autograd glue ------------------------------------------------------------------------------------------

class MlpFn(torch.autograd.Function):
    """raw = UCNeRF(ndc, angle, feat) with gradients for feat and the parameters (PARAM_NAMES order)."""

    @staticmethod
    def forward(ctx, ndc, angle, feat, S, *params):
        P = {n: p.detach() for n, p in zip(PARAM_NAMES, params)}
        for n, p in P.items():
            if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                raise L.UcnError(f"parameter {n} must be a contiguous fp32 CUDA tensor")
        M = feat.shape[0] * feat.shape[1]
        f2 = feat.detach().to(torch.float32).reshape(M, feat.shape[-1]).contiguous()
        raw, saved = mlp_forward_train(CudaOps(), P, ndc.detach().to(torch.float32).reshape(M, 3).contiguous(),
                                       angle.detach().to(torch.float32).contiguous(), f2, int(S))
        ctx.saved, ctx.P, ctx.shape = saved, P, tuple(feat.shape)
        return raw.reshape(feat.shape[0], feat.shape[1], 4)

    @staticmethod
    def backward(ctx, d_raw):
        # every parameter gradient is a view of ONE zero-filled buffer (one fill launch instead of 30); offsets are
        # multiples of four floats so that the 16-byte kernel variants still apply
        offs, total = {}, 0
        for n, p in ctx.P.items():
            offs[n] = total
            total += (p.numel() + 3) & ~3
        flat = torch.zeros((total,), device=d_raw.device, dtype=torch.float32)
        G = {n: flat[offs[n]:offs[n] + p.numel()].view(p.shape) for n, p in ctx.P.items()}
        d_feat = mlp_backward(CudaOps(), ctx.P, ctx.saved, d_raw.to(torch.float32).reshape(-1, 4).contiguous(), G)
        ctx.saved = None
        return (None, None, d_feat.reshape(ctx.shape), None) + tuple(G[n] for n in PARAM_NAMES)


class CompositeFn(torch.autograd.Function):
    """(rgb_map, depth_map) = raw2outputs(raw, z) (network/renderer.py:125-131) with the gradient for raw."""

    @staticmethod
    def forward(ctx, raw, z, white_bkgd):
        R, S = z.shape
        raw4 = raw.detach()[..., :4].to(torch.float32).contiguous()
        z32 = z.detach().to(torch.float32).contiguous()
        rgb, depth = torch.empty((R, 3), device=raw.device), torch.empty((R,), device=raw.device)
        L.check(L.load().ucn_composite(L.ptr(raw4), L.ptr(z32), R, S, int(bool(white_bkgd)), L.ptr(rgb), L.ptr(depth),
                                       None, None, L.stream_ptr()), "ucn_composite")
        ctx.save_for_backward(raw4, z32)
        ctx.white_bkgd = int(bool(white_bkgd))
        return rgb, depth

    @staticmethod
    def backward(ctx, g_rgb, g_depth):
        raw4, z32 = ctx.saved_tensors
        R, S = z32.shape
        g_rgb = (torch.zeros((R, 3), device=raw4.device) if g_rgb is None else g_rgb).to(torch.float32).contiguous()
        g_depth = (torch.zeros((R,), device=raw4.device) if g_depth is None else g_depth).to(torch.float32).contiguous()
        d_raw = torch.empty_like(raw4)
        L.check(L.load().ucn_composite_bwd(L.ptr(raw4), L.ptr(z32), R, S, ctx.white_bkgd, L.ptr(g_rgb), L.ptr(g_depth),
                                           L.ptr(d_raw), L.stream_ptr()), "ucn_composite_bwd")
        return d_raw, None, None


class GatherFn(torch.autograd.Function):
    """input_feat = gen_pts_feats(...) (network/renderer.py:177-212) with gradients for the three feature volumes,
    the source-view features and the confidence map.  `frame` is the PackedFrame the forward gathers from."""

    @staticmethod
    def forward(ctx, frame, pts, ndc1, ndc2, ndc3, vol0, vol1, vol2, img_feat, confidence):
        R, S = pts.shape[:2]
        a = [t.detach().to(torch.float32).contiguous() for t in (pts, ndc1, ndc2, ndc3)]
        feat = torch.empty((R, S, frame.F), device=pts.device)
        L.check(L.load().ucn_gather(C.byref(frame.c), *[L.ptr(t) for t in a], R * S, L.PART_ALL, L.ptr(feat), None,
                                    L.stream_ptr()), "ucn_gather")
        ctx.frame, ctx.a = frame, a
        ctx.shapes = [None if t is None else tuple(t.shape) for t in (vol0, vol1, vol2, img_feat, confidence)]
        return feat

    @staticmethod
    def backward(ctx, d_feat):
        g = d_feat.to(torch.float32).contiguous()
        need = ctx.needs_input_grad[5:10]
        outs = [torch.zeros(sh, device=g.device) if (sh is not None and nd) else None for sh, nd in zip(ctx.shapes, need)]
        n = g.shape[0] * g.shape[1]
        L.check(L.load().ucn_gather_bwd(C.byref(ctx.frame.c), *[L.ptr(t) for t in ctx.a], n, L.ptr(g),
                                        *[L.ptr(o) for o in outs], L.stream_ptr()), "ucn_gather_bwd")
        return (None, None, None, None, None) + tuple(outs)


def wants_grad(*tensors) -> bool:
    return torch.is_grad_enabled() and any(t is not None and getattr(t, "requires_grad", False) for t in tensors)


# ---- data-parallel gradient exchange (SURVEY.md 8e "Collective (train)") --------------------------------------------

def allreduce_grads(params, world_size: int, bucket: "torch.Tensor | None" = None):
    """One all-reduce(sum)/P per step over a flat fp32 bucket of every parameter gradient (NCCL over NVLink on the
    GPU box, gloo in the CPU tests); parameters without a gradient (the three unused layers, Q9) contribute zeros so
    that every rank reduces the same bucket.  Returns the bucket (reusable across steps)."""
    import torch.distributed as dist
    params = list(params)
    n = sum(p.numel() for p in params)
    if bucket is None or bucket.numel() != n:
        bucket = torch.empty((n,), device=params[0].device, dtype=torch.float32)
    # pack: ONE concatenation kernel into the bucket (not one copy per parameter); unpack: one fused multi-tensor copy
    pieces = [p.grad.reshape(-1).to(torch.float32) if p.grad is not None else bucket.new_zeros(p.numel()) for p in params]
    torch.cat(pieces, out=bucket)
    if world_size > 1:
        dist.all_reduce(bucket, op=dist.ReduceOp.SUM)
        bucket.div_(world_size)
        chunks = bucket.split([p.numel() for p in params])
        dst = [p.grad for p in params if p.grad is not None]
        src = [c.view_as(p.grad) for c, p in zip(chunks, params) if p.grad is not None]
        torch._foreach_copy_(dst, src)
    return bucket
