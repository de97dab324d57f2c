"""Drop-in mirror of the reference's PE + uncertainty-conditioned MLP (network/models.py:13-283).

    network/models.py:20   Embedder / :56 get_embedder      network/models.py:74   BaseAdapt_Renderer
    network/models.py:187  UCNeRF                            network/models.py:209  create_ucnerf (factory)

``UCNeRF`` keeps the reference's parameter names and shapes (``state_dict()`` keys ``nerf.*``, including the
three layers forward() never uses, Q9) so released ``.tar`` checkpoints load unchanged; ``forward`` runs
the fused CUDA kernel (PE + 13 linear layers + FiLM + uncertainty blend) -- fp32 SIMT (`mode='fp32'`) or
bf16 tcgen05 (`mode='bf16'`).  In train() mode, when autograd is recording and a parameter or the feature rows require
grad (the training step, train.py:147-190), ``forward_fused`` switches to the fp32 training primitives and their backward
(train_path.MlpFn).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.nn as nn

from . import lib as L
from . import train_path as _tp
from .frame import PackedWeights


def weights_init(m):
    """network/models.py:13-17."""
    if isinstance(m, nn.Linear):
        nn.init.kaiming_normal_(m.weight.data)
        if m.bias is not None:
            nn.init.zeros_(m.bias.data)


class Embedder:
    """network/models.py:20-54.  embed(x) = [x | sin(2^l x), l<L | cos(2^l x), l<L] (Q6 order).
    Inside the MLP kernel the encoding is fused; this standalone version serves foreign callers."""

    def __init__(self, **kwargs):
        self.kwargs = kwargs
        d, L_ = kwargs['input_dims'], kwargs['num_freqs']
        if not kwargs.get('log_sampling', True):
            raise L.UcnError("only log_sampling=True (the reference default) is supported")
        self.num_freqs = L_
        self.freq_bands = 2. ** torch.linspace(0., kwargs['max_freq_log2'], steps=L_)
        self.out_dim = d * (1 + 2 * L_) if kwargs.get('include_input', True) else d * 2 * L_

    def embed(self, inputs):
        fb = self.freq_bands.to(inputs.device)
        scaled = (inputs.unsqueeze(-2) * fb.view(*[1] * (inputs.dim() - 1), -1, 1)).reshape(*inputs.shape[:-1], -1)
        return torch.cat((inputs, torch.sin(scaled), torch.cos(scaled)), dim=-1)


def get_embedder(multires, i=0, input_dims=3):
    """network/models.py:56-71 -> (embed_fn, out_dim)."""
    if i == -1:
        return nn.Identity(), 3
    eo = Embedder(include_input=True, input_dims=input_dims, max_freq_log2=multires - 1, num_freqs=multires,
                  log_sampling=True, periodic_fns=[torch.sin, torch.cos])

    def embed(x, eo=eo):
        return eo.embed(x)
    embed.embedder = eo
    return embed, eo.out_dim


def _is_default_embedders(embed_fn, embeddirs_fn) -> bool:
    e0, e1 = getattr(embed_fn, "embedder", None), getattr(embeddirs_fn, "embedder", None)
    return e0 is not None and e1 is not None and e0.num_freqs == 10 and e1.num_freqs == 4 \
        and e0.kwargs['input_dims'] == 3 and e1.kwargs['input_dims'] == 3


class BaseAdapt_Renderer(nn.Module):
    """Parameter container with the reference's layer names/shapes (network/models.py:74-118)."""

    def __init__(self, D=8, W=256, input_ch=3, input_ch_views=3, output_ch=4, input_ch_feat=8, skips=[4],
                 use_viewdirs=False, fine=False, view_num=4):
        super().__init__()
        self.D, self.W, self.input_ch, self.input_ch_views, self.skips = D, W, input_ch, input_ch_views, skips
        self.view_num = view_num - 1
        self.use_viewdirs = use_viewdirs
        self.in_ch_pts, self.in_ch_views, self.in_ch_feat = input_ch, input_ch_views, input_ch_feat
        self.pts_linears = nn.ModuleList(
            [nn.Linear(input_ch, W, bias=True)] +
            [nn.Linear(W, W, bias=True) if i not in skips else nn.Linear(W + input_ch, W) for i in range(D - 1)])
        self.pts_bias_depth_fine = nn.Linear(24 + 4 * self.view_num, W)
        self.pts_bias_confidence = nn.Linear(8 * self.view_num, W)
        self.pts_bias_confidence_1 = nn.Linear(1, 1)
        self.views_linears = nn.ModuleList([nn.Linear(input_ch_views + W, W // 2)])
        self.view_confi_linears = nn.ModuleList([nn.Linear(input_ch_views + W, W // 2)])
        if not use_viewdirs:
            raise L.UcnError("use_viewdirs=False is not the UC-NeRF configuration (opt.py:86 default True)")
        self.feature_linear = nn.Linear(W, W)
        self.feature_linear_1 = nn.Linear(W, W)
        self.confi_linear = nn.Linear(W, W)
        self.alpha_linear = nn.Linear(W // 2, 1)
        self.alpha_linear_1 = nn.Linear(W, 1)
        self.rgb_linear = nn.Linear(W // 2, 3)
        self.confi_rgb_linear = nn.Linear(W, 3)
        for m in (self.pts_bias_depth_fine, self.pts_linears, self.views_linears, self.view_confi_linears,
                  self.confi_linear, self.pts_bias_confidence_1, self.feature_linear, self.feature_linear_1,
                  self.alpha_linear, self.rgb_linear, self.confi_rgb_linear):
            m.apply(weights_init)


class UCNeRF(nn.Module):
    """network/models.py:187-207, forward on libucnerf.so."""

    def __init__(self, D=8, W=256, input_ch_pts=3, input_ch_views=3, input_ch_feat=8, skips=[4], net_type='v2',
                 fine=False, view_num=4, mode: str = "bf16"):
        super().__init__()
        if (D, W, input_ch_pts, input_ch_views, list(skips)) != (6, 128, 63, 27, [4]):
            raise L.UcnError("the CUDA MLP is specialised for the reference defaults D=6, W=128, multires=10, "
                             "multires_views=4, skips=[4] (opt.py:40-44,90-92; models.py:225)")
        self.in_ch_pts, self.in_ch_views, self.in_ch_feat = input_ch_pts, input_ch_views, input_ch_feat
        self.view_num = view_num
        self.mode = mode
        self.nerf = BaseAdapt_Renderer(D=D, W=W, input_ch_feat=input_ch_feat, input_ch=input_ch_pts, output_ch=4,
                                       skips=skips, input_ch_views=input_ch_views, use_viewdirs=True, fine=fine,
                                       view_num=view_num)
        self._packed: Optional[PackedWeights] = None
        self._packed_key = None

    # -- packed weights follow the parameters (re-packed when any parameter is modified in place) --
    def packed(self) -> PackedWeights:
        params = list(self.parameters())
        key = tuple((p.data_ptr(), p._version) for p in params)
        if self._packed is None or key != self._packed_key:
            dev = params[0].device
            if dev.type != "cuda":
                raise L.UcnError("UCNeRF must live on a CUDA device: there is no CPU path (call .cuda())")
            self._packed = PackedWeights({k: v.detach() for k, v in self.state_dict().items()}, self.view_num - 1, dev)
            self._packed_key = key
        return self._packed

    def forward_uncertainty(self, x):
        return 1 - x

    def forward_alpha(self, x):
        raise L.UcnError("forward_alpha is broken in the reference (self.pts_bias undefined, models.py:127): dead path")

    def forward_fused(self, ndc_pts, viewdirs, feats, mode: Optional[str] = None):
        """ndc_pts [R,S,3], viewdirs [R,3] (per ray), feats [R,S,F] -> raw [R,S,4] = (rgb, sigma)."""
        R, S = feats.shape[:2]
        if self.training and _tp.wants_grad(feats, *self.parameters()):
            # training step (train.py:147-190): primitives with saved activations + their backward (K3b).  Only in
            # train() mode: the eval loop (train.py:251-279, Lightning's validation: eval() + no_grad) renders whole
            # frames, whose saved activations (9.6 KB per sample) would not fit.
            named = dict(self.nerf.named_parameters())
            return _tp.MlpFn.apply(ndc_pts, viewdirs, feats, S, *[named[n] for n in _tp.PARAM_NAMES])
        pw = self.packed()
        raw = torch.empty((R, S, 4), device=feats.device)
        args = [t.to(torch.float32).contiguous() for t in (ndc_pts, feats, viewdirs)]
        L.check(L.load().ucn_mlp_forward(C.byref(pw.c), L.ptr(args[0]), L.ptr(args[1]), L.ptr(args[2]), R, S,
                                         L.ptr(raw), L.MODES[mode or self.mode], L.stream_ptr()), "ucn_mlp_forward")
        return raw

    def forward(self, x):
        """x [..., 63+F+27] = [PE(pts) | input_feat | PE(viewdir)] (network/renderer.py:88-100).  The raw
        3-vectors are the first three channels of each PE block, so the fused kernel re-derives the
        encodings itself; every sample carries its own view direction here (S=1)."""
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        F_ = x2.shape[-1] - self.in_ch_pts - self.in_ch_views
        pts = x2[:, :3]
        feats = x2[:, self.in_ch_pts:self.in_ch_pts + F_]
        dirs = x2[:, self.in_ch_pts + F_:self.in_ch_pts + F_ + 3]
        raw = self.forward_fused(pts.reshape(-1, 1, 3), dirs, feats.reshape(-1, 1, F_))
        return raw.reshape(*lead, 4)


def create_ucnerf(args, pts_embedder=True, dir_embedder=True, device="cuda", mode="bf16"):
    """network/models.py:209-283, the renderer half: embedders, UCNeRF, network_query_fn, render kwargs.
    The CasMVSNet encoder / checkpoint download of the reference factory is outside the hot path
    (SURVEY.md §2 row 10); load weights with ``model.load_state_dict(ckpt['network_fn_state_dict'])``."""
    from .renderer import run_network_mvs
    embed_fn, input_ch = get_embedder(args.multires, getattr(args, "i_embed", 0))
    embeddirs_fn, input_ch_views = get_embedder(args.multires_views, getattr(args, "i_embed", 0))
    model = UCNeRF(D=args.netdepth, W=args.netwidth, input_ch_pts=input_ch, skips=[4], input_ch_views=input_ch_views,
                   input_ch_feat=args.feat_dim, net_type=args.net_type, view_num=args.view_num, mode=mode).to(device)

    def network_query_fn(pts, viewdirs, rays_feats, network_fn):
        return run_network_mvs(pts, viewdirs, rays_feats, network_fn, embed_fn=embed_fn, embeddirs_fn=embeddirs_fn,
                               netchunk=args.netchunk)
    render_kwargs_train = {'network_query_fn': network_query_fn, 'perturb': getattr(args, "perturb", 1.),
                           'N_samples': args.N_samples, 'network_fn': model,
                           'use_viewdirs': True, 'white_bkgd': getattr(args, "white_bkgd", False),
                           'raw_noise_std': getattr(args, "raw_noise_std", 0.)}
    render_kwargs_test = dict(render_kwargs_train)
    render_kwargs_test['perturb'] = False
    return render_kwargs_train, render_kwargs_test, 0, list(model.parameters())
