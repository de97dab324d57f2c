"""Diagnostic (GPU): per-tensor gradient error of the training primitives in both GEMM modes + aligned-operand GEMM check."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from helpers import load_grad_case, rel_err
from oracle import oracle as O
from ucnerf_b200 import lib as L, train_path as TP
DEV = "cuda"
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
ops = TP.CudaOps()
g = torch.Generator(device="cpu").manual_seed(0)
for mode in (1, 0):
    L.check(L.load().ucn_train_gemm_mode(mode))
    for M, K, N in ((1000, 128, 128), (3072, 192, 128), (3072, 64, 128), (3072, 160, 64), (3072, 36, 128)):
        X = torch.randn((M, K), generator=g).to(DEV); W = torch.randn((N, K), generator=g).to(DEV)
        Y = torch.zeros((M, N), device=DEV); ops.linear_fwd(X, W, None, Y)
        e1 = rel_err(Y.cpu().numpy(), (X.double() @ W.double().t()).cpu().numpy())
        Kp = K - 1                                   # ragged reduction on aligned rows (vec path with a masked tail)
        Y2 = torch.zeros((M, N), device=DEV); ops.linear_fwd(X[:, :Kp], W[:, :Kp], None, Y2)
        e2 = rel_err(Y2.cpu().numpy(), (X[:, :Kp].double() @ W[:, :Kp].double().t()).cpu().numpy())
        dY = torch.randn((M, N), generator=g).to(DEV); dX = torch.zeros((M, K), device=DEV)
        ops.linear_bwd_input(dY, W, dX)
        e3 = rel_err(dX.cpu().numpy(), (dY.double() @ W.double()).cpu().numpy())
        dW = torch.zeros((N, K), device=DEV); ops.linear_bwd_weight(dY, X, dW, None)
        e4 = rel_err(dW.cpu().numpy(), (dY.double().t() @ X.double()).cpu().numpy())
        print(f"mode {mode} M{M} K{K} N{N}: fwd {e1:.2e} ragged {e2:.2e} bwd_in {e3:.2e} bwd_w {e4:.2e}")
    for name in ("grad_small_v3",):
        c = load_grad_case(name); gg = c["gg"]
        osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
        py, px = O.chunk_pixels(c["H"], c["W"]); R = c["R"]
        r = O.build_rays_test(osc, py[:R], px[:R], c["t_rand"][:R])
        feat = O.gen_pts_feats(osc, r["pts"], r["ndc_all"]); ang = O.dir_feature(osc, r["rays_d"])
        R, S, F = feat.shape
        P = {n: T(c["weights"]["nerf." + n]) for n in TP.PARAM_NAMES}
        raw, saved = TP.mlp_forward_train(ops, P, T(r["ndc"]["ndc"].reshape(-1, 3)), T(ang), T(feat.reshape(-1, F)), S)
        print(f"mode {mode} raw abs err {np.abs(raw.cpu().numpy().reshape(R, S, 4) - c['g']['raw'][:R]).max():.2e}")
        G = {n: torch.zeros_like(p) for n, p in P.items()}
        d_feat = TP.mlp_backward(ops, P, saved, T(gg["d_raw"].reshape(-1, 4)), G)
        print(f"mode {mode} d_feat {rel_err(d_feat.cpu().numpy().reshape(R, S, F), gg['d_feat']):.2e}")
        print("   " + " ".join(f"{n.replace('_linears', '').replace('.weight', '.w').replace('.bias', '.b')}:{rel_err(G[n].cpu().numpy(), gg['p:nerf.' + n]):.1e}" for n in TP.PARAM_NAMES))
L.check(L.load().ucn_train_gemm_mode(0))
