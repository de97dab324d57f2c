"""CPU: pin the backward restatements (oracle/ucnerf_oracle.c uco_*_bwd, oracle/train_ops.py) and the host-side layer
sequence of the training path (ucnerf_b200/train_path.py) against gradients produced by the reference's own
autograd (tests/golden/grad_*.npz <- tests/golden/make_golden_grad.py)."""
import numpy as np
import pytest
import torch

from helpers import GRAD_CASES, load_grad_case, rel_err
from oracle import oracle as O
from oracle.train_ops import OracleOps
from ucnerf_b200 import train_path as TP


def _forward_inputs(c):
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
    py, px = O.chunk_pixels(c["H"], c["W"])
    R = c["R"]
    r = O.build_rays_test(osc, py[:R], px[:R], c["t_rand"][:R])
    feat = O.gen_pts_feats(osc, r["pts"], r["ndc_all"])
    ang = O.dir_feature(osc, r["rays_d"])
    return osc, r, feat, ang


@pytest.mark.parametrize("name", GRAD_CASES)
def test_composite_backward(name):
    c = load_grad_case(name)
    g = c["gg"]
    osc, r, feat, ang = _forward_inputs(c)
    raw = O.run_network(O.OracleWeights(c["weights"], c["V"]), r["ndc"]["ndc"], feat, ang)
    d_raw = O.raw2outputs_bwd(raw, r["z"], c["g_rgb"], c["g_depth"])
    assert rel_err(d_raw, g["d_raw"]) <= 2e-5


@pytest.mark.parametrize("name", GRAD_CASES)
def test_mlp_backward_layer_sequence(name):
    c = load_grad_case(name)
    g = c["gg"]
    osc, r, feat, ang = _forward_inputs(c)
    R, S, F = feat.shape
    T = torch.from_numpy
    P = {n: T(np.ascontiguousarray(c["weights"]["nerf." + n])) for n in TP.PARAM_NAMES}
    raw, saved = TP.mlp_forward_train(OracleOps(), P, T(r["ndc"]["ndc"].reshape(-1, 3).copy()), T(ang.copy()),
                                      T(feat.reshape(-1, F).copy()), S)
    ref_raw = O.run_network(O.OracleWeights(c["weights"], c["V"]), r["ndc"]["ndc"], feat, ang)
    assert np.abs(raw.numpy().reshape(R, S, 4) - ref_raw).max() <= 2e-5
    G = {n: torch.zeros_like(p) for n, p in P.items()}
    d_feat = TP.mlp_backward(OracleOps(), P, saved, T(g["d_raw"].reshape(-1, 4).copy()), G)
    assert rel_err(d_feat.numpy().reshape(R, S, F), g["d_feat"]) <= 2e-5
    for n in TP.PARAM_NAMES:
        assert rel_err(G[n].numpy(), g["p:nerf." + n]) <= 5e-5, n
    # the three layers forward() never touches get no gradient in the reference either (Q9)
    assert not any(k.startswith("p:nerf.confi_linear") or k.startswith("p:nerf.feature_linear_1") for k in g)


@pytest.mark.parametrize("name", GRAD_CASES)
def test_gather_backward(name):
    c = load_grad_case(name)
    g = c["gg"]
    osc, r, feat, ang = _forward_inputs(c)
    dv, dfe, dco = O.gen_pts_feats_bwd(osc, r["pts"], r["ndc_all"], g["d_feat"])
    for k in range(3):
        assert rel_err(dv[k], g[f"d_vol{k}"][0]) <= 2e-5, k
    assert rel_err(dfe, g["d_img_feats"][:, 0]) <= 2e-5
    assert rel_err(dco, g["d_conf"].reshape(dco.shape)) <= 2e-5
