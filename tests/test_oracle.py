"""CPU: pin the oracle (oracle/) against golden vectors produced by the reference's own code
(tests/golden/make_golden.py).  Positions bit-exact; features/MLP/compositing to fp32 tolerance."""
import numpy as np
import pytest

from helpers import CASES, bits, load_case
from oracle import oracle as O


def test_linspace_matches_torch(golden_dir):
    import torch
    pinned = np.load(f"{golden_dir}/torch_linspace01.npz")
    for k in pinned.files:
        assert np.array_equal(bits(O.linspace01(int(k))), bits(pinned[k])), k
    for n in list(range(1, 130)) + [255, 256]:
        assert np.array_equal(bits(O.linspace01(n)), bits(torch.linspace(0., 1., n).numpy())), n


@pytest.mark.parametrize("name", CASES)
def test_positions_bit_exact(name):
    c = load_case(name)
    g = c["g"]
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
    py, px = O.chunk_pixels(c["H"], c["W"])
    r = O.build_rays_test(osc, py, px, c["t_rand"])
    assert np.array_equal(bits(r["z"]), bits(g["z"]))
    assert np.array_equal(bits(r["rays_d"]), bits(g["rays_d"]))
    assert np.array_equal(bits(r["pts"]), bits(g["pts"]))
    for k in ("stage1", "stage2", "stage3", "ndc"):
        assert np.array_equal(bits(r["ndc"][k]), bits(g[k])), k


@pytest.mark.parametrize("name", CASES)
def test_features_and_render(name):
    c = load_case(name)
    g = c["g"]
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=c["chunk"])
    ow = O.OracleWeights(c["weights"], c["V"])
    py, px = O.chunk_pixels(c["H"], c["W"])
    r = O.build_rays_test(osc, py, px, c["t_rand"])
    feat, idx = O.gen_pts_feats(osc, r["pts"], r["ndc_all"], with_idx=True)
    assert np.abs(feat[::c["stride"]] - g["feat_sub"]).max() <= 1e-5
    # masks are exact 0/1 values
    V = c["V"]
    for v in range(V):
        assert np.array_equal(feat[::c["stride"], :, 24 + 4 * v + 3], g["feat_sub"][..., 24 + 4 * v + 3])
    assert idx.min() >= 0
    ang = O.dir_feature(osc, r["rays_d"])
    assert np.abs(ang - g["angle"]).max() <= 1e-6       # incl. quirk Q4 (frame switch after chunk 0)
    raw = O.run_network(ow, r["ndc"]["ndc"], feat, ang)
    assert np.abs(raw - g["raw"]).max() <= 1e-4
    rgb, depth = O.raw2outputs(g["raw"], g["z"])
    assert np.abs(rgb - g["rgb"]).max() <= 1e-5
    assert np.abs(depth - g["depth"]).max() <= 1e-5
    rgb2, depth2 = O.render_rays(osc, ow, 0, c["H"] * c["W"], c["t_rand"])
    assert np.abs(rgb2 - g["rgb"]).max() <= 1e-4
    assert np.abs(depth2 - g["depth"]).max() <= 1e-4


def test_embed_order():
    x = np.array([[0.3, -0.7, 0.11]], np.float32)
    e = O.embed(x, 4)[0]
    assert e.shape == (27,)
    assert np.allclose(e[:3], x[0])
    assert np.allclose(e[3:6], np.sin(x[0]))           # [x | all sin | all cos], Q6
    assert np.allclose(e[6:9], np.sin(2 * x[0]))
    assert np.allclose(e[15:18], np.cos(x[0]))


@pytest.mark.parametrize("name", ("lin", "lin_jit", "disp_jit"))
def test_ray_marcher_bit_exact(name, golden_dir):
    """data/ray_utils.py:152-197 restated in oracle/ucnerf_oracle.c vs the reference's own output."""
    import sys
    sys.path.insert(0, golden_dir)
    from make_golden_marcher import marcher_inputs
    g = np.load(f"{golden_dir}/ray_marcher.npz")
    rays, rnd, c = marcher_inputs(name)
    xyz, z = O.ray_marcher(rays, c["S"], c["lindisp"], c["perturb"], rnd)
    assert np.array_equal(bits(z), bits(g[name + "_z"]))
    assert np.array_equal(bits(xyz), bits(g[name + "_xyz"]))


@pytest.mark.parametrize("name", ("train_rays_small_v3", "train_rays_s126_v3"))
def test_training_sampler_positions_bit_exact(name):
    """a13: the reference's build_rays (utils/utils.py:400-597) on its own pixel selection and jitter."""
    from helpers import load_train_rays_case
    c = load_train_rays_case(name)
    g = c["tg"]
    osc = O.OracleScene(c["scene"], c["S"], dir_switch=1 << 30)
    pix = g["pix"].astype(np.float32)
    r = O.build_rays_test(osc, pix[0], pix[1], g["t_rand"])
    assert np.array_equal(bits(r["z"]), bits(g["z"]))
    assert np.array_equal(bits(r["rays_d"]), bits(g["rays_d"]))
    assert np.array_equal(bits(r["pts"]), bits(g["pts"]))
    for k in ("stage1", "stage2", "stage3", "ndc"):
        assert np.array_equal(bits(r["ndc"][k]), bits(g[k])), k
    # the sparse-depth rays come last, in the order of the coordinate list (utils/utils.py:490-496)
    assert np.array_equal(g["pix"][:, c["N_rays"]:], c["coords"].T)
    assert np.array_equal(g["colors"], c["target"][:, g["pix"][0], g["pix"][1]].T)
