"""Checkpoint compatibility: the reference's .tar layout (train.py:404-414,
network/models.py:254-266) against tests/golden/ckpt_manifest.json, which tests/golden/make_golden_ckpt.py wrote from
the reference's own UCNeRF / save format."""
import json
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN_DIR, load_case
from ucnerf_b200 import checkpoint, lib as L, synthetic
from ucnerf_b200.models import UCNeRF


def manifest():
    with open(os.path.join(GOLDEN_DIR, "ckpt_manifest.json")) as f:
        return json.load(f)


def make_model(view_num):
    return UCNeRF(D=6, W=128, input_ch_pts=63, skips=[4], input_ch_views=27,
                  input_ch_feat=24 + 12 * (view_num - 1) + 1, view_num=view_num)


@pytest.mark.parametrize("view_num", [4, 6])
def test_state_dict_matches_reference_manifest(view_num):
    want = manifest()[f"view_num_{view_num}"]
    sd = make_model(view_num).state_dict()
    assert set(sd.keys()) == set(want["network_fn_state_dict"].keys())
    for k, (shape, dtype) in want["network_fn_state_dict"].items():
        assert list(sd[k].shape) == shape and str(sd[k].dtype) == dtype, k
    assert sum(v.numel() for v in sd.values()) == want["network_fn_numel"]


def test_reference_format_round_trip(tmp_path):
    """A file laid out as the reference's save_ckpt writes it loads; a file written here has the same layout."""
    wts = synthetic.make_weights(3, seed=5)
    ref_sd = {k: torch.from_numpy(v) for k, v in wts.items()}
    enc_sd = {"feature.conv0.0.conv.weight": torch.arange(24.).reshape(2, 3, 2, 2)}   # opaque to this package
    p = str(tmp_path / "ckpts" / "latest.tar")
    os.makedirs(os.path.dirname(p))
    torch.save({'network_fn_state_dict': ref_sd, 'network_mvs_state_dict': enc_sd}, p)
    m = make_model(4)
    ck = checkpoint.load_ckpt(p, m)
    for k, v in m.state_dict().items():
        assert np.array_equal(v.numpy(), wts[k]), k
    p2 = checkpoint.save_ckpt(str(tmp_path / "out" / "ucnerf.tar"), m, ck[checkpoint.MVS_KEY])
    ck2 = torch.load(p2, map_location="cpu", weights_only=True)
    assert list(ck2.keys()) == manifest()["top_level_keys"]
    assert set(ck2[checkpoint.FN_KEY]) == set(ref_sd)
    for k, v in ck2[checkpoint.FN_KEY].items():
        assert torch.equal(v, ref_sd[k]), k
    assert torch.equal(ck2[checkpoint.MVS_KEY]["feature.conv0.0.conv.weight"], enc_sd["feature.conv0.0.conv.weight"])


def test_errors(tmp_path):
    m = make_model(4)
    with pytest.raises(FileNotFoundError):
        checkpoint.load_ckpt(str(tmp_path / "missing.tar"), m)
    p = str(tmp_path / "bad.tar")
    torch.save({"something_else": {}}, p)
    with pytest.raises(L.UcnError):
        checkpoint.load_ckpt(p, m)
    # a file without the encoder's weights does not load in the reference (strict): refused unless asked for
    with pytest.raises(L.UcnError):
        checkpoint.save_ckpt(str(tmp_path / "no_mvs.tar"), m)
    p3 = checkpoint.save_ckpt(str(tmp_path / "no_mvs.tar"), m, allow_missing_mvs=True)
    assert torch.load(p3, map_location="cpu", weights_only=True)[checkpoint.MVS_KEY] == {}
    # a 6-view checkpoint does not fit a 4-view model: same RuntimeError as the reference's strict load
    torch.save({checkpoint.FN_KEY: make_model(6).state_dict(), checkpoint.MVS_KEY: {}}, p)
    with pytest.raises(RuntimeError):
        checkpoint.load_ckpt(p, m)
    # an encoder module is given but the file carries no encoder state
    torch.save({checkpoint.FN_KEY: make_model(6).state_dict()}, p)
    with pytest.raises(L.UcnError):
        checkpoint.load_ckpt(p, make_model(6), network_mvs=torch.nn.Linear(1, 1))
    checkpoint.load_ckpt(p, make_model(6))      # without an encoder the same file loads


@pytest.mark.gpu
def test_render_from_checkpoint_matches_golden(tmp_path):
    """Weights that went through a reference-layout .tar render the golden frame (fp32 gate 1e-4)."""
    from ucnerf_b200.frame import FrameRenderer, PackedFrame
    c = load_case("small_v3")
    p = str(tmp_path / "latest.tar")
    torch.save({'network_fn_state_dict': {k: torch.from_numpy(v) for k, v in c["weights"].items()},
                'network_mvs_state_dict': {}}, p)
    m = make_model(c["V"] + 1).cuda().eval()
    checkpoint.load_ckpt(p, m)
    fr = PackedFrame.from_scene(c["scene"], c["S"], "cuda:0", dir_switch=c["chunk"])   # golden was rendered in chunks (Q4)
    rgb, depth = FrameRenderer(fr, m.packed(), "fp32").render(t_rand=torch.from_numpy(c["t_rand"]).cuda())
    assert np.abs(rgb.cpu().numpy() - c["g"]["rgb"]).max() <= 1e-4
    assert np.abs(depth.cpu().numpy() - c["g"]["depth"]).max() <= 1e-4
