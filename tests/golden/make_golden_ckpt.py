"""Manifest of the reference's checkpoint layout, made by EXECUTING THE REFERENCE (build container only).

Instantiates the reference's own ``UCNeRF`` (network/models.py:187-207) for view_num 4 and 6, builds the dict its
``save_ckpt`` writes (train.py:404-414: ``{'network_fn_state_dict', 'network_mvs_state_dict'}``), round-trips it through
``torch.save`` / ``torch.load`` and records every key with its shape and dtype, plus the encoder's key count, in
``tests/golden/ckpt_manifest.json``.  The weights themselves are random init and are not stored.

Usage:  python tests/golden/make_golden_ckpt.py
"""
import io
import json
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference  # noqa: E402


def main():
    U, M, Rn = import_reference()
    out = {"format": "torch.save dict, train.py:404-414", "top_level_keys": ["network_fn_state_dict", "network_mvs_state_dict"]}
    for view_num in (4, 6):
        torch.manual_seed(0)
        _, ic = M.get_embedder(10, 0)
        _, icv = M.get_embedder(4, 0)
        feat_dim = 24 + 12 * (view_num - 1) + 1
        model = M.UCNeRF(D=6, W=128, input_ch_pts=ic, skips=[4], input_ch_views=icv, input_ch_feat=feat_dim,
                         view_num=view_num)
        enc = M.CascadeMVSNet(view_num=view_num)
        buf = io.BytesIO()
        torch.save({'network_fn_state_dict': model.state_dict(), 'network_mvs_state_dict': enc.state_dict()}, buf)
        buf.seek(0)
        ck = torch.load(buf, map_location="cpu")
        assert list(ck.keys()) == out["top_level_keys"]
        out[f"view_num_{view_num}"] = {
            "network_fn_state_dict": {k: [list(v.shape), str(v.dtype)] for k, v in ck['network_fn_state_dict'].items()},
            "network_fn_numel": int(sum(v.numel() for v in ck['network_fn_state_dict'].values())),
            "network_mvs_keys": len(ck['network_mvs_state_dict']),
        }
    with open(os.path.join(HERE, "ckpt_manifest.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote ckpt_manifest.json", {k: v["network_fn_numel"] for k, v in out.items() if k.startswith("view")})


if __name__ == "__main__":
    main()
