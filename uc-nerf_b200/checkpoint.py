"""On-disk compatibility with the reference's checkpoints.

    train.py:404-414          save_ckpt: torch.save({'network_fn_state_dict', 'network_mvs_state_dict'}, '<dir>/<name>.tar')
    network/models.py:254-266 create_ucnerf: ckpt = torch.load(args.ckpt); EncodingNet.load_state_dict(...);
                              model.load_state_dict(ckpt['network_fn_state_dict'])

``UCNeRF`` here keeps the reference's parameter names and shapes (``nerf.*``, the three layers its forward never
touches included), so a released ``ucnerf.tar`` loads unchanged and a file written here loads in the reference.  The
CasMVSNet encoder is outside the hot path: its state dict is carried through untouched (handed to the caller's own
encoder module when one is given, kept verbatim on save).  Loading re-packs the kernel-side weight tiles on the next
forward (``UCNeRF.packed`` keys on the parameters' versions).
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

from . import lib as L

FN_KEY = "network_fn_state_dict"
MVS_KEY = "network_mvs_state_dict"


def load_ckpt(path: str, network_fn: torch.nn.Module, network_mvs: Optional[torch.nn.Module] = None,
              map_location="cpu", strict: bool = True) -> Dict:
    """network/models.py:258-266.  Loads ``ckpt['network_fn_state_dict']`` into ``network_fn`` (strict, as the
    reference) and, when an encoder module is passed, ``ckpt['network_mvs_state_dict']`` into it.  Returns the
    checkpoint dict.  Errors are the reference's: a missing file or key raises, a shape mismatch raises."""
    ckpt = torch.load(path, map_location=map_location, weights_only=True)
    if not isinstance(ckpt, dict) or FN_KEY not in ckpt:
        raise L.UcnError(f"{path}: not a UC-NeRF checkpoint (no '{FN_KEY}', train.py:408)")
    if network_mvs is not None:
        if MVS_KEY not in ckpt:
            raise L.UcnError(f"{path}: no '{MVS_KEY}' (train.py:410)")
        network_mvs.load_state_dict(ckpt[MVS_KEY], strict=strict)
    network_fn.load_state_dict(ckpt[FN_KEY], strict=strict)
    return ckpt


def save_ckpt(path: str, network_fn: torch.nn.Module, network_mvs=None, allow_missing_mvs: bool = False) -> str:
    """train.py:404-414.  ``network_mvs`` is the encoder module or its state dict (e.g. the one ``load_ckpt``
    returned).  The reference's loader calls ``EncodingNet.load_state_dict`` strictly (network/models.py:261-262), so a
    file without the encoder's weights does NOT load there: passing None raises unless ``allow_missing_mvs=True`` (an
    empty dict is then written, for renderer-only use with this package's ``load_ckpt``)."""
    d = os.path.dirname(os.path.abspath(path))
    if network_mvs is None and not allow_missing_mvs:
        raise L.UcnError("save_ckpt: network_mvs (encoder module or state dict) is required for a file the reference "
                         "can load; pass allow_missing_mvs=True for a renderer-only checkpoint")
    os.makedirs(d, exist_ok=True)
    if network_mvs is None:
        mvs = {}
    elif isinstance(network_mvs, torch.nn.Module):
        mvs = network_mvs.state_dict()
    else:
        mvs = dict(network_mvs)
    cpu = lambda sd: {k: v.detach().to("cpu") for k, v in sd.items()}
    torch.save({FN_KEY: cpu(network_fn.state_dict()), MVS_KEY: cpu(mvs)}, path)
    return path
