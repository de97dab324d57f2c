"""The allocation-free entry point `ucn_select_training_pixels_static` (include/ucnerf.h; no counterpart in the
reference) against the existing `ucn_select_training_pixels`, both through
the C-ABI.  The selection is integer / index work keyed by a seed, so the bar is the one the other selection tests use:
bit-identical output.  Collected last on purpose: a failure here hides nothing else.
"""
import ctypes as C

import numpy as np
import pytest
import torch

DEV = "cuda"


def _args(H, W, n_rand, patch_num, ps, n_depth, seed_t):
    g = torch.Generator().manual_seed(seed_t)
    conf = torch.rand((H * W,), generator=g).to(DEV)
    coords = torch.stack([torch.randint(0, H, (n_depth,), generator=g),
                          torch.randint(0, W, (n_depth,), generator=g)], 1).to(DEV).contiguous() if n_depth else None
    R = 2 * (patch_num // 2) * ps * ps + n_rand + n_depth
    return conf, coords, R


def _workspace_form(L, lib, conf, coords, H, W, n_rand, patch_num, ps, crop, seed, R):
    ws = torch.zeros(int(lib.ucn_select_workspace_bytes(H, W, patch_num)), dtype=torch.uint8, device=DEV)
    pix = torch.full((2, R), -1.0, device=DEV)
    L.check(lib.ucn_select_training_pixels(L.ptr(conf), H, W, n_rand, patch_num, ps, crop, L.ptr(coords),
                                           0 if coords is None else coords.shape[0], seed, L.ptr(ws), L.ptr(pix),
                                           L.stream_ptr()), "ucn_select_training_pixels")
    return pix


def _static_form(L, lib, conf, coords, H, W, n_rand, patch_num, ps, crop, seed, pix):
    L.check(lib.ucn_select_training_pixels_static(L.ptr(conf), H, W, n_rand, patch_num, ps, crop, L.ptr(coords),
                                                  0 if coords is None else coords.shape[0], seed, L.ptr(pix),
                                                  L.stream_ptr()), "ucn_select_training_pixels_static")


@pytest.mark.gpu
def test_static_scratch_selection_matches_workspace_form_twice_in_a_row():
    from ucnerf_b200 import lib as L
    lib = L.load()
    assert int(lib.ucn_select_static_scratch_bytes()) == 256 + 2048 * 2 * 8
    # two calls in a row on different problems: the second finds the first one's candidates and counter in the array
    cases = [dict(H=512, W=640, n_rand=1024 - 16 * 36, patch_num=16, ps=6, n_depth=37, crop=0, seed=0x1234567890ABCDEF),
             dict(H=96, W=128, n_rand=300, patch_num=8, ps=4, n_depth=0, crop=1, seed=7)]
    prepared = []
    for i, c in enumerate(cases):
        conf, coords, R = _args(c["H"], c["W"], c["n_rand"], c["patch_num"], c["ps"], c["n_depth"], 100 + i)
        want = _workspace_form(L, lib, conf, coords, c["H"], c["W"], c["n_rand"], c["patch_num"], c["ps"], c["crop"],
                               c["seed"], R)
        prepared.append((c, conf, coords, want, torch.full((2, R), -2.0, device=DEV)))
    # warm-up call (module / memset kernel loading), then the measured pair with every buffer already in place
    c, conf, coords, want, got = prepared[0]
    _static_form(L, lib, conf, coords, c["H"], c["W"], c["n_rand"], c["patch_num"], c["ps"], c["crop"], c["seed"], got)
    torch.cuda.synchronize()
    got.fill_(-2.0)
    free0 = torch.cuda.mem_get_info()[0]
    torch_alloc0 = torch.cuda.memory_allocated()
    for c, conf, coords, want, got in prepared:
        _static_form(L, lib, conf, coords, c["H"], c["W"], c["n_rand"], c["patch_num"], c["ps"], c["crop"], c["seed"], got)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == torch_alloc0
    assert torch.cuda.mem_get_info()[0] >= free0, "the static-scratch entry point must not take device memory"
    for c, conf, coords, want, got in prepared:
        assert torch.equal(got, want), "static-scratch selection differs from the workspace form"
        rows, cols = got[0].cpu().numpy(), got[1].cpu().numpy()
        assert rows.min() >= 0 and rows.max() < c["H"] and cols.min() >= 0 and cols.max() < c["W"]
    # and once more, same problem back to back (the counter is cleared per call, not only at load)
    c, conf, coords, want, got = prepared[0]
    for _ in range(2):
        got.fill_(-2.0)
        _static_form(L, lib, conf, coords, c["H"], c["W"], c["n_rand"], c["patch_num"], c["ps"], c["crop"], c["seed"], got)
        assert torch.equal(got, want)


def test_static_scratch_selection_rejects_sizes_beyond_its_array():
    """Argument checks run before any CUDA call, so this half needs no GPU: 2048 x 2048 pixels with 128 patches needs
    2048 blocks x 64 candidates per set, 64 times what the 33 024-byte array holds."""
    from ucnerf_b200 import lib as L
    lib = L.load()
    dummy = (C.c_float * 4)()
    p = C.cast(dummy, C.c_void_p)
    rc = lib.ucn_select_training_pixels_static(p, 2048, 2048, 16, 128, 4, 0, None, 0, 1, p, None)
    assert rc != 0
    msg = lib.ucn_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "static scratch" in msg and "2048" in msg
    assert lib.ucn_select_training_pixels_static(None, 64, 64, 16, 2, 4, 0, None, 0, 1, p, None) != 0
