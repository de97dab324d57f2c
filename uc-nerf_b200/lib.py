"""ctypes binding of libucnerf.so (include/ucnerf.h).  Fails loudly: no library, no product."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libucnerf.so")

UCN_MAX_V = 8
UCN_MAX_S = 1024
MLP_FP32, MLP_BF16_TC, MLP_BF16X3_TC = 0, 1, 2
PART_VOLUMES, PART_VIEWS, PART_ALL = 1, 2, 3
# "bf16" is the production tensor-core mode: split-bf16 operands (meets the 1e-3 RGB gate);
# "bf16_fast" is single-pass bf16 (about 2e-2 worst-case RGB error on random-init weights), opt-in only.
MODES = {"fp32": MLP_FP32, "bf16": MLP_BF16X3_TC, "bf16_fast": MLP_BF16_TC}

_fp = C.c_void_p


class Frame(C.Structure):
    _fields_ = [("H", C.c_int32), ("W", C.c_int32), ("V", C.c_int32), ("S", C.c_int32),
                ("D", C.c_int32 * 3), ("h", C.c_int32 * 3), ("w", C.c_int32 * 3),
                ("K", C.c_float * 9), ("c2w", C.c_float * 16), ("w2c", C.c_float * 16),
                ("near", C.c_float), ("far", C.c_float),
                ("dir_rot0", C.c_float * 9), ("dir_rot1", C.c_float * 9),
                ("dir_switch", C.c_int32), ("reserved", C.c_int32),
                ("src_w2c", C.c_float * (UCN_MAX_V * 16)), ("src_K", C.c_float * (UCN_MAX_V * 9)),
                ("nearfar", _fp * 3), ("vol", _fp * 3), ("conf", _fp), ("rgb", _fp), ("feat", _fp)]


class Weights(C.Structure):
    _fields_ = [("V", C.c_int32), ("reserved", C.c_int32), ("f32", _fp), ("bf16", _fp), ("bf16x3", _fp)]


class StateDict(C.Structure):
    _fields_ = [("V", C.c_int32), ("reserved", C.c_int32), ("pts_w", _fp * 6), ("pts_b", _fp * 6)] + [
        (n + s, _fp) for n in ("film", "fb", "feature", "views", "vconf", "rgb", "alpha", "crgb", "alpha1")
        for s in ("_w", "_b")]


class UcnError(RuntimeError):
    pass


_lib = None

_SIGS = {
    "ucn_version": (C.c_int, []),
    "ucn_last_error": (C.c_char_p, []),
    "ucn_launch_count": (C.c_int64, [C.c_int]),
    "ucn_check_device": (C.c_int, []),
    "ucn_pack_channels_last": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int64, C.c_int, _fp]),
    "ucn_pack_volume": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int, _fp]),
    "ucn_pack_nearfar": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int, _fp]),
    "ucn_pack_volume_rows": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _fp]),
    "ucn_pack_nearfar_rows": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _fp]),
    "ucn_h2d_rows": (C.c_int, [_fp, _fp, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _fp]),
    "ucn_weights_f32_bytes": (C.c_size_t, [C.c_int]),
    "ucn_weights_bf16_bytes": (C.c_size_t, [C.c_int]),
    "ucn_weights_bf16x3_bytes": (C.c_size_t, [C.c_int]),
    "ucn_pack_weights": (C.c_int, [C.POINTER(StateDict), _fp, _fp, _fp, _fp]),
    "ucn_raygen_sample": (C.c_int, [C.POINTER(Frame), _fp, C.c_int, C.c_int, _fp, _fp, C.c_uint64, _fp, _fp, _fp,
                                    _fp, _fp]),
    "ucn_ray_marcher": (C.c_int, [_fp, C.c_int, C.c_int, _fp, _fp, C.c_float, C.c_int, _fp, _fp, _fp]),
    "ucn_gather": (C.c_int, [C.POINTER(Frame), _fp, _fp, _fp, _fp, C.c_int64, C.c_int, _fp, _fp, _fp]),
    "ucn_project_points": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int, C.c_int, _fp, C.c_int64,
                                     _fp, _fp]),
    "ucn_gather_from_z": (C.c_int, [C.POINTER(Frame), _fp, _fp, C.c_int, C.c_int, _fp, _fp, _fp]),
    "ucn_gather_tiles_from_z": (C.c_int, [C.POINTER(Frame), _fp, _fp, C.c_int, C.c_int, _fp, _fp, _fp]),
    "ucn_mlp_forward_tiles": (C.c_int, [C.POINTER(Weights), _fp, _fp, _fp, C.c_int, C.c_int, _fp, C.c_int, _fp]),
    "ucn_operand_tile_bytes": (C.c_size_t, [C.c_int]),
    "ucn_gather_operands_from_z": (C.c_int, [C.POINTER(Frame), _fp, _fp, C.c_int, C.c_int, _fp, _fp, _fp]),
    "ucn_mlp_forward_operands": (C.c_int, [C.POINTER(Weights), _fp, _fp, _fp, C.c_int, C.c_int, _fp, C.c_int, _fp]),
    "ucn_mlp_x3_variant": (C.c_int, [C.c_int]),
    "ucn_dir_feature": (C.c_int, [C.POINTER(Frame), _fp, C.c_int, C.c_int, _fp, _fp]),
    "ucn_mlp_forward": (C.c_int, [C.POINTER(Weights), _fp, _fp, _fp, C.c_int, C.c_int, _fp, C.c_int, _fp]),
    "ucn_composite": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int, _fp, _fp, _fp, _fp, _fp]),
    "ucn_composite_bwd": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int, _fp, _fp, _fp, _fp]),
    "ucn_gather_bwd": (C.c_int, [C.POINTER(Frame), _fp, _fp, _fp, _fp, C.c_int64, _fp, _fp, _fp, _fp, _fp, _fp, _fp]),
    "ucn_train_gemm_mode": (C.c_int, [C.c_int]),
    "ucn_linear_fwd": (C.c_int, [_fp, C.c_int64, _fp, C.c_int64, _fp, C.c_int64, C.c_int, C.c_int, _fp, C.c_int64,
                                 C.c_int, _fp]),
    "ucn_linear_bwd_input": (C.c_int, [_fp, C.c_int64, _fp, C.c_int64, C.c_int64, C.c_int, C.c_int, _fp, C.c_int64,
                                       C.c_int, _fp]),
    "ucn_linear_bwd_weight": (C.c_int, [_fp, C.c_int64, _fp, C.c_int64, C.c_int64, C.c_int, C.c_int, _fp, C.c_int64,
                                        _fp, _fp]),
    "ucn_train_prep": (C.c_int, [_fp, _fp, C.c_int, C.c_int64, _fp, C.c_int64, _fp, C.c_int64, _fp]),
    "ucn_film_relu_fwd": (C.c_int, [_fp, _fp, _fp, C.c_int64, C.c_int64, _fp]),
    "ucn_film_relu_bwd": (C.c_int, [_fp, C.c_int64, _fp, _fp, _fp, _fp, C.c_int, _fp, C.c_int64, _fp]),
    "ucn_mul_fwd": (C.c_int, [_fp, C.c_int64, _fp, _fp, C.c_int64, _fp]),
    "ucn_mul_bwd": (C.c_int, [_fp, _fp, C.c_int64, _fp, _fp, C.c_int, _fp, C.c_int64, _fp]),
    "ucn_relu_bwd": (C.c_int, [_fp, _fp, _fp, C.c_int64, _fp]),
    "ucn_blend_fwd": (C.c_int, [_fp, _fp, _fp, C.c_int64, _fp, C.c_int64, _fp]),
    "ucn_blend_bwd": (C.c_int, [_fp, _fp, _fp, _fp, _fp, C.c_int64, _fp, _fp, _fp, C.c_int64, C.c_int64, _fp]),
    "ucn_frame_workspace_bytes": (C.c_size_t, [C.POINTER(Frame), C.c_int, C.c_int]),
    "ucn_render_rays": (C.c_int, [C.POINTER(Frame), C.POINTER(Weights), C.c_int, C.c_int, _fp, _fp, C.c_uint64,
                                  C.c_int, _fp, C.c_size_t, _fp, _fp, _fp]),
    "ucn_render_losses": (C.c_int, [_fp, _fp, _fp, _fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int, _fp, _fp, _fp, C.c_int,
                                    C.c_int, _fp, _fp, _fp, _fp]),
    "ucn_select_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "ucn_select_training_pixels": (C.c_int, [_fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _fp, C.c_int,
                                             C.c_uint64, _fp, _fp, _fp]),
    "ucn_select_static_scratch_bytes": (C.c_size_t, []),
    "ucn_select_training_pixels_static": (C.c_int, [_fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _fp,
                                                    C.c_int, C.c_uint64, _fp, _fp]),
    "ucn_homo_warp": (C.c_int, [_fp, C.c_int, C.c_int, C.c_int, _fp, _fp, _fp, C.c_int, C.c_int, _fp, _fp, _fp]),
    "ucn_cost_volume_variance": (C.c_int, [_fp, C.c_int, C.c_int, C.c_int, C.c_int, _fp, _fp, C.c_int, C.c_int, C.c_int,
                                           _fp, _fp]),
    "ucn_metrics_workspace_bytes": (C.c_size_t, []),
    "ucn_sq_err_sum": (C.c_int, [_fp, _fp, C.c_int64, _fp, _fp]),
    "ucn_ssim_sum": (C.c_int, [_fp, _fp, C.c_int, C.c_int, C.c_int, C.c_float, _fp, _fp]),
    "ucn_depth_median": (C.c_int, [_fp, _fp, _fp, C.c_int64, C.c_float, C.c_float, _fp, C.c_size_t, _fp, _fp, _fp]),
    "ucn_depth_errors": (C.c_int, [_fp, _fp, _fp, C.c_int64, C.c_float, C.c_float, C.c_float, _fp, _fp]),
    "ucn_peer_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]),
    "ucn_peer_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "ucn_peer_close": (C.c_int, [_fp]),
    "ucn_peer_free": (C.c_int, [_fp]),
}


def exported_symbols():
    """Every symbol include/ucnerf.h declares (used by the CPU-side ABI test)."""
    return sorted(_SIGS)


def load():
    """dlopen libucnerf.so and attach signatures.  Raises if the extension is missing: there is no
    CPU or PyTorch fallback anywhere in this package."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise UcnError(f"{LIB_PATH} not built -- run `python -m ucnerf_b200.build` (no fallback path exists)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().ucn_last_error().decode(errors="replace")
        raise UcnError(f"{what or 'libucnerf'} failed (code {rc}): {msg}")


_PEER_RANGES = []          # [(lo, hi)] address ranges of peer-mapped buffers (shard.PeerFrameBuffer)


def ptr(t):
    """Device pointer of a contiguous fp32/int32 CUDA tensor (or None).  The tensor must live on the CURRENT device
    (the kernels launch on its stream and use its attributes) unless it is a registered peer mapping."""
    if t is None:
        return None
    if not t.is_cuda:
        raise UcnError("libucnerf has no CPU path: tensor must live on a CUDA device")
    if not t.is_contiguous():
        raise UcnError("tensor must be contiguous")
    import torch
    a = t.data_ptr()
    if t.device.index != torch.cuda.current_device() and not any(lo <= a < hi for lo, hi in _PEER_RANGES):
        raise UcnError(f"tensor lives on {t.device} but the current device is cuda:{torch.cuda.current_device()}: "
                       "wrap the call in `with torch.cuda.device(t.device)`")
    return C.c_void_p(a)


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count(reset: bool = False) -> int:
    return int(load().ucn_launch_count(1 if reset else 0))
