"""Knockout timing of the training GEMM (gemm_tc_kernel): forward product M x 128 x K with parts of the CTA disabled.
flags: 16 no pre-split B, 32 one tile per CTA (not persistent)."""
import ctypes as C, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ucnerf_b200 import lib as L
lib = L.load()
lib.ucn_debug_gemm_flags.argtypes = [C.c_int]
M = 258048
for K in (128, 64, 191):
    X = torch.randn(M, K, device="cuda"); W = torch.randn(128, K, device="cuda"); b = torch.randn(128, device="cuda")
    Y = torch.empty(M, 128, device="cuda")
    for fl in (0, 16, 32, 48, 0):
        lib.ucn_debug_gemm_flags(fl)
        def run():
            L.check(lib.ucn_linear_fwd(L.ptr(X), K, L.ptr(W), K, L.ptr(b), M, K, 128, L.ptr(Y), 128, 0, L.stream_ptr()), "fwd")
        for _ in range(3): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize()
        print(f"K={K} flags={fl:2d}: {e0.elapsed_time(e1) / 20 * 1e3:8.1f} us", flush=True)
lib.ucn_debug_gemm_flags(0)

# grad-weight product dW[128][K] += dY^T X, reduction over M split across CTAs (flags >> 12 overrides the split count)
lib.ucn_linear_bwd_weight.restype = C.c_int
for K in (128, 64):
    X = torch.randn(M, K, device="cuda"); dY = torch.randn(M, 128, device="cuda")
    dW = torch.zeros(128, K, device="cuda")
    for sp in (0, 74, 148, 222, 296, 444, 592, 888):
        lib.ucn_debug_gemm_flags((sp << 12) | 128)
        def run():
            L.check(lib.ucn_linear_bwd_weight(L.ptr(dY), 128, L.ptr(X), X.stride(0), M, K, 128, L.ptr(dW), K, None, L.stream_ptr()), "bwdw")
        for _ in range(3): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize()
        print(f"bwd_weight K={K} splits={sp:3d}: {e0.elapsed_time(e1) / 20 * 1e3:8.1f} us", flush=True)
lib.ucn_debug_gemm_flags(0)
