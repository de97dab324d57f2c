"""A/B of the streaming training products: gemm_tc_kernel (debug flag 64) against the warp-specialised gemm_ws_kernel,
forward (M x K -> 128) and grad-input, plus a max-abs check of the two against each other and against torch fp64."""
import ctypes as C, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ucnerf_b200 import lib as L
lib = L.load()
lib.ucn_debug_gemm_flags.argtypes = [C.c_int]
M = int(os.environ.get("GEMM_M", "258048"))
for K, N in [(int(a), int(b)) for a, b in (x.split("x") for x in os.environ.get("GEMM_SHAPES", "128x128,64x128,40x128,128x256,192x128").split(","))]:
    X = torch.randn(M, K, device="cuda")                              # K % 4 == 0: 16-byte aligned rows
    W = torch.randn(N, K, device="cuda"); b = torch.randn(N, device="cuda")
    outs = {}
    for fl in (64, 0, 64, 0):
        lib.ucn_debug_gemm_flags(fl)
        Y = torch.empty(M, N, device="cuda")
        def run():
            L.check(lib.ucn_linear_fwd(L.ptr(X), X.stride(0), L.ptr(W), K, L.ptr(b), M, K, N, L.ptr(Y), N, 0, L.stream_ptr()), "fwd")
        for _ in range(3): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        gb = (M * K + M * N) * 4 / 1e9
        print(f"fwd K={K} N={N} {'old' if fl else 'ws '}: {us:8.1f} us  {gb / us * 1e6 / 1e3:6.2f} TB/s", flush=True)
        outs[fl] = Y
    ref = (X[:4096].double() @ W.double().t() + b.double()).float()
    print(f"   max|ws-old| {float((outs[0] - outs[64]).abs().max()):.3e}   max|ws-fp64| {float((outs[0][:4096] - ref).abs().max()):.3e}"
          f"   max|old-fp64| {float((outs[64][:4096] - ref).abs().max()):.3e}", flush=True)
lib.ucn_debug_gemm_flags(0)

# grad-weight dW[N][K] += dY^T X: gemm_tc_kernel with split reduction (flag 128) against the MN-major pipeline gemm_wsw_kernel
if os.environ.get("GEMM_BWDW", "1") == "1":
    for K, N, ldx in ((128, 128, 128), (64, 128, 64), (191, 128, 191), (128, 256, 128), (37, 128, 37), (155, 128, 155)):
        Xb = torch.randn(M, ldx, device="cuda"); X = Xb[:, :K]
        dY = torch.randn(M, N, device="cuda")
        outs = {}
        for fl in (128, 0, 128, 0):
            lib.ucn_debug_gemm_flags(fl)
            dW = torch.zeros(N, K, device="cuda")
            def run():
                L.check(lib.ucn_linear_bwd_weight(L.ptr(dY), N, L.ptr(Xb), ldx, M, K, N, L.ptr(dW), K, None, L.stream_ptr()), "bwdw")
            run(); torch.cuda.synchronize()
            outs[fl] = dW.clone()
            for _ in range(2): run()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20): run()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / 20 * 1e3
            gb = (M * K + M * N) * 4 / 1e9
            print(f"bwd_weight K={K} N={N} ldx={ldx} {'old' if fl else 'wsw'}: {us:8.1f} us  {gb / us * 1e6 / 1e3:6.2f} TB/s", flush=True)
        ref = (dY.double().t() @ X.double()).float()
        sc = float(ref.abs().max())
        print(f"   rel max|wsw-fp64| {float((outs[0] - ref).abs().max()) / sc:.3e}   rel max|old-fp64| {float((outs[128] - ref).abs().max()) / sc:.3e}", flush=True)
    lib.ucn_debug_gemm_flags(0)
