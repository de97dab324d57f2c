"""Times the device-resident metrics (uc-nerf_b200/metrics.py) on config-2 frames against the host evaluation the
reference does (frame -> .cpu() -> numpy; the oracle restatement of utils/evaluation.py), and reports the HBM
fraction of the three streaming kernels.  One JSON line."""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
from ucnerf_b200 import metrics as M
from oracle import metrics as OM

H, W, NUM = 512, 640, 4
g = torch.Generator("cuda").manual_seed(1)
gt = torch.rand((NUM, H, W), device="cuda", generator=g) * 60 + 0.5
pred = gt * (1 + 0.1 * torch.randn((NUM, H, W), device="cuda", generator=g)).abs() + 0.01
rgb_gt = torch.rand((NUM, H, W, 3), device="cuda", generator=g)
rgb_pr = (rgb_gt + 0.03 * torch.randn((NUM, H, W, 3), device="cuda", generator=g)).clamp(0, 1)

def gpu_eval():
    return M.depth_evaluation(gt, pred, verbose=False), M.get_psnr(list(rgb_pr), list(rgb_gt))

def host_eval():                      # what the reference's eval loop does with a rendered frame
    return (OM.depth_evaluation(gt.cpu().numpy(), pred.cpu().numpy()),
            OM.get_psnr(list(rgb_pr.cpu().numpy()), list(rgb_gt.cpu().numpy())))

for _ in range(3):
    d_gpu, p_gpu = gpu_eval()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    gpu_eval()
torch.cuda.synchronize()
gpu_ms = (time.perf_counter() - t0) * 100
d_cpu, p_cpu = host_eval()
t0 = time.perf_counter()
for _ in range(3):
    host_eval()
cpu_ms = (time.perf_counter() - t0) * 1000 / 3
# kernel-only: events around the raw launches
import ctypes as C
from ucnerf_b200 import lib as L
lib = L.load()
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {}
out8 = torch.zeros(8, dtype=torch.float64, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = gt.numel()
e0.record()
for _ in range(20):
    L.check(lib.ucn_depth_errors(L.ptr(gt), L.ptr(pred), None, n, 1e-4, 100.0, 1.0, C.c_void_p(out8.data_ptr()), L.stream_ptr()))
e1.record(); torch.cuda.synchronize()
err_ms = e0.elapsed_time(e1) / 20
print(json.dumps({"what": "depth_evaluation + get_psnr of 4 frames 640x512 (config 2)", "gpu_ms_incl_host_logic": gpu_ms,
                  "host_numpy_ms_incl_d2h": cpu_ms, "speedup": cpu_ms / gpu_ms,
                  "max_rel_diff_depth": float(np.max(np.abs(d_gpu - d_cpu) / np.maximum(np.abs(d_cpu), 1e-30))),
                  "max_abs_diff_psnr": float(np.max(np.abs(p_gpu - p_cpu))),
                  "depth_err_kernel": {"ms": err_ms, "algorithmic_bytes": 8 * n, "GBps": 8 * n / err_ms / 1e6,
                                       "note": "10.5 MB per launch: smaller than L2 and launch-latency sized; the figure is a lower bound of what the kernel does on resident frames"}}))
