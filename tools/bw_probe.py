"""HBM stream probes on this GPU: write-only (fill), read-only (sum), copy -- the denominators K2 (write-dominated) and K4
(read-dominated) are up against."""
import torch
n = 1 << 30                                   # 4 GiB of fp32
a = torch.empty(n, device="cuda"); b = torch.empty(n, device="cuda")
def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
w = t(lambda: a.fill_(1.0)); print(f"write-only fill_: {4 * n / w / 1e6:.0f} GB/s")
w = t(lambda: a.zero_());    print(f"write-only zero_ (memset): {4 * n / w / 1e6:.0f} GB/s")
r = t(lambda: a.sum());      print(f"read-only sum: {4 * n / r / 1e6:.0f} GB/s")
c = t(lambda: b.copy_(a));   print(f"copy (read+write bytes): {8 * n / c / 1e6:.0f} GB/s")
