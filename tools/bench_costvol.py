"""Times the fused variance cost volume (ucn_cost_volume_variance) at the stage-3 shape of config 2 (8 channels,
8 depth hypotheses, 512x640, 3 source views) with CUDA events, reports its HBM fraction (algorithmic bytes
4*C*(V+1) + 4 per voxel) and the numpy restatement of the reference loop on the host.  One JSON line."""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
from ucnerf_b200 import mvs
from oracle import costvol as OC

C_, D, H, W, V = 8, 8, 512, 640, 3
g = torch.Generator("cuda").manual_seed(2)
feats = [torch.randn((1, C_, H, W), device="cuda", generator=g) for _ in range(V)]
K = torch.tensor([[0.9 * W, 0, W / 2, 0], [0, 0.9 * W, H / 2, 0], [0, 0, 1, 0], [0, 0, 0, 1]])
mats = [K.clone()]
for i in range(V):
    w2c = torch.eye(4); w2c[0, 3] = 0.05 * (i + 1); w2c[1, 3] = -0.02 * i
    mats.append(K @ w2c)
mats = torch.stack(mats); invs = torch.linalg.inv(mats)
dv = torch.rand((1, D, H, W), device="cuda", generator=g) + 0.6
# several depth tensors so that consecutive launches do not re-read the same lines from L2 (inputs 31 MB + 84 MB output each)
for _ in range(3):
    var = mvs.cost_volume_variance(feats, mats, invs, dv)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
times = []
for _ in range(10):
    flush.zero_()
    e0.record(); var = mvs.cost_volume_variance(feats, mats, invs, dv); e1.record(); torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
ms = float(np.median(times))
nvox = D * H * W
abytes = nvox * (4 * C_ * (V + 1) + 4)
pk = os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")
peak = json.load(open(pk)).get("hbm_gbps", 6548.5) if os.path.exists(pk) else 6548.5
projs = np.stack([(mats[i + 1:i + 2] @ invs[0:1])[0, :3].numpy() for i in range(V)])
fh = np.stack([f[0].cpu().numpy() for f in feats]); dh = dv[0].cpu().numpy()
t0 = time.perf_counter(); ref = OC.cost_volume_variance(fh, projs, dh); cpu_ms = (time.perf_counter() - t0) * 1e3
print(json.dumps({"what": "variance cost volume, stage 3 of config 2 (C=8, D=8, 512x640, V=3), L2 flushed between launches",
                  "ms_incl_host_matrix_prep": ms, "algorithmic_bytes": abytes, "GBps": abytes / ms / 1e6, "hbm_peak_GBps": peak,
                  "frac": abytes / ms / 1e6 / peak, "host_numpy_ms": cpu_ms,
                  "bit_identical_to_restatement": bool(np.array_equal(var[0].cpu().numpy().view(np.uint32), ref.view(np.uint32)))}))
