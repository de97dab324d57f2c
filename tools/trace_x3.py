"""Timeline of CTA 0 of mlp_x3_kernel (debug): per-round MMA-issue and epilogue intervals in cycles."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from ucnerf_b200 import lib as L, synthetic
from ucnerf_b200.frame import PackedWeights
lib = L.load()
V1 = os.environ.get("UCN_X3_KERNEL", "2") == "1"
set_trace = lib.ucn_debug_set_trace if V1 else lib.ucn_debug_set_trace_p      # round-1 kernel / pipelined kernel
set_trace.argtypes = [C.c_void_p]
V, S, R = 3, 126, 148 * 2 * 128 * 6 * int(os.environ.get("X3_SCALE", "1")) // 126
F = 24 + 12 * V + 1
pw = PackedWeights(synthetic.make_weights(V, seed=0), V, "cuda")
g = torch.Generator("cuda").manual_seed(0)
ndc = torch.rand((R * S, 3), device="cuda", generator=g)
NT = (R * S + 127) // 128
feat = torch.randn((NT, F, 128), device="cuda", generator=g)          # tile-transposed interchange layout
feat[:, -1, :] = torch.rand((NT, 128), device="cuda", generator=g)
if not V1:
    # operand tiles of the same rows (ucn_gather.cu): [film hi | film lo | img hi | img lo] chunks of [128][8] bf16 + confidence
    kf, kb = 24 + 4 * V, 8 * V
    cf, cb = 3 + (V + 1) // 2, V                                      # stored chunks per half
    def chunks(x, n_chunks, ones):                                    # x [NT, k, 128] fp32 -> hi, lo [NT, chunks, 128, 8] bf16
        pad = torch.zeros((NT, 8 * n_chunks, 128), device="cuda")
        pad[:, :x.shape[1]] = x
        if ones:
            pad[:, x.shape[1]] = 1.0
        hi = pad.to(torch.bfloat16)
        lo = (pad - hi.float()).to(torch.bfloat16)
        f = lambda t: t.view(NT, n_chunks, 8, 128).permute(0, 1, 3, 2).contiguous()
        return f(hi), f(lo)
    fh, fl = chunks(feat[:, :kf], cf, V % 2 == 1)
    ih, il = chunks(feat[:, kf:kf + kb], cb, False)
    rec = torch.cat([t.view(torch.uint8).reshape(NT, -1) for t in (fh, fl, ih, il)] +
                    [feat[:, -1, :].contiguous().view(torch.uint8).reshape(NT, -1)], dim=1).contiguous()
    assert rec.shape[1] == lib.ucn_operand_tile_bytes(V)
ang = torch.randn((R, 3), device="cuda", generator=g)
raw = torch.empty((R * S, 4), device="cuda")
CAP = 2048
buf = torch.zeros(7 * CAP * 2, dtype=torch.int64, device="cuda")
def run():
    if V1:
        L.check(lib.ucn_mlp_forward_tiles(C.byref(pw.c), L.ptr(ndc), L.ptr(feat), L.ptr(ang), R, S, L.ptr(raw), L.MODES["bf16"], L.stream_ptr()), "mlp")
    else:
        L.check(lib.ucn_mlp_forward_operands(C.byref(pw.c), L.ptr(ndc), L.ptr(rec), L.ptr(ang), R, S, L.ptr(raw), L.MODES["bf16"], L.stream_ptr()), "mlp")
if os.environ.get("X3_AB"):                       # same-box A/B of debug-flag variants: X3_AB="0,4" -> median ms of each, interleaved
    sets = [int(x) for x in os.environ["X3_AB"].split(",")]
    set_flags = lib.ucn_debug_set_flags if V1 else lib.ucn_debug_set_flags_p
    times = {f: [] for f in sets}
    for rep in range(7):
        for f in sets:
            L.check(set_flags(f), "flags")
            run(); torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); run(); b.record(); torch.cuda.synchronize()
            times[f].append(a.elapsed_time(b))
    L.check(set_flags(0), "flags")
    for f in sets:
        t = sorted(times[f])
        print(f"flags {f}: median {t[len(t) // 2]:.3f} ms  min {t[0]:.3f}  max {t[-1]:.3f}  ({R * S} samples)")
    sys.exit(0)
flags = int(os.environ.get("X3_DBG", "0"))
L.check((lib.ucn_debug_set_flags if V1 else lib.ucn_debug_set_flags_p)(flags), "flags")
run(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(); e1.record(); torch.cuda.synchronize()
print(f"flags {flags}: kernel {e0.elapsed_time(e1):.3f} ms for {R * S} samples -> {R * S / e0.elapsed_time(e1) / 1e3:.1f} Msamples/s")
L.check(set_trace(C.c_void_p(buf.data_ptr())), "trace")
run(); torch.cuda.synchronize()
L.check(set_trace(None), "trace")
b = buf.cpu().numpy().reshape(7, CAP, 2)
ev = []
for who in range(7):
    for w, t in b[who]:
        if t: ev.append((int(t), int(w) & 0xFF, (int(w) >> 8) & 0xFF, (int(w) >> 16) & 0xFF, int(w) >> 24))
ev.sort(); t0 = ev[0][0]
names = {1: "mma_start", 2: "mma_issued", 3: "epi_wake", 4: "epi_done", 5: "blk_issued", 6: "point", 7: "blk_waited"}
lo, hi = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (0, 200)
for t, tag, r, who, pay in ev[lo:hi]:
    extra = (f"  wfull_wait={pay & 0xFFFFF} go_wait={pay >> 20}" if not V1 else f"  wfull_wait={pay}") if tag == 2 else ""
    if tag == 7:
        extra = f"  go={pay >> 16} wfull={pay & 0xFFFF}"
    print(f"{t - t0:>8} {names[tag]:<11} r{r} {'wg' if tag in (3, 4, 6) else 'slot'}{who}{extra}")
print("events", len(ev), "span cycles", ev[-1][0] - t0)
