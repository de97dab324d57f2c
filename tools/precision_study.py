"""Scratch study (not product code): which operand precision does the tensor-core MLP need to meet the
1e-3 per-pixel RGB gate?  Emulates operand rounding in numpy (fp64 accumulate) on the golden cases."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from helpers import load_case
from oracle import oracle as O


def rnd(x, bits):
    """round fp32 to `bits` explicit mantissa bits (7 = bf16, 10 = fp16/tf32, 23 = fp32)"""
    if bits >= 23:
        return x.astype(np.float32)
    x = np.ascontiguousarray(x, np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    sh = 23 - bits
    u = (u + (1 << (sh - 1)) - 1 + ((u >> sh) & 1)) >> sh << sh
    return u.astype(np.uint32).view(np.float32)


def pe(x, L):
    out = [x]
    for f in range(L):
        out.append(np.sin(x * np.float32(2.0 ** f)))
    for f in range(L):
        out.append(np.cos(x * np.float32(2.0 ** f)))
    return np.concatenate(out, -1).astype(np.float32)


GROUPS = {"pts_bias_depth_fine": "film", "pts_bias_confidence": "fb", "feature_linear": "feat",
          "views_linears.0": "vv", "view_confi_linears.0": "vv"}


def mlp(sd, V, ndc, feat, angle, qa, qw, qh=None, per_group=None):
    """qa/qw: mantissa bits of activation/weight operands; qh: bits for the tiny head layers;
    per_group: {"film"|"trunk"|"fb"|"feat"|"vv": (qa, qw)} overrides per layer group (K3's MMA rounds)"""
    qh = qh or (qa, qw)
    per_group = per_group or {}
    def lin(name, x, a=None, w=None):
        grp = GROUPS.get(name, "trunk" if name.startswith("pts_linears") else None)
        ga, gw = per_group.get(grp, (qa, qw))
        a, w = (ga if a is None else a), (gw if w is None else w)
        Wm, b = sd[f"nerf.{name}.weight"], sd[f"nerf.{name}.bias"]
        return (rnd(x, a).astype(np.float64) @ rnd(Wm, w).astype(np.float64).T + b).astype(np.float32)
    N = ndc.shape[0]
    p63 = pe(ndc, 10)
    p27 = pe(angle, 4)
    kf = 24 + 4 * V
    mod = lin("pts_bias_depth_fine", feat[:, :kf])
    fb = lin("pts_bias_confidence", feat[:, kf:kf + 8 * V])
    u = 1.0 - feat[:, -1:]
    h = p63
    for i in range(6):
        h = np.maximum(lin(f"pts_linears.{i}", h) * mod, 0)
        if i == 4:
            h = np.concatenate([p63, h], -1)
    base_rgb = lin("confi_rgb_linear", h, *qh)
    base_s = lin("alpha_linear_1", h, *qh)
    ft = lin("feature_linear", h * fb)
    h1 = np.concatenate([ft, p27], -1)
    a_rgb = lin("rgb_linear", np.maximum(lin("views_linears.0", h1), 0), *qh)
    a_s = lin("alpha_linear", np.maximum(lin("view_confi_linears.0", h1), 0), *qh)
    rgb = 1 / (1 + np.exp(-(base_rgb * (1 - u) + a_rgb * u)))
    s = np.maximum(a_s * (1 - u) + base_s * u, 0)
    return np.concatenate([rgb, s], -1).astype(np.float32)


def main():
    for name in sys.argv[1:] or ["small_v3", "s126_v3"]:
        c = load_case(name)
        g, sc = c["g"], c["scene"]
        H, W, S, V = c["H"], c["W"], c["S"], c["V"]
        osc = O.OracleScene(sc, S, dir_switch=c["chunk"])
        py, px = O.chunk_pixels(H, W)
        o = O.build_rays_test(osc, py, px, c["t_rand"])
        feat = O.gen_pts_feats(osc, o["pts"], o["ndc_all"])
        R = feat.shape[0]
        ang = np.concatenate([O.dir_feature(osc, o["rays_d"][:c["chunk"]], 0),
                              O.dir_feature(osc, o["rays_d"][c["chunk"]:], c["chunk"])], 0) if R > c["chunk"] else O.dir_feature(osc, o["rays_d"], 0)
        ndc = o["ndc"]["ndc"].reshape(-1, 3)
        angs = np.repeat(ang, S, 0)
        F = feat.reshape(-1, feat.shape[-1])
        ref_raw = mlp(c["weights"], V, ndc, F, angs, 23, 23)
        rgb0, _ = O.raw2outputs(ref_raw.reshape(R, S, 4), o["z"])
        print(name, "fp32-emu vs golden rgb", np.abs(rgb0 - g["rgb"]).max())
        for label, qa, qw, qh in (("bf16 A, bf16 W", 7, 7, None), ("fp16 A, fp16 W", 10, 10, None),
                                  ("bf16x2 A(15), bf16 W", 15, 7, None), ("bf16x2 A, bf16x2 W (15,15)", 15, 15, None),
                                  ("fp16 A, fp16x2 W", 10, 21, None), ("fp16x2 A, fp16 W", 21, 10, None),
                                  ("fp16x2 both", 21, 21, None), ("tf32-ish 13", 13, 13, None),
                                  ("12/12", 12, 12, None), ("14/14", 14, 14, None)):
            raw = mlp(c["weights"], V, ndc, F, angs, qa, qw, qh)
            rgb, dep = O.raw2outputs(raw.reshape(R, S, 4), o["z"])
            e = np.abs(rgb - g["rgb"])
            print(f"  {label:32s} rgb max {e.max():.2e} mean {e.mean():.2e} p99 {np.quantile(e, .99):.2e}  raw max {np.abs(raw-ref_raw).max():.2e}")
        # per-layer-group table: everything split-bf16 (15/15 = three MMA passes) except ONE group (or the whole tail)
        # at a cheaper operand format.  Heads stay fp32 (CUDA cores in K3).  passes: (7,7)/(10,10) = 1, (15,7)/(10,21) = 2
        fmts = (("bf16 1-pass", (7, 7)), ("fp16 1-pass", (10, 10)), ("bf16x2 A, bf16 W (2)", (15, 7)),
                ("bf16 A, bf16x2 W (2)", (7, 15)), ("fp16 A, fp16x2 W (2)", (10, 21)))
        print("  -- one group (or the tail fb+feat+vv) cheaper, rest split-bf16; heads fp32 --")
        for grp in ("film", "trunk", "fb", "feat", "vv", "tail"):
            for label, q in fmts:
                pg = {g_: q for g_ in (("fb", "feat", "vv") if grp == "tail" else (grp,))}
                raw = mlp(c["weights"], V, ndc, F, angs, 15, 15, (23, 23), pg)
                rgb, dep = O.raw2outputs(raw.reshape(R, S, 4), o["z"])
                e = np.abs(rgb - g["rgb"])
                print(f"  {grp:6s} {label:24s} rgb max {e.max():.2e} p99 {np.quantile(e, .99):.2e}  raw max {np.abs(raw-ref_raw).max():.2e}")


if __name__ == "__main__":
    main()
