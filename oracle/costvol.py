"""CPU restatement of the encoder's variance cost volume (TEST INFRASTRUCTURE ONLY -- never imported by the product).

numpy; follows utils/utils.py:1105-1172 (homo_warp) and network/mvs_models.py:603-628 (the per-view loop of
DepthNet.forward).  Pinned by tests/golden/cv_*.npz, which tests/golden/make_golden_costvol.py produced by running the
reference's own DepthNet.forward / homo_warp.
"""
import numpy as np

F32 = np.float32


def _fma(a, b, c):
    """fp32 fused multiply-add (the fp32 x fp32 product is exact in fp64)."""
    return (np.float64(a) * np.float64(b) + np.float64(c)).astype(F32)


def warp_grid(proj, depth_values, H, W, pad=0):
    """Normalised source coordinates [D][Hp][Wp][2] (utils/utils.py:1124-1158).  proj [3][4] fp32 = (src @ ref^-1)[:3];
    depth_values [D][Hp][Wp] (already padded).  fp32 throughout; the K=3 product of `R @ ref_grid` is the sgemm
    micro-kernel's chain  fma(r2, 1, fma(r1, y, r0 * x))  -- the order that reproduces the reference's grid bit for bit
    on all goldens (left-to-right fp32 sums match only 92 %)."""
    D, Hp, Wp = depth_values.shape
    ys, xs = np.meshgrid(np.arange(Hp, dtype=F32) - F32(pad), np.arange(Wp, dtype=F32) - F32(pad), indexing="ij")
    R, T = proj[:, :3].astype(F32), proj[:, 3].astype(F32)
    out = np.empty((3, D, Hp, Wp), F32)
    for i in range(3):
        rot = _fma(R[i, 2], F32(1), _fma(R[i, 1], ys, (R[i, 0] * xs).astype(F32)))      # R @ [x y 1]  (:1138)
        out[i] = rot[None].astype(F32) + (T[i] / depth_values).astype(F32)              # + T / depth
    gx = (out[0] / out[2]).astype(F32) / F32((W - 1) / 2) - F32(1)                      # :1149-1153
    gy = (out[1] / out[2]).astype(F32) / F32((H - 1) / 2) - F32(1)
    return np.stack([gx, gy], -1).astype(F32)


def sample_nearest_border(src, grid):
    """F.grid_sample(mode='nearest', padding_mode='border', align_corners=True) (:1162-1166).  src [C][H][W],
    grid [..., 2] -> [C][...]."""
    C, H, W = src.shape
    x = ((grid[..., 0] + F32(1)) / F32(2)) * F32(W - 1)
    y = ((grid[..., 1] + F32(1)) / F32(2)) * F32(H - 1)
    x = np.clip(x, F32(0), F32(W - 1))
    y = np.clip(y, F32(0), F32(H - 1))
    xi = np.rint(x).astype(np.int64)                                                    # nearbyint: half to even
    yi = np.rint(y).astype(np.int64)
    return src[:, yi, xi]


def homo_warp(src_feat, proj, depth_values, pad=0):
    """-> (warped [C][D][Hp][Wp], grid [D][Hp][Wp][2])."""
    _, H, W = src_feat.shape
    grid = warp_grid(proj, depth_values, H, W, pad)
    return sample_nearest_border(src_feat, grid), grid


def cost_volume_variance(feats, projs, depth_values, pad=0, extra_ones=1):
    """network/mvs_models.py:603-628.  feats [V][C][H][W], projs [V][3][4], depth_values [D][H][W] (unpadded: it is
    replicate-padded here like :597-598).  The reference's in_masks has one slot more than there are views and the
    unused slot stays 1 (`extra_ones`), so count = 1 / (sum of in-image masks + 1)."""
    if pad > 0:
        depth_values = np.pad(depth_values, ((0, 0), (pad, pad), (pad, pad)), mode="edge")
    V, C, H, W = feats.shape
    s = np.zeros((C,) + depth_values.shape, F32)
    sq = np.zeros_like(s)
    n_in = np.full(depth_values.shape, F32(extra_ones), F32)
    for i in range(V):
        w, grid = homo_warp(feats[i], projs[i], depth_values, pad)
        inm = (grid > -1.0) & (grid < 1.0)
        n_in += (inm[..., 0] & inm[..., 1]).astype(F32)
        s = s + w
        sq = sq + w ** 2
    count = (F32(1) / n_in)[None]
    return sq * count - (s * count) ** 2
