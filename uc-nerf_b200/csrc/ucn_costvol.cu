// ucn_costvol.cu -- the step before the render path: homography warp of the source-view
// features into the target frustum and the variance cost volume over the source views.
//
// replaces (reference): utils/utils.py:1105-1172 homo_warp (meshgrid, R @ grid + T / depth, perspective divide,
//   normalisation, F.grid_sample nearest / border / align_corners) and the per-view loop of DepthNet.forward,
//   network/mvs_models.py:603-628 (sum, sum of squares, in-image masks, variance).  The reference materialises
//   a [C][D][H][W] warped volume, its square and a [D][H][W][2] grid per view; here one thread owns one voxel
//   (d, y, x), keeps the V source texel indices in registers and streams the C channels once: V*C gathered reads
//   + C writes per voxel, nothing else touches HBM.  The 3-D regularisation CNN that consumes the volume stays
//   with cuDNN (out of scope).
//
// Arithmetic is the reference's CPU fp32 arithmetic operation for operation (the K=3 product of `R @ grid` is the
// sgemm chain fma(r2, 1, fma(r1, y, r0*x)), true divisions, no contraction anywhere else), so texel indices, masks
// and the variance are bit-exact against the reference-generated goldens.  HBM-bound gather: 4*C*(V+1) B per voxel.
#include "ucn_common.cuh"

namespace {

struct Proj { float m[UCN_MAX_V][12]; };       // row-major [3][4] per view: (src_proj @ ref_proj_inv)[:3]

// normalised source coordinates of voxel (d, row, col) (utils/utils.py:1124-1153)
__device__ __forceinline__ void warp_coord(const float *__restrict__ P, float x, float y, float depth, float half_w,
                                           float half_h, float &gx, float &gy) {
  float s[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float rot = fmaf(P[4 * i + 2], 1.0f, fmaf(P[4 * i + 1], y, __fmul_rn(P[4 * i], x)));
    s[i] = __fadd_rn(rot, __fdiv_rn(P[4 * i + 3], depth));
  }
  gx = __fsub_rn(__fdiv_rn(__fdiv_rn(s[0], s[2]), half_w), 1.0f);
  gy = __fsub_rn(__fdiv_rn(__fdiv_rn(s[1], s[2]), half_h), 1.0f);
}
// F.grid_sample(mode='nearest', padding_mode='border', align_corners=True): texel index of a normalised coordinate
__device__ __forceinline__ int nearest_border(float g, int size) {
  float p = __fmul_rn(__fmul_rn(__fadd_rn(g, 1.0f), 0.5f), (float)(size - 1));
  p = fminf(fmaxf(p, 0.0f), (float)(size - 1));
  return __float2int_rn(p);
}

__global__ void __launch_bounds__(256)
homo_warp_kernel(const float *__restrict__ src, int C, int H, int W, Proj proj, const float *__restrict__ depth,
                 const float *__restrict__ grid_in, int D, int pad, float *__restrict__ out, float *__restrict__ grid_out) {
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  const long long nvox = (long long)D * Hp * Wp;
  const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nvox) return;
  const int col = (int)(v % Wp), row = (int)((v / Wp) % Hp);
  float gx, gy;
  if (grid_in) {
    gx = grid_in[2 * v]; gy = grid_in[2 * v + 1];
  } else {
    warp_coord(proj.m[0], (float)(col - pad), (float)(row - pad), depth[v], (float)(W - 1) / 2.0f, (float)(H - 1) / 2.0f, gx, gy);
  }
  if (grid_out) { grid_out[2 * v] = gx; grid_out[2 * v + 1] = gy; }
  const long long t = (long long)nearest_border(gy, H) * W + nearest_border(gx, W);
  for (int c = 0; c < C; ++c) out[(long long)c * nvox + v] = src[(long long)c * H * W + t];
}

__global__ void __launch_bounds__(256)
cost_volume_variance_kernel(const float *__restrict__ feats, int V, int C, int H, int W, Proj proj,
                            const float *__restrict__ depth, int D, int pad, float extra_ones, float *__restrict__ var) {
  const int Hp = H + 2 * pad, Wp = W + 2 * pad;
  const long long nvox = (long long)D * Hp * Wp;
  const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nvox) return;
  const int col = (int)(v % Wp), row = (int)((v / Wp) % Hp);
  const float x = (float)(col - pad), y = (float)(row - pad), dep = depth[v];
  const float half_w = (float)(W - 1) / 2.0f, half_h = (float)(H - 1) / 2.0f;
  int texel[UCN_MAX_V];
  float n_in = extra_ones;
#pragma unroll
  for (int i = 0; i < UCN_MAX_V; ++i) {
    if (i < V) {
      float gx, gy;
      warp_coord(proj.m[i], x, y, dep, half_w, half_h, gx, gy);
      texel[i] = nearest_border(gy, H) * W + nearest_border(gx, W);
      if (gx > -1.0f && gx < 1.0f && gy > -1.0f && gy < 1.0f) n_in = __fadd_rn(n_in, 1.0f);      // mvs_models.py:616-619
    }
  }
  const float count = __fdiv_rn(1.0f, n_in);                                                      // :624
  const long long plane = (long long)H * W;
  for (int c = 0; c < C; ++c) {
    float s = 0.0f, sq = 0.0f;
#pragma unroll
    for (int i = 0; i < UCN_MAX_V; ++i) {
      if (i < V) {
        const float f = feats[((long long)i * C + c) * plane + texel[i]];
        s = __fadd_rn(s, f);                                                                      // :621-622
        sq = __fadd_rn(sq, __fmul_rn(f, f));
      }
    }
    const float m = __fmul_rn(s, count);
    var[(long long)c * nvox + v] = __fsub_rn(__fmul_rn(sq, count), __fmul_rn(m, m));              // :625
  }
}

inline int load_proj(const float *host, int V, Proj &P) {
  for (int i = 0; i < V; ++i)
    for (int k = 0; k < 12; ++k) P.m[i][k] = host[12 * i + k];
  return 0;
}

}  // namespace

// warped[c][d][y][x] = src[c][nearest texel of the warped voxel], grid_out [D][Hp*Wp][2] (either output grid may be
// NULL).  proj12_host: (src_proj @ ref_proj_inv)[:3] row-major, HOST memory (may be NULL when grid_in is given, in which
// case the coordinates are read from grid_in instead: the reference's `src_grid=` path).  depth_values [D][Hp][Wp].
extern "C" int ucn_homo_warp(const float *src_feat, int C, int H, int W, const float *proj12_host, const float *depth_values,
                             const float *grid_in, int D, int pad, float *warped, float *grid_out, void *stream) {
  UCN_CHECK_ARG(src_feat && warped && (grid_in || (proj12_host && depth_values)), "ucn_homo_warp: null argument");
  UCN_CHECK_ARG(C > 0 && H > 1 && W > 1 && D > 0 && pad >= 0, "ucn_homo_warp: bad shape");
  Proj P = {};
  if (proj12_host) load_proj(proj12_host, 1, P);
  const long long nvox = (long long)D * (H + 2 * pad) * (W + 2 * pad);
  homo_warp_kernel<<<(unsigned)((nvox + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src_feat, C, H, W, P, depth_values, grid_in,
                                                                                     D, pad, warped, grid_out);
  UCN_LAUNCH_CHECK("homo_warp_kernel");
  return UCN_OK;
}

// var[c][d][y][x] = E[f^2] - E[f]^2 over the V warped source views with the reference's count
// 1 / (number of views whose coordinate is inside the image + extra_ones).  feats [V][C][H][W], projs_host [V][12] (HOST).
extern "C" int ucn_cost_volume_variance(const float *feats, int V, int C, int H, int W, const float *projs_host,
                                        const float *depth_values, int D, int pad, int extra_ones, float *var, void *stream) {
  UCN_CHECK_ARG(feats && projs_host && depth_values && var, "ucn_cost_volume_variance: null argument");
  UCN_CHECK_ARG(V >= 1 && V <= UCN_MAX_V, "ucn_cost_volume_variance: 1 <= V <= 8");
  UCN_CHECK_ARG(C > 0 && H > 1 && W > 1 && D > 0 && pad >= 0 && extra_ones >= 0, "ucn_cost_volume_variance: bad shape");
  Proj P = {};
  load_proj(projs_host, V, P);
  const long long nvox = (long long)D * (H + 2 * pad) * (W + 2 * pad);
  cost_volume_variance_kernel<<<(unsigned)((nvox + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      feats, V, C, H, W, P, depth_values, D, pad, (float)extra_ones, var);
  UCN_LAUNCH_CHECK("cost_volume_variance_kernel");
  return UCN_OK;
}
