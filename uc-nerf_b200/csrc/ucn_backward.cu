// ucn_backward.cu -- K4b / K2b: backward of alpha compositing and of the fused gathers (training step, config 4).
//
// The reference trains through autograd (train.py:147-190): d loss / d (rgb_map, depth_map) flows back through
// raw2outputs (network/renderer.py:25-36,125-131), the MLP (ucn_mlp_train.cu) and the two grid_sample groups of
// gen_pts_feats (utils/utils.py:776-788,859-890) into the three feature volumes, the source-view FPN features and
// the confidence map.  Sample positions carry no gradient: depth hypotheses are detached
// (network/mvs_models.py:716-717), source images are data.
//
// K4b  warp per ray.  With q_i = 1 - alpha_i + 1e-10, T_i = prod_{j<i} q_j, w_i = alpha_i T_i and the per-sample
//      cotangent v_i = <g_rgb, rgb_i> + g_depth z_i (- sum g_rgb under white_bkgd):
//        d rgb_i = w_i g_rgb        d sigma_i = exp(-sigma_i) (T_i v_i - (sum_{k>i} w_k v_k) / q_i)
//      (the cumprod backward ATen uses when no factor is zero; q_i >= 1e-10).  Two passes over the ray: total of
//      w v, then the emit pass with the running prefix.
// K2b  thread per sample, the forward's index arithmetic verbatim (ucn_common.cuh), fp32 atomics straight into the
//      reference layouts (NCDHW volumes, [V][8][H][W] features, [H][W] confidence): grads need no unpacking.
#include "ucn_common.cuh"

namespace {

constexpr int kWarps = 8;

__global__ void __launch_bounds__(kWarps * 32)
composite_bwd_kernel(const float4 *__restrict__ raw, const float *__restrict__ z, int R, int S, int white_bkgd,
                     const float *__restrict__ g_rgb, const float *__restrict__ g_depth, float4 *__restrict__ d_raw) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (r >= R) return;
  const float gr = g_rgb[3 * r], gg = g_rgb[3 * r + 1], gb = g_rgb[3 * r + 2], gd = g_depth[r];
  const float gbg = white_bkgd ? -(gr + gg + gb) : 0.0f;
  float total = 0.f, prefix = 0.f;
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
    float T = 1.0f;
    for (int base = 0; base < S; base += 32) {
      const int i = base + lane;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      float zi = 0.f;
      if (i < S) { s = __ldg(raw + (size_t)r * S + i); zi = __ldg(z + (size_t)r * S + i); }
      const float e = i < S ? expf(-s.w) : 1.0f;
      const float alpha = i < S ? 1.0f - e : 0.0f;
      const float qi = i < S ? (1.0f - alpha) + 1e-10f : 1.0f;
      float p = qi;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float up = __shfl_up_sync(0xffffffffu, p, o);
        if (lane >= o) p *= up;
      }
      float excl = __shfl_up_sync(0xffffffffu, p, 1);
      if (lane == 0) excl = 1.0f;
      const float Ti = T * excl, w = alpha * Ti;
      const float v = gr * s.x + gg * s.y + gb * s.z + gd * zi + gbg;
      float wv = w * v, inc = wv;                       // inclusive sum of w v over the group
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
      }
      if (pass == 0) {
        total += __shfl_sync(0xffffffffu, inc, 31);
      } else {
        const float suffix = total - (prefix + inc);    // sum_{k>i} w_k v_k
        if (i < S) d_raw[(size_t)r * S + i] = make_float4(w * gr, w * gg, w * gb, e * (Ti * v - suffix / qi));
        prefix += __shfl_sync(0xffffffffu, inc, 31);
      }
      T *= __shfl_sync(0xffffffffu, p, 31);
    }
  }
}

__device__ __forceinline__ void red(float *p, float v) { if (v != 0.0f) atomicAdd(p, v); }

// d_vol NCDHW [8][D][h][w]; g[c * gs] = cotangent of channel c
__device__ __forceinline__ void tri8_bwd(float *__restrict__ d_vol, int D, int h, int w, float cx, float cy, float cz,
                                         const float *__restrict__ g) {
  const ucn::Lerp1 ax = ucn::axis(ucn::to_grid(cx), w, false);
  const ucn::Lerp1 ay = ucn::axis(ucn::to_grid(cy), h, false);
  const ucn::Lerp1 az = ucn::axis(ucn::to_grid(cz), D, false);
  const size_t plane = (size_t)h * w, chan = plane * D;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int dx = c & 1, dy = (c >> 1) & 1, dz = c >> 2;
    if ((dx && !ax.has1) || (dy && !ay.has1) || (dz && !az.has1)) continue;      // ATen within_bounds test
    const float wgt = __fmul_rn(__fmul_rn(dx ? ax.w1 : ax.w0, dy ? ay.w1 : ay.w0), dz ? az.w1 : az.w0);
    if (wgt == 0.0f) continue;
    float *t = d_vol + (size_t)(az.i0 + dz) * plane + (size_t)(ay.i0 + dy) * w + (ax.i0 + dx);
#pragma unroll
    for (int k = 0; k < 8; ++k) red(t + k * chan, wgt * g[k]);
  }
}

__global__ void __launch_bounds__(128)
gather_bwd_kernel(const __grid_constant__ ucn_frame f, const float *__restrict__ pts, const float *__restrict__ ndc1,
                  const float *__restrict__ ndc2, const float *__restrict__ ndc3, int64_t n,
                  const float *__restrict__ d_feat, float *d_vol0, float *d_vol1, float *d_vol2, float *d_img_feat,
                  float *d_conf) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n) return;
  const int V = f.V, F = 24 + 12 * V + 1;
  const float *g = d_feat + q * F;
  float gv[8];
  if (d_vol0) {
#pragma unroll
    for (int k = 0; k < 8; ++k) gv[k] = g[k];
    tri8_bwd(d_vol0, f.D[0], f.h[0], f.w[0], ndc1[3 * q], ndc1[3 * q + 1], ndc1[3 * q + 2], gv);
  }
  if (d_vol1) {
#pragma unroll
    for (int k = 0; k < 8; ++k) gv[k] = g[8 + k];
    tri8_bwd(d_vol1, f.D[1], f.h[1], f.w[1], ndc2[3 * q], ndc2[3 * q + 1], ndc2[3 * q + 2], gv);
  }
  if (d_vol2) {
#pragma unroll
    for (int k = 0; k < 8; ++k) gv[k] = g[16 + k];
    tri8_bwd(d_vol2, f.D[2], f.h[2], f.w[2], ndc3[3 * q], ndc3[3 * q + 1], ndc3[3 * q + 2], gv);
  }
  if (d_conf) {
    const ucn::Lerp1 ax = ucn::axis(ucn::to_grid(ndc3[3 * q]), f.W, false), ay = ucn::axis(ucn::to_grid(ndc3[3 * q + 1]), f.H, false);
    const float gc = g[F - 1];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int dx = c & 1, dy = c >> 1;
      if ((dx && !ax.has1) || (dy && !ay.has1)) continue;
      red(d_conf + (size_t)(ay.i0 + dy) * f.W + (ax.i0 + dx), gc * __fmul_rn(dx ? ax.w1 : ax.w0, dy ? ay.w1 : ay.w0));
    }
  }
  if (d_img_feat) {
    const float iw = (float)(f.W - 1), ih = (float)(f.H - 1);
    const float p[3] = {pts[3 * q], pts[3 * q + 1], pts[3 * q + 2]};
    const size_t HW = (size_t)f.H * f.W;
    for (int v = 0; v < V; ++v) {
      float nx, ny, pz;
      ucn::project(f.src_w2c + 16 * v, f.src_K + 9 * v, iw, ih, p, nx, ny, pz);
      const ucn::Lerp1 ax = ucn::axis(ucn::to_grid(nx), f.W, true), ay = ucn::axis(ucn::to_grid(ny), f.H, true);
      const float *gf = g + 24 + 4 * V + 8 * v;
      float *dst = d_img_feat + (size_t)v * 8 * HW;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int dx = c & 1, dy = c >> 1;
        if ((dx && !ax.has1) || (dy && !ay.has1)) continue;
        const float wgt = __fmul_rn(dx ? ax.w1 : ax.w0, dy ? ay.w1 : ay.w0);
        if (wgt == 0.0f) continue;
        float *t = dst + (size_t)(ay.i0 + dy) * f.W + (ax.i0 + dx);
#pragma unroll
        for (int k = 0; k < 8; ++k) red(t + k * HW, wgt * gf[k]);
      }
    }
  }
}

}  // namespace

extern "C" int ucn_composite_bwd(const float *raw, const float *z, int R, int S, int white_bkgd, const float *g_rgb,
                                 const float *g_depth, float *d_raw, void *stream) {
  UCN_CHECK_ARG(raw && z && g_rgb && g_depth && d_raw, "ucn_composite_bwd: null argument");
  UCN_CHECK_ARG(S > 0, "ucn_composite_bwd: S <= 0");
  if (R <= 0) return UCN_OK;
  composite_bwd_kernel<<<(R + kWarps - 1) / kWarps, kWarps * 32, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4 *>(raw), z, R, S, white_bkgd, g_rgb, g_depth, reinterpret_cast<float4 *>(d_raw));
  UCN_LAUNCH_CHECK("composite_bwd_kernel");
  return UCN_OK;
}

extern "C" int ucn_gather_bwd(const ucn_frame *f, const float *pts, const float *ndc_s1, const float *ndc_s2,
                              const float *ndc_s3, int64_t n_samples, const float *d_feat, float *d_vol0, float *d_vol1,
                              float *d_vol2, float *d_img_feat, float *d_conf, void *stream) {
  UCN_CHECK_ARG(f && d_feat && ndc_s1 && ndc_s2 && ndc_s3, "ucn_gather_bwd: null argument");
  UCN_CHECK_ARG(!d_img_feat || pts, "ucn_gather_bwd: world points are needed for the source-view gradients");
  UCN_CHECK_ARG(f->V >= 1 && f->V <= UCN_MAX_V, "ucn_gather_bwd: 1 <= V <= 8 source views");
  if (n_samples <= 0) return UCN_OK;
  gather_bwd_kernel<<<(unsigned)((n_samples + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      *f, pts, ndc_s1, ndc_s2, ndc_s3, n_samples, d_feat, d_vol0, d_vol1, d_vol2, d_img_feat, d_conf);
  UCN_LAUNCH_CHECK("gather_bwd_kernel");
  return UCN_OK;
}
