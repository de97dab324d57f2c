/*
 * ucnerf_oracle.c -- CPU restatement of the UC-NeRF per-ray rendering hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and only as the checker / CPU baseline.  The product path (uc-nerf_b200/) never calls it.
 *
 * Parity pin: every function below is checked bit-for-bit (positions) or to fp32 tolerance
 * (features, MLP, compositing) against outputs of the reference's own Python code, executed in
 * the build container by tests/golden/make_golden.py and committed under tests/golden/.
 *
 * Each function cites the reference lines (paths relative to the reference repo) it follows.
 * Arithmetic order is part of the contract: this file is compiled with -ffp-contract=off and
 * uses fmaf() exactly where the reference's sgemm (MKL / cuBLAS, K=3) fuses multiply-adds.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "ucnerf_oracle.h"

/* [N,3] x [3,3]^T product as torch.matmul computes it (k-ordered FMA chain; verified against
 * torch CPU/MKL in tests/golden/make_golden.py).  M is row-major, y_j = sum_k x_k * M[j][k]. */
static inline void mm3t(const float *M, int ld, const float x[3], float y[3]) {
  for (int j = 0; j < 3; ++j) {
    float acc = x[0] * M[j * ld + 0];
    acc = fmaf(x[1], M[j * ld + 1], acc);
    acc = fmaf(x[2], M[j * ld + 2], acc);
    y[j] = acc;
  }
}

static int cmp_float(const void *a, const void *b) {
  float x = *(const float *)a, y = *(const float *)b;
  return (x > y) - (x < y);
}

/* world -> camera -> pixel -> [0,1] normalised xy (+ raw projective depth).
 * reference: utils/utils.py:333-347 (get_ndc_coordinate). */
static inline void project(const float *w2c, const float *K, float iw, float ih, const float p[3],
                           float *nx, float *ny, float *pz) {
  float pc[3], pp[3];
  mm3t(w2c, 4, p, pc);                       /* matmul(points, R.t()) */
  pc[0] = pc[0] + w2c[3];                    /* + T */
  pc[1] = pc[1] + w2c[7];
  pc[2] = pc[2] + w2c[11];
  if (fabsf(pc[2]) < 1e-4f) pc[2] = 1e-4f;   /* :341-342 */
  mm3t(K, 3, pc, pp);                        /* @ intrinsic.t() */
  *nx = (pp[0] / pp[2] + 0.0f) / iw;         /* :345-347, inv_scale = (W-1, H-1) */
  *ny = (pp[1] / pp[2] + 0.0f) / ih;
  *pz = pp[2];
}

/* ------------------------------------------------------------------------------------------
 * Ray generation + depth-bounded sampling + target-view NDC.
 * reference: utils/utils.py:248-268 (get_rays_mvs), :393-397 (sample_points_uniform),
 *            :639-739 (build_rays_test), :323-373 (get_ndc_coordinate).
 * pix_y/pix_x are the float pixel coordinates the reference slices from its meshgrid.
 * ------------------------------------------------------------------------------------------ */
void uco_raygen_sample(const uco_scene *sc, const float *pix_y, const float *pix_x, int R,
                       const float *t_vals, const float *t_rand, float *z_out, float *rays_d_out,
                       float *pts_out, float *ndc_out /* [4][R][S][3] or NULL */) {
  const int S = sc->S, n = S / 3;
  const float iw = (float)(sc->W - 1), ih = (float)(sc->H - 1);
#pragma omp parallel for schedule(static)
  for (int r = 0; r < R; ++r) {
    float zs[UCO_MAX_S];
    const float ys = pix_y[r], xs = pix_x[r];
    float dirs[3] = {(xs - sc->K[2]) / sc->K[0], (ys - sc->K[5]) / sc->K[4], 1.0f};
    float d[3];
    mm3t(sc->c2w, 4, dirs, d);                       /* dirs @ c2w[:3,:3].t() */
    const float o[3] = {sc->c2w[3], sc->c2w[7], sc->c2w[11]};
    const long py = (long)ys, px = (long)xs;           /* .long() */
    float nr[3], fr[3];
    for (int k = 0; k < 3; ++k) {                      /* trunc-div pixel index, :659-683 */
      const int div = k == 0 ? 4 : (k == 1 ? 2 : 1);
      const long qy = py / div, qx = px / div;
      const long hw = (long)sc->h[k] * sc->w[k];
      nr[k] = sc->depth[k][qy * sc->w[k] + qx];
      fr[k] = sc->depth[k][(long)(sc->D[k] - 1) * hw + qy * sc->w[k] + qx];
      for (int i = 0; i < n; ++i) {                    /* near*(1-t) + far*t, :396 */
        const float t = t_vals[i];
        zs[k * n + i] = nr[k] * (1.0f - t) + fr[k] * t;
      }
    }
    qsort(zs, (size_t)S, sizeof(float), cmp_float);    /* torch.sort, :700 */
    float zj[UCO_MAX_S];
    for (int i = 0; i < S; ++i) {                      /* stratified jitter, :701-706 */
      const float lower = i == 0 ? zs[0] : 0.5f * (zs[i] + zs[i - 1]);
      const float upper = i == S - 1 ? zs[S - 1] : 0.5f * (zs[i + 1] + zs[i]);
      zj[i] = lower + (upper - lower) * t_rand[(size_t)r * S + i];
    }
    if (rays_d_out) { rays_d_out[3 * r] = d[0]; rays_d_out[3 * r + 1] = d[1]; rays_d_out[3 * r + 2] = d[2]; }
    for (int i = 0; i < S; ++i) {
      const size_t q = (size_t)r * S + i;
      z_out[q] = zj[i];
      float p[3] = {o[0] + zj[i] * d[0], o[1] + zj[i] * d[1], o[2] + zj[i] * d[2]};   /* :716 */
      if (pts_out) { pts_out[3 * q] = p[0]; pts_out[3 * q + 1] = p[1]; pts_out[3 * q + 2] = p[2]; }
      if (ndc_out) {
        float nx, ny, pz;
        project(sc->w2c, sc->K, iw, ih, p, &nx, &ny, &pz);
        const size_t plane = (size_t)R * S * 3;
        for (int k = 0; k < 4; ++k) {                  /* stage1, stage2, stage3, ndc; :349-367 */
          const float nn = k < 3 ? nr[k] : sc->near, ff = k < 3 ? fr[k] : sc->far;
          float *dst = ndc_out + k * plane + 3 * q;
          dst[0] = nx; dst[1] = ny; dst[2] = (pz - nn) / (ff - nn);
        }
      }
    }
  }
}

/* ATen grid_sampler helpers (torch/include/ATen/native/GridSampler.h:27-36,58-60; the CUDA
 * twin is ATen/native/cuda/GridSampler.cuh).  padding_mode = border everywhere on this path. */
static inline float unnorm(float g, int size, int align_corners) {
  if (align_corners) return ((g + 1.0f) / 2.0f) * (float)(size - 1);
  return ((g + 1.0f) * (float)size - 1.0f) / 2.0f;
}
static inline float clipc(float x, int size) { return fminf((float)(size - 1), fmaxf(x, 0.0f)); }

/* trilinear gather of C=8 channels from an NCDHW volume; returns floor indices. */
static inline void tri8(const float *vol, int D, int h, int w, float gx, float gy, float gz, float *out,
                        int32_t *idx) {
  const float ix = clipc(unnorm(gx, w, 0), w), iy = clipc(unnorm(gy, h, 0), h), iz = clipc(unnorm(gz, D, 0), D);
  const float fx = floorf(ix), fy = floorf(iy), fz = floorf(iz);
  const int x0 = (int)fx, y0 = (int)fy, z0 = (int)fz;
  const float tx = ix - fx, ty = iy - fy, tz = iz - fz;      /* == ix - ix_tnw */
  const float ux = (fx + 1.0f) - ix, uy = (fy + 1.0f) - iy, uz = (fz + 1.0f) - iz;  /* ix_bse - ix */
  if (idx) { idx[0] = x0; idx[1] = y0; idx[2] = z0; }
  const size_t cs = (size_t)D * h * w;
  for (int c = 0; c < 8; ++c) {
    const float *v = vol + c * cs;
    float acc = 0.0f;
    for (int corner = 0; corner < 8; ++corner) {               /* tnw,tne,tsw,tse,bnw,bne,bsw,bse */
      const int dx = corner & 1, dy = (corner >> 1) & 1, dz = corner >> 2;
      const int x = x0 + dx, y = y0 + dy, zz = z0 + dz;
      if (x >= w || y >= h || zz >= D) continue;               /* within_bounds_3d */
      const float wgt = (dx ? tx : ux) * (dy ? ty : uy) * (dz ? tz : uz);
      acc += v[((size_t)zz * h + y) * w + x] * wgt;
    }
    out[c] = acc;
  }
}

/* bilinear gather of C channels from an [C,H,W] map. */
static inline void bil(const float *img, int C, int H, int W, float gx, float gy, int align, float *out,
                       int32_t *idx) {
  const float ix = clipc(unnorm(gx, W, align), W), iy = clipc(unnorm(gy, H, align), H);
  const float fx = floorf(ix), fy = floorf(iy);
  const int x0 = (int)fx, y0 = (int)fy;
  const float tx = ix - fx, ty = iy - fy, ux = (fx + 1.0f) - ix, uy = (fy + 1.0f) - iy;
  if (idx) { idx[0] = x0; idx[1] = y0; }
  for (int c = 0; c < C; ++c) {
    const float *v = img + (size_t)c * H * W;
    float acc = 0.0f;
    for (int corner = 0; corner < 4; ++corner) {               /* nw, ne, sw, se */
      const int dx = corner & 1, dy = corner >> 1;
      const int x = x0 + dx, y = y0 + dy;
      if (x >= W || y >= H) continue;
      acc += v[(size_t)y * W + x] * ((dx ? tx : ux) * (dy ? ty : uy));
    }
    out[c] = acc;
  }
}

/* ------------------------------------------------------------------------------------------
 * Feature row assembly: volume gathers + confidence + per-source-view projection/gathers.
 * reference: utils/utils.py:833-893 (index_point_feature), :742-799 (build_color_volume),
 *            network/renderer.py:177-212 (gen_pts_feats; layout of input_feat).
 * idx layout per sample (int32): [s1 x,y,z | s2 x,y,z | s3 x,y,z | conf x,y | (x,y,mask) per view]
 * ------------------------------------------------------------------------------------------ */
void uco_gather(const uco_scene *sc, const float *pts, const float *ndc /* [4][R][S][3] */, int R,
                float *feat /* [R][S][F] */, int32_t *idx /* [R][S][11+3V] or NULL */) {
  const int S = sc->S, V = sc->V, F = 24 + 12 * V + 1, NI = 11 + 3 * V;
  const float iw = (float)(sc->W - 1), ih = (float)(sc->H - 1);
  const size_t plane = (size_t)R * S * 3;
  const size_t HW = (size_t)sc->H * sc->W;
#pragma omp parallel for schedule(static)
  for (long q = 0; q < (long)R * S; ++q) {
    float *row = feat + (size_t)q * F;
    int32_t *ir = idx ? idx + (size_t)q * NI : NULL;
    for (int k = 0; k < 3; ++k) {
      const float *c = ndc + k * plane + 3 * q;
      tri8(sc->vol[k], sc->D[k], sc->h[k], sc->w[k], c[0] * 2 - 1.0f, c[1] * 2 - 1.0f, c[2] * 2 - 1.0f,
           row + 8 * k, ir ? ir + 3 * k : NULL);
    }
    {   /* confidence: 5-D sampler over a depth-1 "volume" == 2-D bilinear, align_corners=False (Q13) */
      const float *c = ndc + 2 * plane + 3 * q;
      bil(sc->conf, 1, sc->H, sc->W, c[0] * 2 - 1.0f, c[1] * 2 - 1.0f, 0, row + F - 1, ir ? ir + 9 : NULL);
    }
    const float *p = pts + 3 * q;
    for (int v = 0; v < V; ++v) {
      float nx, ny, pz;
      project(sc->src_w2c + 16 * v, sc->src_K + 9 * v, iw, ih, p, &nx, &ny, &pz);
      const float gx = nx * 2.0f - 1.0f, gy = ny * 2.0f - 1.0f;
      int32_t ii[2];
      bil(sc->imgs + (size_t)v * 3 * HW, 3, sc->H, sc->W, gx, gy, 1, row + 24 + 4 * v, ii);
      const float m = (gx > -1.0f && gx < 1.0f && gy > -1.0f && gy < 1.0f) ? 1.0f : 0.0f;   /* :791-792 */
      row[24 + 4 * v + 3] = m;
      bil(sc->feats + (size_t)v * 8 * HW, 8, sc->H, sc->W, gx, gy, 1, row + 24 + 4 * V + 8 * v, NULL);
      if (ir) { ir[11 + 3 * v] = ii[0]; ir[11 + 3 * v + 1] = ii[1]; ir[11 + 3 * v + 2] = (int32_t)m; }
    }
  }
}

/* view-direction feature. reference: network/renderer.py:232-239, :163-174.
 * Rays with global index < dir_switch use dir_rot0, the rest dir_rot1 (quirk Q4: rendering() strips
 * the target pose in place on its first call, so later chunks rotate into source view 1's frame). */
void uco_dir_feature(const uco_scene *sc, const float *rays_d, int R, int ray_offset, float *angle) {
  for (int r = 0; r < R; ++r) {
    const float *d = rays_d + 3 * r;
    const float nrm = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    const float u[3] = {d[0] / nrm, d[1] / nrm, d[2] / nrm};
    mm3t((r + ray_offset) < sc->dir_switch ? sc->dir_rot0 : sc->dir_rot1, 3, u, angle + 3 * r);
  }
}

/* positional encoding, order [x | sin(f0 x) .. sin(f_{L-1} x) | cos(f0 x) .. cos(f_{L-1} x)].
 * reference: network/models.py:37-40,50-54. */
void uco_embed(const float *x, int L, float *out) {
  out[0] = x[0]; out[1] = x[1]; out[2] = x[2];
  for (int l = 0; l < L; ++l) {
    const float f = (float)(1 << l);
    for (int c = 0; c < 3; ++c) {
      const float a = x[c] * f;
      out[3 + 3 * l + c] = sinf(a);
      out[3 + 3 * L + 3 * l + c] = cosf(a);
    }
  }
}

/* alpha compositing. reference: network/renderer.py:25-36 (raw2alpha), :109-140 (raw2outputs). */
void uco_composite(const float *raw, const float *z, int R, int S, float *rgb, float *depth) {
#pragma omp parallel for schedule(static)
  for (int r = 0; r < R; ++r) {
    float T = 1.0f, acc[3] = {0, 0, 0}, dm = 0.0f;
    for (int i = 0; i < S; ++i) {
      const float *s = raw + ((size_t)r * S + i) * 4;
      const float alpha = 1.0f - expf(-s[3]);          /* dists ignored (Q3) */
      const float wgt = alpha * T;
      acc[0] += wgt * s[0]; acc[1] += wgt * s[1]; acc[2] += wgt * s[2];
      dm += wgt * z[(size_t)r * S + i];
      T = T * (1.0f - alpha + 1e-10f);
    }
    rgb[3 * r] = acc[0]; rgb[3 * r + 1] = acc[1]; rgb[3 * r + 2] = acc[2];
    depth[r] = dm;
  }
}

/* ------------------------------------------------------------------------------------------
 * Backward restatements (training step, train.py:147-190 through autograd).  Serial loops, double accumulators:
 * these are the checkers of ucn_composite_bwd / ucn_gather_bwd and are pinned to the reference's autograd
 * gradients (tests/golden/grad_*.npz).
 * ------------------------------------------------------------------------------------------ */
/* backward of uco_composite: cumprod backward as ATen computes it for non-zero factors (reverse cumsum of
 * grad*out divided by the input), network/renderer.py:25-36,125-131. */
void uco_composite_bwd(const float *raw, const float *z, int R, int S, const float *g_rgb, const float *g_depth,
                       float *d_raw) {
  for (int r = 0; r < R; ++r) {
    double T = 1.0, suffix = 0.0;
    double Ti[UCO_MAX_S], wv[UCO_MAX_S];
    for (int i = 0; i < S; ++i) {
      const float *s = raw + ((size_t)r * S + i) * 4;
      const double alpha = 1.0 - exp(-(double)s[3]);
      const double v = (double)g_rgb[3 * r] * s[0] + (double)g_rgb[3 * r + 1] * s[1] + (double)g_rgb[3 * r + 2] * s[2] +
                       (double)g_depth[r] * z[(size_t)r * S + i];
      Ti[i] = T;
      wv[i] = alpha * T * v;
      T *= (1.0 - alpha) + 1e-10;
    }
    for (int i = S - 1; i >= 0; --i) {
      const float *s = raw + ((size_t)r * S + i) * 4;
      float *d = d_raw + ((size_t)r * S + i) * 4;
      const double e = exp(-(double)s[3]), alpha = 1.0 - e, q = (1.0 - alpha) + 1e-10, w = alpha * Ti[i];
      const double vv = (double)g_rgb[3 * r] * s[0] + (double)g_rgb[3 * r + 1] * s[1] + (double)g_rgb[3 * r + 2] * s[2] +
                        (double)g_depth[r] * z[(size_t)r * S + i];
      d[0] = (float)(w * g_rgb[3 * r]); d[1] = (float)(w * g_rgb[3 * r + 1]); d[2] = (float)(w * g_rgb[3 * r + 2]);
      d[3] = (float)(e * (Ti[i] * vv - suffix / q));
      suffix += wv[i];
    }
  }
}

static inline void tri8_bwd(float *dvol, int D, int h, int w, float gx, float gy, float gz, const float *g) {
  const float ix = clipc(unnorm(gx, w, 0), w), iy = clipc(unnorm(gy, h, 0), h), iz = clipc(unnorm(gz, D, 0), D);
  const float fx = floorf(ix), fy = floorf(iy), fz = floorf(iz);
  const int x0 = (int)fx, y0 = (int)fy, z0 = (int)fz;
  const float tx = ix - fx, ty = iy - fy, tz = iz - fz;
  const float ux = (fx + 1.0f) - ix, uy = (fy + 1.0f) - iy, uz = (fz + 1.0f) - iz;
  const size_t cs = (size_t)D * h * w;
  for (int corner = 0; corner < 8; ++corner) {
    const int dx = corner & 1, dy = (corner >> 1) & 1, dz = corner >> 2;
    const int x = x0 + dx, y = y0 + dy, zz = z0 + dz;
    if (x >= w || y >= h || zz >= D) continue;
    const float wgt = (dx ? tx : ux) * (dy ? ty : uy) * (dz ? tz : uz);
    for (int c = 0; c < 8; ++c) dvol[c * cs + ((size_t)zz * h + y) * w + x] += wgt * g[c];
  }
}
static inline void bil_bwd(float *dimg, int C, int H, int W, float gx, float gy, int align, const float *g) {
  const float ix = clipc(unnorm(gx, W, align), W), iy = clipc(unnorm(gy, H, align), H);
  const float fx = floorf(ix), fy = floorf(iy);
  const int x0 = (int)fx, y0 = (int)fy;
  const float tx = ix - fx, ty = iy - fy, ux = (fx + 1.0f) - ix, uy = (fy + 1.0f) - iy;
  for (int corner = 0; corner < 4; ++corner) {
    const int dx = corner & 1, dy = corner >> 1;
    const int x = x0 + dx, y = y0 + dy;
    if (x >= W || y >= H) continue;
    const float wgt = (dx ? tx : ux) * (dy ? ty : uy);
    for (int c = 0; c < C; ++c) dimg[(size_t)c * H * W + (size_t)y * W + x] += wgt * g[c];
  }
}
/* backward of uco_gather (grid_sample backward w.r.t. the sampled tensors, utils/utils.py:776-788,859-890):
 * d_feat [R][S][F] -> d_vol[k] [8,D_k,h_k,w_k], d_feats [V,8,H,W], d_conf [H,W] (accumulated; caller zeroes). */
void uco_gather_bwd(const uco_scene *sc, const float *pts, const float *ndc, int R, const float *d_feat,
                    float *d_vol0, float *d_vol1, float *d_vol2, float *d_feats, float *d_conf) {
  const int S = sc->S, V = sc->V, F = 24 + 12 * V + 1;
  const float iw = (float)(sc->W - 1), ih = (float)(sc->H - 1);
  const size_t plane = (size_t)R * S * 3, HW = (size_t)sc->H * sc->W;
  float *dv[3] = {d_vol0, d_vol1, d_vol2};
  for (long q = 0; q < (long)R * S; ++q) {
    const float *g = d_feat + (size_t)q * F;
    for (int k = 0; k < 3; ++k) {
      const float *c = ndc + k * plane + 3 * q;
      tri8_bwd(dv[k], sc->D[k], sc->h[k], sc->w[k], c[0] * 2 - 1.0f, c[1] * 2 - 1.0f, c[2] * 2 - 1.0f, g + 8 * k);
    }
    {
      const float *c = ndc + 2 * plane + 3 * q;
      bil_bwd(d_conf, 1, sc->H, sc->W, c[0] * 2 - 1.0f, c[1] * 2 - 1.0f, 0, g + F - 1);
    }
    const float *p = pts + 3 * q;
    for (int v = 0; v < V; ++v) {
      float nx, ny, pz;
      project(sc->src_w2c + 16 * v, sc->src_K + 9 * v, iw, ih, p, &nx, &ny, &pz);
      bil_bwd(d_feats + (size_t)v * 8 * HW, 8, sc->H, sc->W, nx * 2.0f - 1.0f, ny * 2.0f - 1.0f, 1, g + 24 + 4 * V + 8 * v);
    }
  }
}

/* data/ray_utils.py:152-197 ray_marcher.  rays [R][8] = (o, d, near, far); t = linspace(0,1,S); rnd [R][S] uniforms
 * (used when perturb > 0).  xyz [R][S][3], z [R][S].  Every elementwise op of the reference rounds on its own. */
static inline float marcher_z(float near, float far, float t, int lindisp) {
  const float omt = 1.0f - t;
  if (!lindisp) { const float a = near * omt, b = far * t; return a + b; }
  const float in = 1.0f / near, jf = 1.0f / far;
  const float a = in * omt, b = jf * t;
  const float s = a + b;
  return 1.0f / s;
}
void uco_ray_marcher(const float *rays, int R, int S, const float *t, const float *rnd, float perturb, int lindisp,
                     float *xyz, float *z) {
  for (int r = 0; r < R; ++r) {
    const float *ray = rays + (size_t)r * 8;
    for (int j = 0; j < S; ++j) {
      float zj = marcher_z(ray[6], ray[7], t[j], lindisp);
      if (perturb > 0.0f) {
        float lower = zj, upper = zj;
        if (j > 0) { const float sm = marcher_z(ray[6], ray[7], t[j - 1], lindisp) + zj; lower = 0.5f * sm; }
        if (j < S - 1) { const float sm = zj + marcher_z(ray[6], ray[7], t[j + 1], lindisp); upper = 0.5f * sm; }
        const float w = upper - lower, pr = perturb * rnd[(size_t)r * S + j];
        const float wp = w * pr;
        zj = lower + wp;
      }
      z[(size_t)r * S + j] = zj;
      for (int c = 0; c < 3; ++c) { const float dz = ray[3 + c] * zj; xyz[((size_t)r * S + j) * 3 + c] = ray[c] + dz; }
    }
  }
}
