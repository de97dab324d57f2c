/*
 * ucnerf_oracle_mlp.c -- CPU restatement of PE + the uncertainty-conditioned MLP, and the chained
 * whole-path driver used as CPU baseline.  TEST INFRASTRUCTURE ONLY (see ucnerf_oracle.c).
 *
 * reference: network/renderer.py:78-106 (run_network_mvs: [PE(ndc) | input_feat | PE(angle)]),
 *            network/models.py:138-184 (BaseAdapt_Renderer.forward), :187-207 (UCNeRF).
 * fp32 throughout, like the reference (which runs SIMT sgemm, TF32 off).  This file is tolerance-
 * checked (1e-4), so it may be compiled with FMA contraction / AVX2.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "ucnerf_oracle.h"

#define BLK 48

/* torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU baseline legs ask for all host cores explicitly. */
void uco_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int uco_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* Y[B][N] = X[B][K] @ Wt[K][N] + b ; 4-row register blocking, N innermost (vectorises). */
static void linear(const float *X, int ldx, int K, const float *Wt, const float *b, float *Y, int ldy, int N,
                   int B) {
  int i = 0;
  for (; i + 4 <= B; i += 4) {
    float *y0 = Y + (size_t)i * ldy, *y1 = y0 + ldy, *y2 = y1 + ldy, *y3 = y2 + ldy;
    for (int j = 0; j < N; ++j) { y0[j] = b[j]; y1[j] = b[j]; y2[j] = b[j]; y3[j] = b[j]; }
    const float *x0 = X + (size_t)i * ldx, *x1 = x0 + ldx, *x2 = x1 + ldx, *x3 = x2 + ldx;
    for (int k = 0; k < K; ++k) {
      const float a0 = x0[k], a1 = x1[k], a2 = x2[k], a3 = x3[k];
      const float *w = Wt + (size_t)k * N;
      for (int j = 0; j < N; ++j) {
        const float wj = w[j];
        y0[j] += a0 * wj; y1[j] += a1 * wj; y2[j] += a2 * wj; y3[j] += a3 * wj;
      }
    }
  }
  for (; i < B; ++i) {
    float *y = Y + (size_t)i * ldy;
    const float *x = X + (size_t)i * ldx;
    for (int j = 0; j < N; ++j) y[j] = b[j];
    for (int k = 0; k < K; ++k) {
      const float a = x[k];
      const float *w = Wt + (size_t)k * N;
      for (int j = 0; j < N; ++j) y[j] += a * w[j];
    }
  }
}

static float *transpose(const float *W, int out, int in) {   /* torch [out,in] -> [in,out] */
  float *t = (float *)malloc(sizeof(float) * (size_t)out * in);
  for (int o = 0; o < out; ++o)
    for (int i = 0; i < in; ++i) t[(size_t)i * out + o] = W[(size_t)o * in + i];
  return t;
}

typedef struct {
  float *pts[8], *film, *fb, *feature, *views, *vconf, *rgb, *alpha, *crgb, *alpha1;
} wt_t;

static void wt_make(const uco_weights *w, wt_t *t) {
  const int Wd = w->width, P = w->in_pts, V = w->V;
  for (int i = 0; i < w->depth; ++i) {
    const int k = i == 0 ? P : (i - 1 == w->skip ? Wd + P : Wd);
    t->pts[i] = transpose(w->pts_w[i], Wd, k);
  }
  t->film = transpose(w->film_w, Wd, 24 + 4 * V);
  t->fb = transpose(w->fb_w, Wd, 8 * V);
  t->feature = transpose(w->feature_w, Wd, Wd);
  t->views = transpose(w->views_w, Wd / 2, Wd + w->in_views);
  t->vconf = transpose(w->vconf_w, Wd / 2, Wd + w->in_views);
  t->rgb = transpose(w->rgb_w, 3, Wd / 2);
  t->alpha = transpose(w->alpha_w, 1, Wd / 2);
  t->crgb = transpose(w->crgb_w, 3, Wd);
  t->alpha1 = transpose(w->alpha1_w, 1, Wd);
}
static void wt_free(const uco_weights *w, wt_t *t) {
  for (int i = 0; i < w->depth; ++i) free(t->pts[i]);
  free(t->film); free(t->fb); free(t->feature); free(t->views); free(t->vconf);
  free(t->rgb); free(t->alpha); free(t->crgb); free(t->alpha1);
}

/* one block of B <= BLK consecutive samples q0.. ; angle is per ray (q / S). */
static void mlp_block(const uco_weights *w, const wt_t *t, const float *ndc_pts, const float *feat,
                      const float *angle, long q0, int B, int S, float *raw) {
  const int Wd = w->width, P = w->in_pts, PV = w->in_views, V = w->V, F = 24 + 12 * V + 1;
  const int LP = (P - 3) / 6, LV = (PV - 3) / 6;
  const int ldh = P + Wd;                                   /* [pe | h] so the skip concat is free */
  float pe_h[BLK * (63 + 128)], mod[BLK * 128], fbv[BLK * 128], tmp[BLK * 128], h1[BLK * (128 + 27)];
  float br[BLK * 64], bo[BLK * 4], u[BLK], base[BLK * 4];
  for (int i = 0; i < B; ++i) {
    uco_embed(ndc_pts + 3 * (q0 + i), LP, pe_h + (size_t)i * ldh);
    u[i] = 1.0f - feat[(size_t)(q0 + i) * F + F - 1];      /* models.py:149 */
  }
  linear(feat + (size_t)q0 * F, F, 24 + 4 * V, t->film, w->film_b, mod, Wd, Wd, B);          /* :150 */
  linear(feat + (size_t)q0 * F + 24 + 4 * V, F, 8 * V, t->fb, w->fb_b, fbv, Wd, Wd, B);      /* :151 */
  for (int l = 0; l < w->depth; ++l) {                                                        /* :153-157 */
    const float *X = l == 0 ? pe_h : (l - 1 == w->skip ? pe_h : pe_h + P);
    const int K = l == 0 ? P : (l - 1 == w->skip ? P + Wd : Wd);
    linear(X, ldh, K, t->pts[l], w->pts_b[l], tmp, Wd, Wd, B);
    for (int i = 0; i < B; ++i)
      for (int j = 0; j < Wd; ++j) {
        const float v = tmp[i * Wd + j] * mod[i * Wd + j];
        pe_h[(size_t)i * ldh + P + j] = v > 0.0f ? v : 0.0f;
      }
  }
  const float *h = pe_h + P;
  linear(h, ldh, Wd, t->crgb, w->crgb_b, base, 4, 3, B);                                      /* :161 */
  for (int i = 0; i < B; ++i) {                                                               /* :162 */
    float a = w->alpha1_b[0];
    for (int k = 0; k < Wd; ++k) a += h[(size_t)i * ldh + k] * t->alpha1[k];
    base[i * 4 + 3] = a;
  }
  for (int i = 0; i < B; ++i)
    for (int j = 0; j < Wd; ++j) tmp[i * Wd + j] = h[(size_t)i * ldh + j] * fbv[i * Wd + j];
  const int ld1 = Wd + PV;
  linear(tmp, Wd, Wd, t->feature, w->feature_b, h1, ld1, Wd, B);                              /* :164 */
  for (int i = 0; i < B; ++i) uco_embed(angle + 3 * ((q0 + i) / S), LV, h1 + (size_t)i * ld1 + Wd);
  linear(h1, ld1, ld1, t->views, w->views_b, br, Wd / 2, Wd / 2, B);                          /* :168-170 */
  for (int i = 0; i < B * (Wd / 2); ++i) br[i] = br[i] > 0.0f ? br[i] : 0.0f;
  linear(br, Wd / 2, Wd / 2, t->rgb, w->rgb_b, bo, 4, 3, B);
  linear(h1, ld1, ld1, t->vconf, w->vconf_b, br, Wd / 2, Wd / 2, B);                          /* :172-174 */
  for (int i = 0; i < B * (Wd / 2); ++i) br[i] = br[i] > 0.0f ? br[i] : 0.0f;
  for (int i = 0; i < B; ++i) {
    float a = w->alpha_b[0];
    for (int k = 0; k < Wd / 2; ++k) a += br[i * (Wd / 2) + k] * t->alpha[k];
    bo[i * 4 + 3] = a;
  }
  for (int i = 0; i < B; ++i) {                                                               /* :177-180 */
    const float un = u[i], cf = 1.0f - un;
    float *o = raw + (size_t)(q0 + i) * 4;
    for (int c = 0; c < 3; ++c) {
      const float x = base[i * 4 + c] * cf + bo[i * 4 + c] * un;
      o[c] = 1.0f / (1.0f + expf(-x));
    }
    const float s = bo[i * 4 + 3] * cf + base[i * 4 + 3] * un;
    o[3] = s > 0.0f ? s : 0.0f;
  }
}

void uco_mlp(const uco_weights *w, const float *ndc_pts, const float *feat, const float *angle, int R, int S,
             float *raw) {
  wt_t t;
  wt_make(w, &t);
  const long N = (long)R * S;
#pragma omp parallel for schedule(dynamic, 4)
  for (long q0 = 0; q0 < N; q0 += BLK) {
    const int B = (int)(N - q0 < BLK ? N - q0 : BLK);
    mlp_block(w, &t, ndc_pts, feat, angle, q0, B, S, raw);
  }
  wt_free(w, &t);
}

/* Whole path over rays [ray0, ray0+R) of the row-major pixel grid, processed in chunks so the staged
 * intermediates stay cache/RAM friendly.  Mirrors the reference's eval loop (train.py:251-275). */
void uco_render_rays(const uco_scene *sc, const uco_weights *wt, int ray0, int R, const float *t_vals,
                     const float *t_rand, float *rgb, float *depth) {
  const int S = sc->S, F = 24 + 12 * sc->V + 1, CH = 1024;
  float *py = (float *)malloc(sizeof(float) * CH), *px = (float *)malloc(sizeof(float) * CH);
  float *z = (float *)malloc(sizeof(float) * (size_t)CH * S);
  float *rd = (float *)malloc(sizeof(float) * (size_t)CH * 3), *ang = (float *)malloc(sizeof(float) * (size_t)CH * 3);
  float *pts = (float *)malloc(sizeof(float) * (size_t)CH * S * 3);
  float *ndc = (float *)malloc(sizeof(float) * (size_t)CH * S * 12);
  float *feat = (float *)malloc(sizeof(float) * (size_t)CH * S * F);
  float *raw = (float *)malloc(sizeof(float) * (size_t)CH * S * 4);
  for (int c0 = 0; c0 < R; c0 += CH) {
    const int n = R - c0 < CH ? R - c0 : CH;
    for (int i = 0; i < n; ++i) {
      const int g = ray0 + c0 + i;
      py[i] = (float)(g / sc->W); px[i] = (float)(g % sc->W);
    }
    uco_raygen_sample(sc, py, px, n, t_vals, t_rand + (size_t)c0 * S, z, rd, pts, ndc);
    uco_gather(sc, pts, ndc, n, feat, NULL);
    uco_dir_feature(sc, rd, n, ray0 + c0, ang);
    uco_mlp(wt, ndc + (size_t)3 * n * S * 3, feat, ang, n, S, raw);
    uco_composite(raw, z, n, S, rgb + (size_t)3 * c0, depth + c0);
  }
  free(py); free(px); free(z); free(rd); free(ang); free(pts); free(ndc); free(feat); free(raw);
}
