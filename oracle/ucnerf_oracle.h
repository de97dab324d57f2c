/* ucnerf_oracle.h -- C interface of the CPU oracle (TEST INFRASTRUCTURE ONLY, see ucnerf_oracle.c).
 * Arrays use the REFERENCE's layouts (NCDHW volumes, [V,3,H,W] images, torch [out,in] weights). */
#ifndef UCNERF_ORACLE_H
#define UCNERF_ORACLE_H
#include <stdint.h>

#define UCO_MAX_S 768
#define UCO_MAX_V 8

typedef struct {
  int32_t H, W, V, S;
  int32_t D[3], h[3], w[3];
  float K[9], c2w[16], w2c[16];
  float near, far;
  float dir_rot0[9], dir_rot1[9];
  int32_t dir_switch;
  int32_t _pad;
  const float *depth[3];   /* [D_k,h_k,w_k]   outputs['stageK']['depth_values'][0]          */
  const float *vol[3];     /* [8,D_k,h_k,w_k] outputs['stageK']['volume_feature_no_ref'][0] */
  const float *conf;       /* [H,W]                                                         */
  const float *imgs;       /* [V,3,H,W]                                                     */
  const float *feats;      /* [V,8,H,W]                                                     */
  const float *src_w2c;    /* [V,16]                                                        */
  const float *src_K;      /* [V,9]                                                         */
} uco_scene;

typedef struct {
  int32_t V, in_pts, in_views, width, depth, skip;   /* 63, 27, 128, 6, 4 */
  const float *pts_w[8], *pts_b[8];
  const float *film_w, *film_b;        /* nerf.pts_bias_depth_fine   [W, 24+4V] */
  const float *fb_w, *fb_b;            /* nerf.pts_bias_confidence   [W, 8V]    */
  const float *feature_w, *feature_b;  /* nerf.feature_linear        [W, W]     */
  const float *views_w, *views_b;      /* nerf.views_linears.0       [W/2, W+27]*/
  const float *vconf_w, *vconf_b;      /* nerf.view_confi_linears.0  [W/2, W+27]*/
  const float *rgb_w, *rgb_b;          /* nerf.rgb_linear            [3, W/2]   */
  const float *alpha_w, *alpha_b;      /* nerf.alpha_linear          [1, W/2]   */
  const float *crgb_w, *crgb_b;        /* nerf.confi_rgb_linear      [3, W]     */
  const float *alpha1_w, *alpha1_b;    /* nerf.alpha_linear_1        [1, W]     */
} uco_weights;

void uco_raygen_sample(const uco_scene *sc, const float *pix_y, const float *pix_x, int R,
                       const float *t_vals, const float *t_rand, float *z_out, float *rays_d_out,
                       float *pts_out, float *ndc_out);
void uco_gather(const uco_scene *sc, const float *pts, const float *ndc, int R, float *feat, int32_t *idx);
void uco_dir_feature(const uco_scene *sc, const float *rays_d, int R, int ray_offset, float *angle);
void uco_embed(const float *x, int L, float *out);
void uco_composite(const float *raw, const float *z, int R, int S, float *rgb, float *depth);
/* network/models.py:138-184 + network/renderer.py:78-106 */
void uco_mlp(const uco_weights *wt, const float *ndc_pts, const float *feat, const float *angle, int R,
             int S, float *raw);
/* whole path for n rays starting at global ray index ray0 of the row-major pixel grid */
void uco_render_rays(const uco_scene *sc, const uco_weights *wt, int ray0, int R, const float *t_vals,
                     const float *t_rand, float *rgb, float *depth);
void uco_ray_marcher(const float *rays, int R, int S, const float *t, const float *rnd, float perturb, int lindisp,
                     float *xyz, float *z);
/* backward restatements (training step): checkers of ucn_composite_bwd / ucn_gather_bwd */
void uco_composite_bwd(const float *raw, const float *z, int R, int S, const float *g_rgb, const float *g_depth,
                       float *d_raw);
void uco_gather_bwd(const uco_scene *sc, const float *pts, const float *ndc, int R, const float *d_feat,
                    float *d_vol0, float *d_vol1, float *d_vol2, float *d_feats, float *d_conf);
int uco_num_threads(void);
void uco_set_num_threads(int n);
#endif
