/*
 * ucnerf.h -- C ABI of libucnerf.so: the B200-native (sm_100a) UC-NeRF per-ray rendering hot path.
 *
 * This is the drop-in boundary.  The reference (wrld/UC-NeRF) has no FFI/plugin layer: its hot path is
 * plain Python functions over torch tensors, so each entry point below names the reference function
 * (file:line, relative to the reference repo) whose work it replaces; the Python mirror of those
 * functions (uc-nerf_b200/*.py: same names, arguments, return tuples) binds them through ctypes.
 * INTEGRATION.md shows the binding a reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, else a negative UCN_E_* code or a positive cudaError_t;
 *     ucn_last_error() returns a human-readable string for the calling thread.
 *   - all `const float*` / `float*` arguments are DEVICE pointers unless the name ends in `_host`.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); no call synchronises
 *     except the *_host entry points, which copy H2D/D2H and synchronise `stream` before returning.
 *   - kernels never allocate; callers own every buffer.  There is NO CPU fallback.
 */
#ifndef UCNERF_H
#define UCNERF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UCN_MAX_V 8        /* source views                         */
#define UCN_MAX_S 1024     /* samples per ray (multiple of 3)      */
#define UCN_FEAT_C 8       /* channels of every feature map/volume */

enum {
  UCN_OK = 0,
  UCN_E_ARG = -1,          /* bad argument (null pointer, S%3!=0, V>UCN_MAX_V, ...) */
  UCN_E_UNSUPPORTED = -2,  /* device is not sm_100 or mode not compiled in          */
  UCN_E_WORKSPACE = -3     /* workspace too small (see ucn_frame_workspace_bytes)   */
};

/* MLP arithmetic.
 *   FP32      SIMT fp32 (parity mode, 1e-4 gate).
 *   BF16X3_TC tcgen05, split-bf16 operands (hi + lo, three MMAs per K step), fp32 accumulate in TMEM: the
 *             production tensor-core path; meets the 1e-3 per-pixel RGB / 0.05 dB gate.
 *   BF16_TC   tcgen05, single-pass bf16 operands: fastest, but plain bf16 rounding of a 6-layer FiLM MLP
 *             gives ~2e-2 worst-case RGB error on random-init weights, so it is opt-in only. */
enum { UCN_MLP_FP32 = 0, UCN_MLP_BF16_TC = 1, UCN_MLP_BF16X3_TC = 2 };

/* One target view ("frame"): cameras by value, packed maps by device pointer.
 * Packed layouts (produced by ucn_pack_*):
 *   nearfar[k] : [h_k][w_k][2]        first / last depth-hypothesis plane of stage k
 *   vol[k]     : [h_k][w_k][D_k][8]   cost-feature volume of stage k, depth innermost then channels (ucn_pack_volume)
 *   conf       : [H][W]               photometric confidence (as is)
 *   rgb        : [V][H][W][4]         source images, rgb + zero pad
 *   feat       : [V][H][W][8]         source FPN features, channels-last                        */
typedef struct ucn_frame {
  int32_t H, W, V, S;
  int32_t D[3], h[3], w[3];
  float K[9];               /* target intrinsics           (utils/utils.py:639-645)          */
  float c2w[16];            /* target camera-to-world      (tgt_to_world)                    */
  float w2c[16];            /* world-to-target             (world_to_ref)                    */
  float near, far;          /* near_fars_ref[ref_idx]      (utils/utils.py:657,713-714)      */
  float dir_rot0[9];        /* rotation of the view-dir feature for rays <  dir_switch       */
  float dir_rot1[9];        /* ... and for rays >= dir_switch (reference quirk: renderer.py:241-243) */
  int32_t dir_switch;
  int32_t reserved;
  float src_w2c[UCN_MAX_V * 16];
  float src_K[UCN_MAX_V * 9];
  const float *nearfar[3];
  const float *vol[3];
  const float *conf;
  const float *rgb;
  const float *feat;
} ucn_frame;

/* Packed UC-NeRF MLP weights (produced by ucn_pack_weights from the torch state_dict arrays). */
typedef struct ucn_weights {
  int32_t V;                /* source views the film/fb layers were built for */
  int32_t reserved;
  const float *f32;         /* fp32 pack: transposed [K][N] matrices + biases  */
  const void *bf16;         /* bf16 pack: UMMA K-major core-matrix tiles (may be NULL if unused) */
  const void *bf16x3;       /* split-bf16 (hi/lo) pack for UCN_MLP_BF16X3_TC (may be NULL if unused) */
} ucn_weights;

/* Pointers to the reference state_dict tensors (torch layout [out,in], fp32, DEVICE memory).
 * network/models.py:87-104; the three layers forward() never touches are not needed. */
typedef struct ucn_state_dict {
  int32_t V, reserved;
  const float *pts_w[6], *pts_b[6];      /* nerf.pts_linears.{0..5}                    */
  const float *film_w, *film_b;          /* nerf.pts_bias_depth_fine    [128, 24+4V]   */
  const float *fb_w, *fb_b;              /* nerf.pts_bias_confidence    [128, 8V]      */
  const float *feature_w, *feature_b;    /* nerf.feature_linear         [128,128]      */
  const float *views_w, *views_b;        /* nerf.views_linears.0        [64,155]       */
  const float *vconf_w, *vconf_b;        /* nerf.view_confi_linears.0   [64,155]       */
  const float *rgb_w, *rgb_b;            /* nerf.rgb_linear             [3,64]         */
  const float *alpha_w, *alpha_b;        /* nerf.alpha_linear           [1,64]         */
  const float *crgb_w, *crgb_b;          /* nerf.confi_rgb_linear       [3,128]        */
  const float *alpha1_w, *alpha1_b;      /* nerf.alpha_linear_1         [1,128]        */
} ucn_state_dict;

/* ---- library ------------------------------------------------------------------------------- */
int ucn_version(void);
const char *ucn_last_error(void);
/* number of kernel launches issued by this library since process start / last reset */
int64_t ucn_launch_count(int reset);
/* 0 if the current device is sm_100 (B200), else UCN_E_UNSUPPORTED */
int ucn_check_device(void);

/* ---- K0: packing (once per target view / per weight load) ----------------------------------- */
/* src [B][C][N] -> dst [B][N][Cpad] (zero padded).  Replaces nothing in the reference: it is the
 * relayout that lets the gathers below issue one 16/32-byte vector load per texel instead of C
 * channel-strided 4-byte loads (NCDHW: utils/utils.py:869-890; NCHW: :776-788). */
int ucn_pack_channels_last(const float *src, float *dst, int B, int Cn, int64_t N, int Cpad, void *stream);
/* feature volume NCDHW [8][D][h][w] -> [h][w][D][8] (outputs['stageK']['volume_feature_no_ref'],
 * network/mvs_models.py:649,741-746): all samples of a ray stay in the same <=4 (x,y) columns of the target-frustum
 * volume (utils/utils.py:852-857), so depth-innermost turns the trilinear gather of a warp into contiguous runs. */
int ucn_pack_volume(const float *src, float *dst, int D, int h, int w, void *stream);
/* depth_values [D][h][w] -> [h][w][2] = (plane 0, plane D-1)   (utils/utils.py:668-683) */
int ucn_pack_nearfar(const float *depth_values, float *dst, int D, int h, int w, void *stream);
/* Row-band variants for ray-tile sharding: the volumes live in the TARGET frustum
 * (train.py:235 world_to_ref = pose_ref['w2cs'][0] is the target's own camera), so the rays of image rows [r0, r1)
 * touch only stage rows [r0/ds - 1, r1/ds + 1].  These pack source rows [row0, row0 + rows) of planes that hold h rows
 * each into a dense band; the caller stores `band - row0 * row_bytes` in ucn_frame.vol / .nearfar so that the
 * kernels keep indexing absolute rows (ucn_frame.h stays the FULL height: border clamping is unchanged). */
int ucn_pack_volume_rows(const float *src, float *dst, int D, int h, int w, int row0, int rows, void *stream);
int ucn_pack_nearfar_rows(const float *depth_values, float *dst, int D, int h, int w, int row0, int rows, void *stream);
/* host -> device copy of rows [row0, row0+rows) of `planes` fp32 [h][w] planes into a dense [planes][rows][w] band
 * (one strided DMA; src_host should be pinned).  The per-rank share of the per-pose map upload of a sharded sweep. */
int ucn_h2d_rows(float *dst, const float *src_host, int64_t planes, int h, int w, int row0, int rows, void *stream);
size_t ucn_weights_f32_bytes(int V);
size_t ucn_weights_bf16_bytes(int V);
size_t ucn_weights_bf16x3_bytes(int V);
/* any destination may be NULL (that pack is skipped) */
int ucn_pack_weights(const ucn_state_dict *sd, float *dst_f32, void *dst_bf16, void *dst_bf16x3, void *stream);

/* ---- K1: ray generation + depth-bounded sampling (+ target-view NDC) -------------------------
 * replaces get_rays_mvs (utils/utils.py:217-271, eval branch and explicit-pixel branch),
 * sample_points_uniform (:393-397), build_rays_test (:600-739) and get_ndc_coordinate (:323-373).
 * Rays are pixels ray0..ray0+R-1 of the row-major grid, or (pix_yx != NULL) explicit float
 * (row, col) pairs laid out [2][R].  t_vals = linspace(0,1,S/3) (device).  t_rand [R][S] uniform
 * jitter; if NULL an in-kernel Philox4x32-10 stream keyed by (seed, global ray, sample) is used.
 * Outputs (any may be NULL except z): z [R][S], rays_d [R][3], pts [R][S][3], ndc [4][R][S][3]
 * (stage1, stage2, stage3, ndc). */
int ucn_raygen_sample(const ucn_frame *f, const float *pix_yx, int ray0, int R, const float *t_vals,
                      const float *t_rand, uint64_t seed, float *z, float *rays_d, float *pts, float *ndc,
                      void *stream);

/* data/ray_utils.py:152-197 ray_marcher (not on the live eval path; kept as a thin mode of K1):
 * rays [R][8] = (origin, direction, near, far); t_vals = linspace(0,1,S); t_rand [R][S] uniforms (needed when
 * perturb > 0).  Outputs xyz [R][S][3], z [R][S]. */
int ucn_ray_marcher(const float *rays, int R, int S, const float *t_vals, const float *t_rand, float perturb,
                    int lindisp, float *xyz, float *z, void *stream);

/* ---- K2: projection + fused gathers -> feature rows ----------------------------------------
 * replaces index_point_feature (utils/utils.py:833-893), build_color_volume (:742-799),
 * get_ndc_coordinate(sample_2d=True) (:323-347) and gen_pts_feats (network/renderer.py:177-212).
 * feat [R*S][F], F = 24+12V+1, layout of input_feat (renderer.py:198-207).
 * idx (optional) int32 [R*S][11+3V]: floor indices of every gather + in-bounds mask, for parity. */
#define UCN_PART_VOLUMES 1   /* the three volume gathers + confidence (index_point_feature) */
#define UCN_PART_VIEWS 2     /* per-source-view rgb/mask/feature gathers (build_color_volume) */
#define UCN_PART_ALL 3
/* `parts` selects which gathers run (others leave zeros), so the two reference functions can also be
 * called on their own. */
int ucn_gather(const ucn_frame *f, const float *pts, const float *ndc_s1, const float *ndc_s2,
               const float *ndc_s3, int64_t n_samples, int parts, float *feat, int32_t *idx, void *stream);
/* world points [n][3] -> [n][3] = (x/(W-1), y/(H-1), projective depth) for one camera:
 * get_ndc_coordinate(..., sample_2d=True) (utils/utils.py:333-347,368-369). */
int ucn_project_points(const float *w2c_host16, const float *K_host9, int H, int W, const float *pts, int64_t n,
                       float *out, void *stream);
/* same, positions recomputed from z (bit-identical to K1's): also emits ndc 'ndc' [R*S][3] */
int ucn_gather_from_z(const ucn_frame *f, const float *z, const float *pix_yx, int ray0, int R, float *feat,
                      float *ndc_pts, void *stream);

/* same gathers, feature rows written in the tile-transposed interchange layout of the tensor-core MLP:
 * feat_tiles [ceil(R*S/128)][F][128] (tile t, channel c, row r = sample 128 t + r; rows past R*S are zero).
 * Both K2's stores and K3's per-row loads are then one coalesced 128-byte line per warp and channel. */
int ucn_gather_tiles_from_z(const ucn_frame *f, const float *z, const float *pix_yx, int ray0, int R,
                            float *feat_tiles, float *ndc_pts, void *stream);

/* view-direction feature: gen_dir_feature + prologue of rendering (network/renderer.py:163-174,232-239) */
int ucn_dir_feature(const ucn_frame *f, const float *rays_d, int R, int ray_offset, float *angle, void *stream);

/* ---- K3: positional encoding + uncertainty-conditioned MLP ----------------------------------
 * replaces Embedder.embed (network/models.py:50-54), run_network_mvs/batchify
 * (network/renderer.py:57-106) and BaseAdapt_Renderer.forward (network/models.py:138-184).
 * ndc_pts [N][3], feat [N][F], angle [R][3] (per ray; N = R*S) -> raw [N][4] = (rgb, sigma). */
int ucn_mlp_forward(const ucn_weights *w, const float *ndc_pts, const float *feat, const float *angle, int R,
                    int S, float *raw, int mode, void *stream);

/* same network on tile-transposed feature rows (ucn_gather_tiles_from_z); UCN_MLP_BF16X3_TC only */
int ucn_mlp_forward_tiles(const ucn_weights *w, const float *ndc_pts, const float *feat_tiles, const float *angle,
                          int R, int S, float *raw, int mode, void *stream);
/* Operand tiles: the interchange format between K2 and the tensor-core MLP (UCN_MLP_BF16X3_TC).  One record of
 * ucn_operand_tile_bytes(V) bytes per 128-sample tile: [film hi | film lo | img hi | img lo] K-chunks of 128 rows x 8 bf16
 * (hi = bf16(x), lo = bf16(x - hi); film = mvs feats 24 | (r,g,b,mask) x V | 1 when V is odd: 3 + ceil(V/2) chunks,
 * img = image feats 8 x V: V chunks; network/models.py:143-150) followed by the 128 sampled confidences (fp32).  Chunks
 * that would hold only the bias ones column or zero K padding are not stored (K3 writes those constants).  A chunk is
 * already the UMMA K-major core-matrix layout, so K3 copies it into shared memory / tensor memory verbatim.  V <= 7. */
size_t ucn_operand_tile_bytes(int V);
/* K2 writing operand tiles (same gathers as ucn_gather_tiles_from_z; network/renderer.py:177-212) */
int ucn_gather_operands_from_z(const ucn_frame *f, const float *z, const float *pix_yx, int ray0, int R,
                               void *operand_tiles, float *ndc_pts, void *stream);
/* K3 reading operand tiles (network/models.py:138-184): TMEM-resident activations, pipelined rounds */
int ucn_mlp_forward_operands(const ucn_weights *w, const float *ndc_pts, const void *operand_tiles, const float *angle,
                             int R, int S, float *raw, int mode, void *stream);
/* which kernel pair ucn_render_rays uses for UCN_MLP_BF16X3_TC: 2 (default) = operand tiles + ucn_mlp_forward_operands,
 * 1 = round-1 pair (tile-transposed fp32 rows + ucn_mlp_forward_tiles).  set = 0 queries; returns the current value. */
int ucn_mlp_x3_variant(int set);

/* ---- K4: alpha compositing -------------------------------------------------------------------
 * replaces raw2alpha / raw2outputs (network/renderer.py:25-36,109-140).  rgb [R][3], depth [R];
 * acc/weights optional (acc [R], weights [R][S]); white_bkgd as in renderer.py:138-139. */
int ucn_composite(const float *raw, const float *z, int R, int S, int white_bkgd, float *rgb, float *depth,
                  float *acc, float *weights, void *stream);

/* ---- training step: backward kernels and the fp32 training primitives -----------
 * The reference trains through autograd (train.py:147-190).  Sample positions carry no gradient (depth hypotheses
 * are detached, network/mvs_models.py:716-717); gradients reach the MLP parameters, the three feature volumes, the
 * source-view FPN features and the confidence map.  uc-nerf_b200/train_path.py wires these into torch.autograd. */
/* K4b: backward of raw2outputs (network/renderer.py:25-36,125-131): cotangents g_rgb [R][3], g_depth [R] -> d_raw [R][S][4] */
int ucn_composite_bwd(const float *raw, const float *z, int R, int S, int white_bkgd, const float *g_rgb,
                      const float *g_depth, float *d_raw, void *stream);
/* K2b: backward of the grid_sample groups of gen_pts_feats (utils/utils.py:776-788,859-890): d_feat [n][F] is
 * scatter-added (fp32 atomics) into gradients laid out like the reference tensors: d_volK [8][D_k][h_k][w_k],
 * d_img_feat [V][8][H][W], d_conf [H][W]; any destination may be NULL.  Only dims and cameras of `f` are read. */
int ucn_gather_bwd(const ucn_frame *f, const float *pts, const float *ndc_s1, const float *ndc_s2, const float *ndc_s3,
                   int64_t n_samples, const float *d_feat, float *d_vol0, float *d_vol1, float *d_vol2,
                   float *d_img_feat, float *d_conf, void *stream);
/* K3b primitives: nn.Linear forward / grad-input / grad-weight on row-major fp32 operands with explicit leading
 * dimensions; W and dW in the reference's [out][in] layout (network/models.py:87-104). */
/* 0 (default): products that fill a 128-wide tile run on tcgen05 with split-bf16 operands (a_hi*b_hi + a_lo*b_hi +
 * a_hi*b_lo, fp32 accumulate in TMEM); 1: fp32 CUDA cores only. */
int ucn_train_gemm_mode(int mode);
int ucn_linear_fwd(const float *X, int64_t ldx, const float *W, int64_t ldw, const float *b, int64_t M, int K, int N,
                   float *Y, int64_t ldy, int flags /* 1: accumulate into Y, 2: ReLU */, void *stream);
int ucn_linear_bwd_input(const float *dY, int64_t ldy, const float *W, int64_t ldw, int64_t M, int K, int N, float *dX,
                         int64_t ldx, int flags /* 1: accumulate into dX */, void *stream);
/* dW += dY^T X, db += column sums of dY (atomic accumulation: zero the destinations first; db may be NULL) */
int ucn_linear_bwd_weight(const float *dY, int64_t ldy, const float *X, int64_t ldx, int64_t M, int K, int N, float *dW,
                          int64_t lddw, float *db, void *stream);
/* ph[:, 0:63] = PE_10(ndc [N][3]); ftv[:, 128:155] = PE_4(angle[ray]) (network/models.py:50-54, renderer.py:88-100) */
int ucn_train_prep(const float *ndc, const float *angle, int S, int64_t N, float *ph, int64_t ldph, float *ftv,
                   int64_t ldftv, void *stream);
/* h = relu(pre * mod) over [M][128] (network/models.py:154-155) */
int ucn_film_relu_fwd(const float *pre, const float *mod, float *h, int64_t ldh, int64_t M, void *stream);
/* backward: dpre = dh * mod * [pre*mod > 0], dmod (+)= dh * pre * [pre*mod > 0]; db (optional) += column sums of dpre
 * (the bias gradient of the layer that produced pre) */
int ucn_film_relu_bwd(const float *dh, int64_t lddh, const float *pre, const float *mod, float *dpre, float *dmod,
                      int accumulate, float *db, int64_t M, void *stream);
/* o = a * b over [M][128] (h * feats_bias, network/models.py:164); backward: da (+)= do*b, db = do*a */
int ucn_mul_fwd(const float *a, int64_t lda, const float *b, float *o, int64_t M, void *stream);
int ucn_mul_bwd(const float *d_o, const float *a, int64_t lda, const float *b, float *da, int accumulate, float *db,
                int64_t M, void *stream);
int ucn_relu_bwd(const float *dy, const float *y, float *dx, int64_t n, void *stream);
/* uncertainty-weighted blend of the base and adapt heads (network/models.py:149,177-180); conf = input_feat[..., -1] */
int ucn_blend_fwd(const float *base4, const float *adapt4, const float *conf, int64_t ldconf, float *raw, int64_t M,
                  void *stream);
int ucn_blend_bwd(const float *d_raw, const float *raw, const float *base4, const float *adapt4, const float *conf,
                  int64_t ldconf, float *d_base4, float *d_adapt4, float *d_conf, int64_t lddconf, int64_t M,
                  void *stream);

/* ---- whole frame --------------------------------------------------------------------------- */
size_t ucn_frame_workspace_bytes(const ucn_frame *f, int R, int mode);
/* K1 -> K2 -> K3 -> K4 for rays [ray0, ray0+R) on `stream`, everything device-resident.
 * replaces the chunk loop of train.py:251-275.  out_rgb [R][3], out_depth [R]. */
int ucn_render_rays(const ucn_frame *f, const ucn_weights *w, int ray0, int R, const float *t_vals,
                    const float *t_rand, uint64_t seed, int mode, void *workspace, size_t workspace_bytes,
                    float *out_rgb, float *out_depth, void *stream);

/* ---- renderer-side training loss, value and gradient in one launch ------------------- */
/* replaces train.py:164-188 (img2mse, weighted sparse-depth MSE, weighted sum), network/mvs_models.py:491-509
 * (EdgePreservingSmoothnessLoss), utils/loss.py:54-112 (GradientLoss) and their autograd graph.
 * terms5 = {loss, img, nerf_depth, smooth, scale_invariant}; d_rgb [R][3], d_depth [R] = d loss / d (rgb, depth).
 * rays [0, patch_num*patch_size^2): patches (first half smoothness, second half scale-invariant); [N_rays, R): sparse depth */
int ucn_render_losses(const float *rgb, const float *depth, const float *target_rgb, const int64_t *pix_rows,
                      const int64_t *pix_cols, int R, int N_rays, int patch_num, int patch_size,
                      const float *sparse_depths, const float *sparse_weights, const float *dpt, int H, int W,
                      float *terms5, float *d_rgb, float *d_depth, void *stream);

/* ---- the step before the path: homography warp + variance cost volume -------------- */
/* replaces utils/utils.py:1105-1172 (homo_warp) and the per-view loop of network/mvs_models.py:603-628.
 * B = 1.  depth_values [D][H+2pad][W+2pad] (already replicate-padded), matrices (src_proj @ ref_proj_inv)[:3]
 * row-major in HOST memory.  warped [C][D][Hp][Wp]; grids [D][Hp*Wp][2] normalised coordinates. */
int ucn_homo_warp(const float *src_feat, int C, int H, int W, const float *proj12_host, const float *depth_values,
                  const float *grid_in, int D, int pad, float *warped, float *grid_out, void *stream);
int ucn_cost_volume_variance(const float *feats, int V, int C, int H, int W, const float *projs_host,
                             const float *depth_values, int D, int pad, int extra_ones, float *var, void *stream);

/* ---- evaluation metrics of a rendered frame, device-resident ---------------------- */
/* replaces the host-side numpy evaluation of utils/utils.py:18-22 (get_psnr) and utils/evaluation.py:8-74
 * (compute_errors, depth_evaluation).  Accumulators are fp64 and live in device memory; the caller zeroes them. */
size_t ucn_metrics_workspace_bytes(void);
/* *out_sum += sum (a[i]-b[i])^2;  PSNR = -10 log10(sum / n) */
int ucn_sq_err_sum(const float *a, const float *b, int64_t n, double *out_sum, void *stream);
/* SSIM as the reference's rgb_evaluation takes it (utils/evaluation.py:93: skimage.metrics.structural_similarity with
 * data_range = 1 and the library's defaults -- scikit-image >= 0.21 per requirements.txt:8, not vendored: 7 x 7 uniform
 * window, sample covariance, K1 = 0.01, K2 = 0.03, mean over the map cropped by 3 pixels, then over channels).
 * a, b [C][H][W] fp32; *out_sum += the sum of the SSIM map over channels and interior pixels (fp64; caller zeroes it);
 * mean SSIM = sum / (C (H-6) (W-6)). */
int ucn_ssim_sum(const float *a, const float *b, int C, int H, int W, float data_range, double *out_sum, void *stream);
/* exact medians of valid gt / pred depths (valid: min_depth < gt < max_depth and mask != 0, mask may be NULL):
 * out_mid4 = {gt, gt, pred, pred} values of ranks (c-1)/2 and c/2 (numpy's median is their mean), *out_count = c */
int ucn_depth_median(const float *gt, const float *pred, const unsigned char *mask, int64_t n, float min_depth,
                     float max_depth, void *workspace, size_t workspace_bytes, float *out_mid4, int64_t *out_count,
                     void *stream);
/* out8 += {sum |gt-p|/gt, sum (gt-p)^2/gt, sum (gt-p)^2, sum (log gt - log p)^2, #[t<1.25], #[t<1.25^2],
 * #[t<1.25^3], #valid},  p = clamp(pred*ratio, min_depth, max_depth),  t = max(gt/p, p/gt) */
int ucn_depth_errors(const float *gt, const float *pred, const unsigned char *mask, int64_t n, float min_depth,
                     float max_depth, float ratio, double *out8, void *stream);

/* ---- training sampler: ray selection in one launch (utils/utils.py:450-479,169-215,217-247,274-320) ------------- */
/* pix_yx [2][R] fp32 (row, col), R = 2*(patch_num/2)*patch_size^2 + n_rand + n_depth: patch_num/2 patches drawn by a
 * confidence-weighted multinomial without replacement, patch_num/2 by (1 - confidence) (patch offset clamped to
 * [0, H/ps - 2], uniform shift in [0, ps) per patch), n_rand uniform pixels (inside the centre crop when crop != 0),
 * then the n_depth sparse-depth coordinates coords_rc [n_depth][2] int64 (row, col).  Same distribution as the
 * reference's torch.multinomial / randint sequence (the exponential race torch runs itself), keyed by `seed`.
 * workspace: ucn_select_workspace_bytes() bytes, zero-filled ONCE by the caller before its first use. */
size_t ucn_select_workspace_bytes(int H, int W, int patch_num);
int ucn_select_training_pixels(const float *confidence, int H, int W, int n_rand, int patch_num, int patch_size, int crop,
                               const long long *coords_rc, int n_depth, uint64_t seed, void *workspace, float *pix_yx,
                               void *stream);

/* The same selection for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  The kernel's workspace
 * is a file-scope `static __device__` array of ucn_select_static_scratch_bytes() = 33 024 bytes that every GPU loading
 * the library holds until exit; the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine), the
 * caller passes inputs and outputs only.  LIMIT: ceil(H*W / 2048) * (patch_num / 2) <= 2048 (the same bound
 * ucn_select_training_pixels has; e.g. patch_num 16 up to 524 288 pixels); larger requests return UCN_E_ARG.
 * All calls on one GPU share the array: calls that overlap on the GPU (two streams) race -- serialise them.
 * Output: bit-identical to ucn_select_training_pixels for the same arguments and seed. */
size_t ucn_select_static_scratch_bytes(void);
int ucn_select_training_pixels_static(const float *confidence, int H, int W, int n_rand, int patch_num, int patch_size,
                                      int crop, const long long *coords_rc, int n_depth, uint64_t seed, float *pix_yx,
                                      void *stream);

/* ---- multi-GPU: peer-mapped frame buffer (the reference is single-GPU, train.py:433) ------ */
/* One process per GPU.  The root allocates the [H*W][4]-float frame buffer and exports a 64-byte CUDA IPC handle;
 * every other process opens it with its own device current and passes tile pointers inside it as ucn_render_rays'
 * out_rgb / out_depth, so K4's stores land in the root's HBM over NVLink (no gather collective). */
#define UCN_PEER_HANDLE_BYTES 64
int ucn_peer_alloc(size_t bytes, void **dptr, unsigned char *handle);
int ucn_peer_open(const unsigned char *handle, void **dptr);
int ucn_peer_close(void *dptr);
int ucn_peer_free(void *dptr);

#ifdef __cplusplus
}
#endif
#endif /* UCNERF_H */
