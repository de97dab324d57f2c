// ucn_weights.cuh -- packed weight layouts shared by the pack kernel and the two MLP kernels.
//
// fp32 pack (SIMT parity mode): every matrix transposed to [K][N] row-major with K padded so that the
// activation columns the kernel keeps in shared memory line up (zero rows where the activation buffer
// has a zero pad column):
//   film  [24+4V][128]  fb [8V][128]
//   pts0  [64][128]   (row 63 zero)          pts1..4 [128][128]
//   pts5  [192][128]  (rows 0..62 = PE, row 63 zero, rows 64..191 = h)      skip concat, models.py:156-157
//   feat  [128][128]
//   vv    [156][128]  columns 0..63 = views_linears.0, 64..127 = view_confi_linears.0; rows 0..127 =
//                     feature, 128..154 = PE(view dir), row 155 zero
//   base  [128][4]    columns = confi_rgb_linear (3) | alpha_linear_1 (1)
//   adapt [128][4]    rows 0..63 -> rgb_linear (cols 0..2), rows 64..127 -> alpha_linear (col 3)
// Biases follow each matrix (N floats).
#pragma once
#include <stdint.h>

namespace ucn {

struct F32Layout {
  int V, Kfilm, Kfb;
  size_t film, film_b, fb, fb_b, pts[6], pts_b[6], feat, feat_b, vv, vv_b, base, base_b, adapt, adapt_b, total;
  __host__ __device__ explicit F32Layout(int V_) : V(V_), Kfilm(24 + 4 * V_), Kfb(8 * V_) {
    size_t o = 0;
    auto take = [&](size_t n) { size_t r = o; o += (n + 3) & ~size_t(3); return r; };
    film = take((size_t)Kfilm * 128); film_b = take(128);
    fb = take((size_t)Kfb * 128); fb_b = take(128);
    const int Ks[6] = {64, 128, 128, 128, 128, 192};
    for (int i = 0; i < 6; ++i) { pts[i] = take((size_t)Ks[i] * 128); pts_b[i] = take(128); }
    feat = take(128 * 128); feat_b = take(128);
    vv = take(156 * 128); vv_b = take(128);
    base = take(128 * 4); base_b = take(4);
    adapt = take(128 * 4); adapt_b = take(4);
    total = o;
  }
};

// bf16 pack (tcgen05 mode): each B operand is an [N][K] K-major matrix stored as UMMA "interleaved"
// (no-swizzle) core matrices: element (n,k) at ((k/8)*N + n)*8 + (k%8)  [bf16 units]; i.e. LBO = N*16 B
// between K-chunks of 8, SBO = 128 B between 8-row groups.  K is padded to a multiple of 16 with zeros.
//   film [128][Kf16]  fb [128][Kb16]  pts0 [128][64]  pts1..4 [128][128]  pts5 [128][192]
//   feat [128][128]   vv [128][160]   base [16][128]  adapt [16][128]
// followed by an fp32 bias block: film_b, fb_b, pts_b[6], feat_b, vv_b (128 each), base_b, adapt_b (4 each, padded to 16).
struct BF16Layout {
  int V, Kfilm, Kfb;          // padded K
  size_t film, fb, pts[6], feat, vv, base, adapt, bias, total_bytes;   // byte offsets
  size_t b_film, b_fb, b_pts[6], b_feat, b_vv, b_base, b_adapt;        // float offsets inside the bias block
  __host__ __device__ explicit BF16Layout(int V_) : V(V_) {
    Kfilm = (24 + 4 * V_ + 1 + 15) & ~15;    // +1: row k = 24+4V holds the bias (A operand carries a ones column)
    Kfb = (8 * V_ + 1 + 15) & ~15;           // +1: row k = 8V holds the bias
    size_t o = 0;
    auto take = [&](size_t n_elems) { size_t r = o; o += (n_elems * 2 + 127) & ~size_t(127); return r; };
    film = take((size_t)128 * Kfilm);
    fb = take((size_t)128 * Kfb);
    const int Ks[6] = {64, 128, 128, 128, 128, 192};
    for (int i = 0; i < 6; ++i) pts[i] = take((size_t)128 * Ks[i]);
    feat = take(128 * 128);
    vv = take(128 * 160);
    base = take(16 * 128);
    adapt = take(16 * 128);
    bias = o;
    size_t b = 0;
    auto takeb = [&](size_t n) { size_t r = b; b += n; return r; };
    b_film = takeb(128); b_fb = takeb(128);
    for (int i = 0; i < 6; ++i) b_pts[i] = takeb(128);
    b_feat = takeb(128); b_vv = takeb(128); b_base = takeb(16); b_adapt = takeb(16);
    total_bytes = o + b * sizeof(float);
  }
};


// split-bf16 pack (tcgen05 "x3" mode, the parity-grade tensor-core path): every weight w is stored as
// hi = bf16(w) and lo = bf16(w - hi); the kernel issues A_hi*W_hi + A_lo*W_hi + A_hi*W_lo per K step,
// which carries ~16 significant bits per operand with fp32 accumulation.  A matrix [128][K] (K padded to
// 16) is a sequence of K/16 "units"; unit u = [hi tile 4096 B][lo tile 4096 B], each tile the UMMA
// no-swizzle K-major core-matrix layout of a 128 x 16 block: element (n,kk) at ((kk/8)*128 + n)*8 + kk%8.
//   film [Kfilm]  (row k = 24+4V carries the bias: the A operand has a ones column there)
//   pts0 [64]  pts1..4 [128]  pts5h [128] (skip layer, hidden part)  pts5pe [64] (skip layer, PE part)
//   fb [Kfb] (bias row at k = 8V)   feat [128]   vvh [128] (views|view_confi on the feature part)
// followed by an fp32 block: b_pts[6*128] b_feat[128] b_vv[128] w_base[128][4] w_adapt[128][4] b_base[4]
// b_adapt[4] (kSmemFloats floats, copied to shared memory) and vd[28][128] (view-direction rows of
// views|view_confi, applied per ray on the CUDA cores).
struct X3Layout {
  static constexpr int kUnit = 8192;
  static constexpr int kSmemFloats = 6 * 128 + 128 + 128 + 512 + 512 + 4 + 4;
  int V, Kfilm, Kfb;
  size_t film, pts0, pts14[4], pts5h, pts5pe, fb, feat, vvh, f32, total_bytes;   // byte offsets
  // float offsets inside the fp32 block
  static constexpr int b_pts = 0, b_feat = 768, b_vv = 896, w_base = 1024, w_adapt = 1536, b_base = 2048,
                       b_adapt = 2052, vd = 2056;
  __host__ __device__ explicit X3Layout(int V_) : V(V_) {
    Kfilm = (24 + 4 * V_ + 1 + 15) & ~15;
    Kfb = (8 * V_ + 1 + 15) & ~15;
    size_t o = 0;
    auto take = [&](int K) { size_t r = o; o += (size_t)(K / 16) * kUnit; return r; };
    film = take(Kfilm); pts0 = take(64);
    for (int i = 0; i < 4; ++i) pts14[i] = take(128);
    pts5h = take(128); pts5pe = take(64); fb = take(Kfb); feat = take(128); vvh = take(128);
    f32 = o;
    total_bytes = o + (size_t)(vd + 28 * 128) * sizeof(float);
  }
};

}  // namespace ucn
