// ucn_gather.cu -- K2: per-sample projection into the source views + fused bilinear/trilinear gathers,
// assembled into the reference's input_feat row layout.
//
// replaces (reference): utils/utils.py:833-893 index_point_feature, :742-799 build_color_volume,
//   :323-347 get_ndc_coordinate(sample_2d=True); network/renderer.py:177-212 gen_pts_feats.
//
// Data layout (see ucn_pack.cu): all maps are channels-last so one texel is one 16/32-byte vector load
// (volume/feature texel = 8 fp32 = one 32 B sector; rgb texel = float4); volumes are [h][w][D][8], depth
// innermost, because the 32 consecutive samples of a ray that a warp owns share their (x,y) columns and differ
// only in depth: their corner texels are then neighbours in memory (a few 128-byte lines per load instruction
// instead of one line per lane).  One thread owns one sample.  Row-major output ([N][F], the reference layout):
// a 128-thread block stages its rows in shared memory and writes one contiguous 128*F*4-byte span.  Tiled output
// ([tile][F][128], the tensor-core MLP's interchange layout): every channel is stored straight from registers,
// one coalesced 128-byte line per warp and channel, no shared memory.
//
// Roofline: HBM-bound in this staged form.  Algorithmic bytes per sample = 4F (row out) + 12 (ndc out)
// + 4 (z in) = 4F+16, plus the maps once per frame (they are L2-resident: 62 MB at 320x256).
#include "ucn_common.cuh"

namespace {

constexpr int kThreads = 128;

struct Acc8 { float v[8]; };

// one 8-channel texel = 32 bytes = ONE 256-bit load (LDG.E.256, sm_100): half the load instructions and L1 requests
// of two float4 loads
__device__ __forceinline__ void ldg8(const float *__restrict__ p, float4 &lo, float4 &hi) {
  asm("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
      : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w) : "l"(p));
}
__device__ __forceinline__ void fma8(Acc8 &a, const float *__restrict__ texel, float w) {
  float4 lo, hi;
  ldg8(texel, lo, hi);
  a.v[0] = fmaf(lo.x, w, a.v[0]); a.v[1] = fmaf(lo.y, w, a.v[1]);
  a.v[2] = fmaf(lo.z, w, a.v[2]); a.v[3] = fmaf(lo.w, w, a.v[3]);
  a.v[4] = fmaf(hi.x, w, a.v[4]); a.v[5] = fmaf(hi.y, w, a.v[5]);
  a.v[6] = fmaf(hi.z, w, a.v[6]); a.v[7] = fmaf(hi.w, w, a.v[7]);
}

// trilinear, border padding, align_corners=False, over a [h][w][D][8] volume; channel c -> out[c * os].
// Branch-free: where the reference skips an out-of-range corner (ATen within_bounds test), the corner index is clamped
// instead and its weight is exactly 0 there (the coordinate was clipped to size-1, so w1 = x - floor(x) = 0): the
// fma adds +0.  32-bit texel indices (launcher checks every map has < 2^28 texels).
__device__ __forceinline__ Acc8 tri8(const float *__restrict__ vol, int D, int h, int w, float cx, float cy, float cz,
                                     int32_t *__restrict__ idx) {
  const ucn::Lerp1 ax = ucn::axis(ucn::to_grid(cx), w, false);
  const ucn::Lerp1 ay = ucn::axis(ucn::to_grid(cy), h, false);
  const ucn::Lerp1 az = ucn::axis(ucn::to_grid(cz), D, false);
  if (idx) { idx[0] = ax.i0; idx[1] = ay.i0; idx[2] = az.i0; }
  const uint32_t x1 = ax.has1 ? 1u : 0u, y1 = ay.has1 ? (uint32_t)w : 0u, z1 = az.has1 ? 1u : 0u;
  const uint32_t col00 = ((uint32_t)ay.i0 * (uint32_t)w + (uint32_t)ax.i0) * (uint32_t)D + (uint32_t)az.i0;
  const uint32_t Du = (uint32_t)D;
  Acc8 a = {};
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int dx = c & 1, dy = (c >> 1) & 1, dz = c >> 2;
    const float wgt = __fmul_rn(__fmul_rn(dx ? ax.w1 : ax.w0, dy ? ay.w1 : ay.w0), dz ? az.w1 : az.w0);
    const uint32_t t = col00 + ((dy ? y1 : 0u) + (dx ? x1 : 0u)) * Du + (dz ? z1 : 0u);
    fma8(a, vol + (size_t)t * 8, wgt);
  }
  return a;
}

// ---- operand tiles (OUT = 2): what the tensor-core MLP's shared-memory / TMEM operands are made of -----------------
// One 128-sample tile = one record of ucn_operand_tile_bytes(V) bytes:
//   film hi | film lo   3 + ceil(V/2) chunks each: [mvs feats 24 | (r,g,b,mask) x V | 1 (V odd)]   (models.py:143-150)
//   img  hi | img  lo   V chunks each:             [img feats 8 x V]                              (models.py:147)
//   confidence          128 fp32                                                                   (models.py:149)
// Only chunks that carry data are stored: the chunk that holds nothing but the ones column (V even: film; always: img) and
// the all-zero K padding up to a multiple of 16 are constants K3 writes itself.
// a chunk = 8 consecutive K values of the 128 rows as bf16, [row][8] (2 KB) -- the UMMA no-swizzle K-major core-matrix
// layout, so K3 copies chunks verbatim (cp.async / 128-bit loads); hi = bf16(x), lo = bf16(x - hi), the split K3's
// round-1 kernel did itself.  The "1" is the ones column that folds the layer's bias into the MMA.
struct OperGeom { int kf8, kb8; };                         // stored (data) chunks per half
__host__ __device__ __forceinline__ OperGeom oper_geom(int V) { return {3 + (V + 1) / 2, V}; }
__device__ __forceinline__ uint32_t bf2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void emit_chunk(unsigned char *__restrict__ hi_base, unsigned char *__restrict__ lo_base, int chunk,
                                           const float *v) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    h[i] = bf2(v[2 * i], v[2 * i + 1]);
    l[i] = bf2(v[2 * i] - __uint_as_float(h[i] << 16), v[2 * i + 1] - __uint_as_float(h[i] & 0xFFFF0000u));
  }
  *reinterpret_cast<uint4 *>(hi_base + (size_t)chunk * 2048) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4 *>(lo_base + (size_t)chunk * 2048) = make_uint4(l[0], l[1], l[2], l[3]);
}

template <bool FROM_Z, int OUT>        // OUT: 0 row-major [N][F], 1 tile-transposed [tile][F][128], 2 operand tiles
__global__ void __launch_bounds__(kThreads, OUT ? 8 : 7)
gather_kernel(const __grid_constant__ ucn_frame f, const float *__restrict__ pts_or_z,
              const float *__restrict__ ndc1, const float *__restrict__ ndc2, const float *__restrict__ ndc3,
              const float *__restrict__ pix_yx, int ray0, int R, int64_t n_samples, int parts,
              float *__restrict__ feat, float *__restrict__ ndc_out, int32_t *__restrict__ idx_out) {
  extern __shared__ float rows[];                       // [kThreads][F + 1]  (+1: odd stride, no bank conflicts)
  const int V = f.V, F = 24 + 12 * V + 1, ld = F + 1;
  const int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  // channel c of this sample goes to row[c * os]: shared-memory staging row, or (tiled) straight to the
  // [F][128] tile of this block in global memory
  // (TILED is a template parameter so that the 61 channel stores compile to STG / STS instead of generic ST, and the
  //  tiled instance, which needs no shared memory, is held to 64 registers: 8 blocks per SM instead of 7)
  constexpr bool TILED = OUT == 1, OPER = OUT == 2;
  constexpr int os = TILED ? kThreads : 1;
  float *row = TILED ? feat + (int64_t)blockIdx.x * F * kThreads + threadIdx.x : rows + (size_t)threadIdx.x * ld;
  const OperGeom og = oper_geom(V);
  unsigned char *rec = reinterpret_cast<unsigned char *>(feat) + (size_t)blockIdx.x * ((2 * og.kf8 + 2 * og.kb8) * 2048 + 512) +
                       threadIdx.x * 16;                       // this row's 16 bytes of chunk 0 of the tile's record
  unsigned char *film_hi = rec, *film_lo = rec + og.kf8 * 2048, *img_hi = rec + 2 * og.kf8 * 2048,
                *img_lo = img_hi + og.kb8 * 2048;
  float cbuf[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};          // OPER: (r,g,b,mask) of two views = one film chunk
  if (q < n_samples) {
    const float iw = (float)(f.W - 1), ih = (float)(f.H - 1);
    float p[3], c1[3], c2[3], c3[3];
    if (FROM_Z) {
      const int r = (int)((uint32_t)q / (uint32_t)f.S);         // launcher: n_samples < 2^31
      float ys, xs, d[3], nr[3], fr[3], nx, ny, pz;
      ucn::pixel_of_ray(f, pix_yx, ray0, R, r, ys, xs);
      ucn::ray_dir(f, ys, xs, d);
      ucn::stage_near_far(f, ys, xs, nr, fr);
      const float zj = pts_or_z[q];
      p[0] = __fadd_rn(f.c2w[3], __fmul_rn(zj, d[0]));
      p[1] = __fadd_rn(f.c2w[7], __fmul_rn(zj, d[1]));
      p[2] = __fadd_rn(f.c2w[11], __fmul_rn(zj, d[2]));
      ucn::project(f.w2c, f.K, iw, ih, p, nx, ny, pz);
      c1[0] = c2[0] = c3[0] = nx;
      c1[1] = c2[1] = c3[1] = ny;
      c1[2] = __fdiv_rn(__fsub_rn(pz, nr[0]), __fsub_rn(fr[0], nr[0]));
      c2[2] = __fdiv_rn(__fsub_rn(pz, nr[1]), __fsub_rn(fr[1], nr[1]));
      c3[2] = __fdiv_rn(__fsub_rn(pz, nr[2]), __fsub_rn(fr[2], nr[2]));
      if (ndc_out) {
        ndc_out[3 * q] = nx; ndc_out[3 * q + 1] = ny;
        ndc_out[3 * q + 2] = __fdiv_rn(__fsub_rn(pz, f.near), __fsub_rn(f.far, f.near));
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        p[c] = pts_or_z[3 * q + c];
        c1[c] = ndc1[3 * q + c]; c2[c] = ndc2[3 * q + c]; c3[c] = ndc3[3 * q + c];
      }
    }
    int32_t *ir = idx_out ? idx_out + q * (11 + 3 * V) : nullptr;
    if (!OPER && parts != UCN_PART_ALL) {
      for (int c = 0; c < F; ++c) row[c * os] = 0.0f;
      if (ir) for (int c = 0; c < 11 + 3 * V; ++c) ir[c] = 0;
    }
    if (parts & UCN_PART_VOLUMES) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float *cc = k == 0 ? c1 : (k == 1 ? c2 : c3);
      const Acc8 a = tri8(f.vol[k], f.D[k], f.h[k], f.w[k], cc[0], cc[1], cc[2], ir ? ir + 3 * k : nullptr);
      if (OPER) emit_chunk(film_hi, film_lo, k, a.v);
      else {
#pragma unroll
        for (int c = 0; c < 8; ++c) row[(8 * k + c) * os] = a.v[c];
      }
    }
    {   // confidence: depth-1 "volume" => bilinear, align_corners=False (renderer.py:195, utils.py:859-865)
      const ucn::Lerp1 ax = ucn::axis(ucn::to_grid(c3[0]), f.W, false), ay = ucn::axis(ucn::to_grid(c3[1]), f.H, false);
      float acc = 0.0f;
      const uint32_t t00 = (uint32_t)ay.i0 * (uint32_t)f.W + (uint32_t)ax.i0;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int dx = c & 1, dy = c >> 1;
        const uint32_t t = t00 + ((dy && ay.has1) ? (uint32_t)f.W : 0u) + ((dx && ax.has1) ? 1u : 0u);
        acc = fmaf(__ldg(f.conf + t), __fmul_rn(dx ? ax.w1 : ax.w0, dy ? ay.w1 : ay.w0), acc);
      }
      if (OPER) reinterpret_cast<float *>(rec - threadIdx.x * 16 + (2 * og.kf8 + 2 * og.kb8) * 2048)[threadIdx.x] = acc;
      else row[(F - 1) * os] = acc;
      if (ir) { ir[9] = ax.i0; ir[10] = ay.i0; }
    }
    }
    const size_t HW = (size_t)f.H * f.W;
    for (int v = 0; v < ((parts & UCN_PART_VIEWS) ? V : 0); ++v) {   // utils/utils.py:766-798
      float nx, ny, pz;
      ucn::project(f.src_w2c + 16 * v, f.src_K + 9 * v, iw, ih, p, nx, ny, pz);
      const float gx = ucn::to_grid(nx), gy = ucn::to_grid(ny);
      const ucn::Lerp1 ax = ucn::axis(gx, f.W, true), ay = ucn::axis(gy, f.H, true);
      const float m = (gx > -1.0f && gx < 1.0f && gy > -1.0f && gy < 1.0f) ? 1.0f : 0.0f;
      float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
      Acc8 a = {};
      const float *rgb = f.rgb + (size_t)v * HW * 4, *ft = f.feat + (size_t)v * HW * 8;
      const uint32_t t00 = (uint32_t)ay.i0 * (uint32_t)f.W + (uint32_t)ax.i0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int dx = k & 1, dy = k >> 1;
        const float wgt = __fmul_rn(dx ? ax.w1 : ax.w0, dy ? ay.w1 : ay.w0);
        const size_t t = t00 + ((dy && ay.has1) ? (uint32_t)f.W : 0u) + ((dx && ax.has1) ? 1u : 0u);
        const float4 px = ucn::ldg4(rgb + t * 4);
        c.x = fmaf(px.x, wgt, c.x); c.y = fmaf(px.y, wgt, c.y); c.z = fmaf(px.z, wgt, c.z);
        fma8(a, ft + t * 8, wgt);
      }
      if (OPER) {
        const int h4 = (v & 1) * 4;
        cbuf[h4] = c.x; cbuf[h4 + 1] = c.y; cbuf[h4 + 2] = c.z; cbuf[h4 + 3] = m;
        if (v & 1) emit_chunk(film_hi, film_lo, 3 + (v >> 1), cbuf);
        emit_chunk(img_hi, img_lo, v, a.v);
      } else {
        float *o = row + (24 + 4 * v) * os;
        o[0] = c.x; o[os] = c.y; o[2 * os] = c.z; o[3 * os] = m;
        float *of = row + (24 + 4 * V + 8 * v) * os;
#pragma unroll
        for (int k = 0; k < 8; ++k) of[k * os] = a.v[k];
      }
      if (ir) { ir[11 + 3 * v] = ax.i0; ir[12 + 3 * v] = ay.i0; ir[13 + 3 * v] = (int32_t)m; }
    }
  }
  if (OPER) {
    // the ones column (bias) closes the film / img operands; K padding and the rows past the last sample are zeros
    const float zero8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (q < n_samples) {
      float tail[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      int kc = 3 + (V >> 1);
      if (V & 1) {                                           // last view sits in the low half of its chunk; the 1 follows it
        tail[0] = cbuf[0]; tail[1] = cbuf[1]; tail[2] = cbuf[2]; tail[3] = cbuf[3]; tail[4] = 1.0f;
      } else tail[0] = 1.0f;
      if (V & 1) emit_chunk(film_hi, film_lo, kc, tail);     // (V even: the ones column is a chunk of its own -- not stored)
    } else {
      for (int kc = 0; kc < og.kf8; ++kc) emit_chunk(film_hi, film_lo, kc, zero8);
      for (int kc = 0; kc < og.kb8; ++kc) emit_chunk(img_hi, img_lo, kc, zero8);
      reinterpret_cast<float *>(rec - threadIdx.x * 16 + (2 * og.kf8 + 2 * og.kb8) * 2048)[threadIdx.x] = 0.0f;
    }
    return;
  }
  if (TILED) {
    // tile-transposed rows for the tensor-core MLP: block == one 128-sample tile, laid out [F][128] so that both
    // the stores above and the MLP's per-row loads are one coalesced 128-byte line per warp and channel
    if (q >= n_samples)
      for (int c = 0; c < F; ++c) row[(int64_t)c * kThreads] = 0.0f;
    return;
  }
  __syncthreads();
  // coalesced write-out of the block's contiguous [rows_here][F] span
  const int64_t q0 = (int64_t)blockIdx.x * kThreads;
  const int rows_here = (int)min((int64_t)kThreads, n_samples - q0);
  float *dst = feat + q0 * F;
  for (int e = threadIdx.x; e < rows_here * F; e += kThreads) dst[e] = rows[(e / F) * ld + e % F];
}

}  // namespace

static int launch_gather(bool from_z, const ucn_frame *f, const float *a, const float *n1, const float *n2,
                         const float *n3, const float *pix_yx, int ray0, int R, int64_t n, int parts, int tiled,
                         float *feat, float *ndc_out, int32_t *idx, void *stream) {
  UCN_CHECK_ARG(f && a && feat, "ucn_gather: null argument");
  UCN_CHECK_ARG(f->V >= 1 && f->V <= UCN_MAX_V, "ucn_gather: 1 <= V <= 8 source views");
  UCN_CHECK_ARG(n < (1ll << 31), "ucn_gather: more than 2^31 samples in one call; split the ray range");
  for (int k = 0; k < 3; ++k)
    UCN_CHECK_ARG((long long)f->D[k] * f->h[k] * f->w[k] < (1ll << 28), "ucn_gather: volume has 2^28 texels or more");
  UCN_CHECK_ARG((long long)f->H * f->W < (1ll << 28), "ucn_gather: image has 2^28 pixels or more");
  if (n <= 0) return UCN_OK;
  const int F = 24 + 12 * f->V + 1;
  const size_t smem = tiled ? 0 : (size_t)kThreads * (F + 1) * sizeof(float);
  const unsigned blocks = (unsigned)((n + kThreads - 1) / kThreads);
  const cudaStream_t st = (cudaStream_t)stream;
  static int attr_a[kUcnMaxDevices], attr_b[kUcnMaxDevices];
  if (!tiled) {
    cudaError_t e = ucn_ensure_dyn_smem(gather_kernel<true, 0>, 100 * 1024, attr_a);
    if (e == cudaSuccess) e = ucn_ensure_dyn_smem(gather_kernel<false, 0>, 100 * 1024, attr_b);
    if (e != cudaSuccess) return ucn_cuda_status(e, "gather smem attribute");
  }
  if (from_z && tiled == 2)
    gather_kernel<true, 2><<<blocks, kThreads, 0, st>>>(*f, a, n1, n2, n3, pix_yx, ray0, R, n, parts, feat, ndc_out, idx);
  else if (from_z && tiled)
    gather_kernel<true, 1><<<blocks, kThreads, 0, st>>>(*f, a, n1, n2, n3, pix_yx, ray0, R, n, parts, feat, ndc_out, idx);
  else if (from_z)
    gather_kernel<true, 0><<<blocks, kThreads, smem, st>>>(*f, a, n1, n2, n3, pix_yx, ray0, R, n, parts, feat, ndc_out, idx);
  else if (tiled)
    gather_kernel<false, 1><<<blocks, kThreads, 0, st>>>(*f, a, n1, n2, n3, pix_yx, ray0, R, n, parts, feat, ndc_out, idx);
  else
    gather_kernel<false, 0><<<blocks, kThreads, smem, st>>>(*f, a, n1, n2, n3, pix_yx, ray0, R, n, parts, feat, ndc_out, idx);
  UCN_LAUNCH_CHECK("gather_kernel");
  return UCN_OK;
}

extern "C" int ucn_gather(const ucn_frame *f, const float *pts, const float *ndc_s1, const float *ndc_s2,
                          const float *ndc_s3, int64_t n_samples, int parts, float *feat, int32_t *idx,
                          void *stream) {
  UCN_CHECK_ARG(ndc_s1 && ndc_s2 && ndc_s3, "ucn_gather: null ndc");
  UCN_CHECK_ARG(parts >= 1 && parts <= 3, "ucn_gather: parts must be 1, 2 or 3");
  return launch_gather(false, f, pts, ndc_s1, ndc_s2, ndc_s3, nullptr, 0, 0, n_samples, parts, 0, feat, nullptr, idx,
                       stream);
}

extern "C" int ucn_gather_from_z(const ucn_frame *f, const float *z, const float *pix_yx, int ray0, int R,
                                 float *feat, float *ndc_pts, void *stream) {
  UCN_CHECK_ARG(f && f->S > 0, "ucn_gather_from_z: bad frame");
  return launch_gather(true, f, z, nullptr, nullptr, nullptr, pix_yx, ray0, R, (int64_t)R * f->S, UCN_PART_ALL, 0, feat,
                       ndc_pts, nullptr, stream);
}

extern "C" int ucn_gather_tiles_from_z(const ucn_frame *f, const float *z, const float *pix_yx, int ray0, int R,
                                       float *feat_tiles, float *ndc_pts, void *stream) {
  UCN_CHECK_ARG(f && f->S > 0, "ucn_gather_tiles_from_z: bad frame");
  return launch_gather(true, f, z, nullptr, nullptr, nullptr, pix_yx, ray0, R, (int64_t)R * f->S, UCN_PART_ALL, 1,
                       feat_tiles, ndc_pts, nullptr, stream);
}

extern "C" size_t ucn_operand_tile_bytes(int V) {
  const OperGeom og = oper_geom(V);
  return (size_t)(2 * og.kf8 + 2 * og.kb8) * 2048 + 512;
}

extern "C" int ucn_gather_operands_from_z(const ucn_frame *f, const float *z, const float *pix_yx, int ray0, int R,
                                          void *operand_tiles, float *ndc_pts, void *stream) {
  UCN_CHECK_ARG(f && f->S > 0, "ucn_gather_operands_from_z: bad frame");
  UCN_CHECK_ARG(f->V <= 7, "ucn_gather_operands_from_z: the tensor-core MLP takes at most 7 source views");
  return launch_gather(true, f, z, nullptr, nullptr, nullptr, pix_yx, ray0, R, (int64_t)R * f->S, UCN_PART_ALL, 2,
                       static_cast<float *>(operand_tiles), ndc_pts, nullptr, stream);
}

namespace {
struct Cam { float w2c[16], K[9]; };
__global__ void project_kernel(const __grid_constant__ Cam cam, float iw, float ih, const float *__restrict__ pts,
                               int64_t n, float *__restrict__ out) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n) return;
  const float p[3] = {pts[3 * q], pts[3 * q + 1], pts[3 * q + 2]};
  float nx, ny, pz;
  ucn::project(cam.w2c, cam.K, iw, ih, p, nx, ny, pz);
  out[3 * q] = nx; out[3 * q + 1] = ny; out[3 * q + 2] = pz;
}
}  // namespace

extern "C" int ucn_project_points(const float *w2c_host16, const float *K_host9, int H, int W, const float *pts,
                                  int64_t n, float *out, void *stream) {
  UCN_CHECK_ARG(w2c_host16 && K_host9 && pts && out, "ucn_project_points: null argument");
  if (n <= 0) return UCN_OK;
  Cam cam;
  for (int i = 0; i < 16; ++i) cam.w2c[i] = w2c_host16[i];
  for (int i = 0; i < 9; ++i) cam.K[i] = K_host9[i];
  project_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(cam, (float)(W - 1), (float)(H - 1), pts, n, out);
  UCN_LAUNCH_CHECK("project_kernel");
  return UCN_OK;
}
