// ucn_pack.cu -- K0: relayout of the encoder's maps and of the MLP weights into kernel-friendly form.
// Runs once per target view (maps) / once per weight load; HBM-bound, one pass over the data.
//
// The reference keeps NCDHW / NCHW maps and samples them with F.grid_sample, i.e. C channel-strided
// 4-byte loads per texel (utils/utils.py:776-788,859-890).  Channels-last makes a texel one vector load.
#include <limits.h>
#include <string.h>

#include <array>
#include <map>
#include <mutex>

#include <cuda_bf16.h>

#include "ucn_common.cuh"
#include "ucn_weights.cuh"

// host copies of the head weights [w_base 128x4 | w_adapt 128x4] per split-bf16 pack (keyed by the pack's device address;
// refreshed by every ucn_pack_weights into that address)
namespace {
std::mutex g_headw_mu;
std::map<const void *, std::array<float, 1024>> g_headw;
}  // namespace
void ucn_headw_store(const void *x3_pack, const float *w1024) {
  std::lock_guard<std::mutex> lk(g_headw_mu);
  std::array<float, 1024> a;
  memcpy(a.data(), w1024, sizeof(float) * 1024);
  g_headw[x3_pack] = a;
}
bool ucn_headw_lookup(const void *x3_pack, float *w1024) {
  std::lock_guard<std::mutex> lk(g_headw_mu);
  auto it = g_headw.find(x3_pack);
  if (it == g_headw.end()) return false;
  memcpy(w1024, it->second.data(), sizeof(float) * 1024);
  return true;
}

namespace {

__global__ void pack_cl_kernel(const float *__restrict__ src, float *__restrict__ dst, int Cn, int64_t N, int Cpad,
                               int64_t total) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // over B*N
  if (i >= total) return;
  const int64_t b = i / N, n = i - b * N;
  const float *s = src + (b * Cn) * N + n;
  float *d = dst + i * Cpad;
  if (Cpad == 4) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    v.x = s[0];
    if (Cn > 1) v.y = s[N];
    if (Cn > 2) v.z = s[2 * N];
    if (Cn > 3) v.w = s[3 * N];
    *reinterpret_cast<float4 *>(d) = v;
  } else if (Cpad == 8 && Cn == 8) {
    float4 lo = make_float4(s[0], s[N], s[2 * N], s[3 * N]);
    float4 hi = make_float4(s[4 * N], s[5 * N], s[6 * N], s[7 * N]);
    reinterpret_cast<float4 *>(d)[0] = lo;
    reinterpret_cast<float4 *>(d)[1] = hi;
  } else {
    for (int c = 0; c < Cpad; ++c) d[c] = c < Cn ? s[(int64_t)c * N] : 0.0f;
  }
}

// feature volume NCDHW [8][D][h][w] -> [h][w][D][8]: depth innermost, so the 126 samples of a ray (which all fall in
// the same <=4 (x,y) columns of the target-frustum volume, utils/utils.py:852-857) read one contiguous D*32-byte run
// per column instead of D planes h*w*32 bytes apart.  Block = one image row y x 16 columns x, staged through shared
// memory so that both the NCDHW reads (along x) and the packed writes (16*D*8 contiguous floats) are coalesced.
constexpr int kVolTileX = 16;
// (row band: block row y of the OUTPUT is source row y + row0; h is the source's row count)
__global__ void pack_volume_kernel(const float *__restrict__ src, float *__restrict__ dst, int D, int h, int w, int row0) {
  extern __shared__ float tile[];                       // [kVolTileX][D*8 + 1]
  const int y = blockIdx.y, x0 = blockIdx.x * kVolTileX, ld = D * 8 + 1;
  const int nx = min(kVolTileX, w - x0);
  for (int idx = threadIdx.x; idx < 8 * D * kVolTileX; idx += blockDim.x) {
    const int x = idx % kVolTileX, z = (idx / kVolTileX) % D, c = idx / (kVolTileX * D);
    if (x < nx) tile[x * ld + z * 8 + c] = src[(((size_t)c * D + z) * h + y + row0) * w + x0 + x];
  }
  __syncthreads();
  float *out = dst + ((size_t)y * w + x0) * D * 8;
  for (int e = threadIdx.x; e < nx * D * 8; e += blockDim.x) out[e] = tile[(e / (D * 8)) * ld + e % (D * 8)];
}

// (row band: n texels starting at texel `first` of planes that hold hw texels each)
__global__ void pack_nearfar_kernel(const float *__restrict__ dv, float *__restrict__ dst, int D, int hw, int first, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  reinterpret_cast<float2 *>(dst)[i] = make_float2(dv[first + i], dv[(size_t)(D - 1) * hw + first + i]);
}

// dst_f32[(row(k))*ldn + n_off + n] = src[n*K + k]; row(k) = (k >= shift_from ? k+1 : k) + k_off
__global__ void pack_w_f32_kernel(const float *__restrict__ src, int Nout, int K, float *__restrict__ dst, int ldn,
                                  int n_off, int k_off, int shift_from) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nout * K) return;
  const int n = i / K, k = i - n * K;
  const int row = (k >= shift_from ? k + 1 : k) + k_off;
  dst[(size_t)row * ldn + n_off + n] = src[i];
}

// UMMA interleaved K-major tile: element (n,k) -> ((k/8)*Ntile + n)*8 + k%8
__global__ void pack_w_bf16_kernel(const float *__restrict__ src, int Nout, int K, __nv_bfloat16 *__restrict__ dst,
                                   int Ntile, int n_off, int k_off, int shift_from) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nout * K) return;
  const int n = i / K + n_off, k0 = i % K;
  const int k = (k0 >= shift_from ? k0 + 1 : k0) + k_off;
  dst[((size_t)(k >> 3) * Ntile + n) * 8 + (k & 7)] = __float2bfloat16_rn(src[i]);
}

__global__ void pack_vd_kernel(const float *__restrict__ src, float *__restrict__ vd, int n_off) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;          // src [64][155]
  if (i >= 64 * 27) return;
  const int n = i / 27, k = i % 27;
  vd[k * 128 + n_off + n] = src[n * 155 + 128 + k];
}

// split-bf16 unit layout (ucn::X3Layout): element (n,k) -> unit k/16: hi tile, then lo tile at +4096 B
__global__ void pack_w_x3_kernel(const float *__restrict__ src, int Nout, int K, int k_src0, int k_cnt,
                                 unsigned char *__restrict__ dst, int n_off, int k_off) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nout * k_cnt) return;
  const int n = i / k_cnt + n_off, ks = i % k_cnt;
  const float w = src[(size_t)(i / k_cnt) * K + k_src0 + ks];
  const int k = ks + k_off, u = k >> 4, kk = k & 15;
  const __nv_bfloat16 hi = __float2bfloat16_rn(w);
  const __nv_bfloat16 lo = __float2bfloat16_rn(w - __bfloat162float(hi));
  __nv_bfloat16 *t = reinterpret_cast<__nv_bfloat16 *>(dst + (size_t)u * ucn::X3Layout::kUnit);
  const int e = ((kk >> 3) * 128 + n) * 8 + (kk & 7);
  t[e] = hi;
  t[2048 + e] = lo;
}

}  // namespace

extern "C" int ucn_pack_channels_last(const float *src, float *dst, int B, int Cn, int64_t N, int Cpad, void *stream) {
  UCN_CHECK_ARG(src && dst && B > 0 && Cn > 0 && N > 0 && Cpad >= Cn, "ucn_pack_channels_last: bad argument");
  const int64_t total = (int64_t)B * N;
  pack_cl_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src, dst, Cn, N, Cpad, total);
  UCN_LAUNCH_CHECK("pack_cl_kernel");
  return UCN_OK;
}

extern "C" int ucn_pack_volume_rows(const float *src, float *dst, int D, int h, int w, int row0, int rows, void *stream) {
  UCN_CHECK_ARG(src && dst && D > 0 && D <= 256 && h > 0 && w > 0 && row0 >= 0 && rows >= 0 && row0 + rows <= h,
                "ucn_pack_volume: bad argument");
  if (rows == 0) return UCN_OK;
  const dim3 grid((w + kVolTileX - 1) / kVolTileX, rows);
  const size_t smem = (size_t)kVolTileX * (D * 8 + 1) * sizeof(float);
  static int attr[kUcnMaxDevices];
  if (smem > 48 * 1024) {
    cudaError_t e = ucn_ensure_dyn_smem(pack_volume_kernel, (int)smem, attr);
    if (e != cudaSuccess) return ucn_cuda_status(e, "pack_volume smem attribute");
  }
  pack_volume_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(src, dst, D, h, w, row0);
  UCN_LAUNCH_CHECK("pack_volume_kernel");
  return UCN_OK;
}
extern "C" int ucn_pack_volume(const float *src, float *dst, int D, int h, int w, void *stream) {
  return ucn_pack_volume_rows(src, dst, D, h, w, 0, h, stream);
}

extern "C" int ucn_pack_nearfar_rows(const float *depth_values, float *dst, int D, int h, int w, int row0, int rows,
                                     void *stream) {
  UCN_CHECK_ARG(depth_values && dst && D > 0 && h > 0 && w > 0 && row0 >= 0 && rows >= 0 && row0 + rows <= h,
                "ucn_pack_nearfar: bad argument");
  if (rows == 0) return UCN_OK;
  const int n = rows * w;
  pack_nearfar_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(depth_values, dst, D, h * w, row0 * w, n);
  UCN_LAUNCH_CHECK("pack_nearfar_kernel");
  return UCN_OK;
}
extern "C" int ucn_pack_nearfar(const float *depth_values, float *dst, int D, int h, int w, void *stream) {
  return ucn_pack_nearfar_rows(depth_values, dst, D, h, w, 0, h, stream);
}

// Host -> device copy of a row band of `planes` row-major [h][w] fp32 planes (one strided DMA: cudaMemcpy2DAsync with
// one "row" per plane).  src is the HOST tensor (pinned for an asynchronous copy), dst the dense [planes][rows][w] band.
extern "C" int ucn_h2d_rows(float *dst, const float *src_host, int64_t planes, int h, int w, int row0, int rows, void *stream) {
  UCN_CHECK_ARG(dst && src_host && planes > 0 && h > 0 && w > 0 && row0 >= 0 && rows >= 0 && row0 + rows <= h,
                "ucn_h2d_rows: bad argument");
  if (rows == 0) return UCN_OK;
  const size_t width = (size_t)rows * w * sizeof(float);
  cudaError_t e = cudaMemcpy2DAsync(dst, width, src_host + (size_t)row0 * w, (size_t)h * w * sizeof(float), width,
                                    (size_t)planes, cudaMemcpyHostToDevice, (cudaStream_t)stream);
  return e == cudaSuccess ? UCN_OK : ucn_cuda_status(e, "ucn_h2d_rows: cudaMemcpy2DAsync");
}

extern "C" size_t ucn_weights_f32_bytes(int V) { return ucn::F32Layout(V).total * sizeof(float); }
extern "C" size_t ucn_weights_bf16_bytes(int V) { return ucn::BF16Layout(V).total_bytes; }
extern "C" size_t ucn_weights_bf16x3_bytes(int V) { return ucn::X3Layout(V).total_bytes; }

static int pack_weights_x3(const ucn_state_dict *sd, unsigned char *dst, cudaStream_t st) {
  const int V = sd->V;
  const ucn::X3Layout X(V);
  cudaError_t e = cudaMemsetAsync(dst, 0, X.total_bytes, st);
  if (e != cudaSuccess) return ucn_cuda_status(e, "memset x3 pack");
  auto mat = [&](const float *src, int Nout, int K, int k_src0, int k_cnt, size_t off, int n_off, int k_off) {
    const int n = Nout * k_cnt;
    pack_w_x3_kernel<<<(n + 255) / 256, 256, 0, st>>>(src, Nout, K, k_src0, k_cnt, dst + off, n_off, k_off);
    ucn_count_launch();
  };
  float *f32 = reinterpret_cast<float *>(dst + X.f32);
  auto vec = [&](const float *src, int n, int off) {
    cudaMemcpyAsync(f32 + off, src, n * sizeof(float), cudaMemcpyDeviceToDevice, st);
  };
  const int Kf = 24 + 4 * V, Kb = 8 * V;
  mat(sd->film_w, 128, Kf, 0, Kf, X.film, 0, 0);
  mat(sd->film_b, 128, 1, 0, 1, X.film, 0, Kf);                  // bias row (A carries a ones column at k = Kf)
  mat(sd->fb_w, 128, Kb, 0, Kb, X.fb, 0, 0);
  mat(sd->fb_b, 128, 1, 0, 1, X.fb, 0, Kb);
  mat(sd->pts_w[0], 128, 63, 0, 63, X.pts0, 0, 0);
  for (int i = 1; i < 5; ++i) mat(sd->pts_w[i], 128, 128, 0, 128, X.pts14[i - 1], 0, 0);
  mat(sd->pts_w[5], 128, 191, 63, 128, X.pts5h, 0, 0);           // skip layer: input = [PE63 | h]  (models.py:156)
  mat(sd->pts_w[5], 128, 191, 0, 63, X.pts5pe, 0, 0);
  mat(sd->feature_w, 128, 128, 0, 128, X.feat, 0, 0);
  mat(sd->views_w, 64, 155, 0, 128, X.vvh, 0, 0);
  mat(sd->vconf_w, 64, 155, 0, 128, X.vvh, 64, 0);
  for (int i = 0; i < 6; ++i) vec(sd->pts_b[i], 128, X.b_pts + 128 * i);
  vec(sd->feature_b, 128, X.b_feat);
  vec(sd->views_b, 64, X.b_vv);
  vec(sd->vconf_b, 64, X.b_vv + 64);
  // heads and view-direction rows stay fp32 (CUDA-core epilogues): w_base [k][4], w_adapt [k][4], vd [k][128]
  pack_w_f32_kernel<<<2, 256, 0, st>>>(sd->crgb_w, 3, 128, f32 + X.w_base, 4, 0, 0, INT_MAX);
  pack_w_f32_kernel<<<1, 256, 0, st>>>(sd->alpha1_w, 1, 128, f32 + X.w_base, 4, 3, 0, INT_MAX);
  pack_w_f32_kernel<<<1, 256, 0, st>>>(sd->rgb_w, 3, 64, f32 + X.w_adapt, 4, 0, 0, INT_MAX);
  pack_w_f32_kernel<<<1, 256, 0, st>>>(sd->alpha_w, 1, 64, f32 + X.w_adapt, 4, 3, 64, INT_MAX);
  ucn_count_launch(4);
  vec(sd->crgb_b, 3, X.b_base); vec(sd->alpha1_b, 1, X.b_base + 3);
  vec(sd->rgb_b, 3, X.b_adapt); vec(sd->alpha_b, 1, X.b_adapt + 3);
  // vd[k][n] = W[n][128 + k]: rows 128..154 of the K dimension of views_linears.0 / view_confi_linears.0
  pack_vd_kernel<<<(27 * 64 + 255) / 256, 256, 0, st>>>(sd->views_w, f32 + X.vd, 0);
  pack_vd_kernel<<<(27 * 64 + 255) / 256, 256, 0, st>>>(sd->vconf_w, f32 + X.vd, 64);
  ucn_count_launch(2);
  e = cudaGetLastError();
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_pack_weights(x3)");
  // the 4-wide head weights also travel to K3 as a kernel parameter (constant bank): keep a host copy per pack
  float host[1024];
  e = cudaMemcpyAsync(host, f32 + X.w_base, sizeof(host), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_pack_weights(x3): head weights");
  ucn_headw_store(dst, host);
  return UCN_OK;
}

extern "C" int ucn_pack_weights(const ucn_state_dict *sd, float *dst, void *dst_bf16, void *dst_x3, void *stream_) {
  UCN_CHECK_ARG(sd && (dst || dst_bf16 || dst_x3), "ucn_pack_weights: null argument");
  UCN_CHECK_ARG(sd->V >= 1 && sd->V <= UCN_MAX_V, "ucn_pack_weights: 1 <= V <= 8");
  cudaStream_t st = (cudaStream_t)stream_;
  const int V = sd->V;
  const ucn::F32Layout L(V);
  const ucn::BF16Layout B(V);
  __nv_bfloat16 *hb = reinterpret_cast<__nv_bfloat16 *>(dst_bf16);
  float *bias16 = dst_bf16 ? reinterpret_cast<float *>(reinterpret_cast<char *>(dst_bf16) + B.bias) : nullptr;
  cudaError_t e;
  if (dst && (e = cudaMemsetAsync(dst, 0, L.total * sizeof(float), st)) != cudaSuccess) return ucn_cuda_status(e, "memset f32 pack");
  if (hb && (e = cudaMemsetAsync(hb, 0, B.total_bytes, st)) != cudaSuccess) return ucn_cuda_status(e, "memset bf16 pack");

  auto mat = [&](const float *src, int Nout, int K, size_t off32, int ldn, size_t off16, int Ntile, int n_off,
                 int k_off, int shift_from) {
    const int n = Nout * K, blocks = (n + 255) / 256;
    if (dst) { pack_w_f32_kernel<<<blocks, 256, 0, st>>>(src, Nout, K, dst + off32, ldn, n_off, k_off, shift_from); ucn_count_launch(); }
    if (hb) {
      pack_w_bf16_kernel<<<blocks, 256, 0, st>>>(src, Nout, K, reinterpret_cast<__nv_bfloat16 *>(reinterpret_cast<char *>(hb) + off16),
                                                 Ntile, n_off, k_off, shift_from);
      ucn_count_launch();
    }
  };
  auto vec = [&](const float *src, int n, size_t off32, size_t off16) {
    if (dst) cudaMemcpyAsync(dst + off32, src, n * sizeof(float), cudaMemcpyDeviceToDevice, st);
    if (bias16) cudaMemcpyAsync(bias16 + off16, src, n * sizeof(float), cudaMemcpyDeviceToDevice, st);
  };
  const int NS = INT_MAX;
  if (hb) {   // film / fb biases ride in the MMA: bias row at k = K_real of the bf16 tile (ones column in A)
    pack_w_bf16_kernel<<<1, 128, 0, st>>>(sd->film_b, 128, 1, reinterpret_cast<__nv_bfloat16 *>(reinterpret_cast<char *>(hb) + B.film), 128, 0, 24 + 4 * V, NS);
    pack_w_bf16_kernel<<<1, 128, 0, st>>>(sd->fb_b, 128, 1, reinterpret_cast<__nv_bfloat16 *>(reinterpret_cast<char *>(hb) + B.fb), 128, 0, 8 * V, NS);
    ucn_count_launch(2);
  }
  mat(sd->film_w, 128, 24 + 4 * V, L.film, 128, B.film, 128, 0, 0, NS);  vec(sd->film_b, 128, L.film_b, B.b_film);
  mat(sd->fb_w, 128, 8 * V, L.fb, 128, B.fb, 128, 0, 0, NS);              vec(sd->fb_b, 128, L.fb_b, B.b_fb);
  for (int i = 0; i < 6; ++i) {
    const int K = i == 0 ? 63 : (i == 5 ? 191 : 128);
    mat(sd->pts_w[i], 128, K, L.pts[i], 128, B.pts[i], 128, 0, 0, i == 5 ? 63 : NS);
    vec(sd->pts_b[i], 128, L.pts_b[i], B.b_pts[i]);
  }
  mat(sd->feature_w, 128, 128, L.feat, 128, B.feat, 128, 0, 0, NS);       vec(sd->feature_b, 128, L.feat_b, B.b_feat);
  mat(sd->views_w, 64, 155, L.vv, 128, B.vv, 128, 0, 0, NS);              vec(sd->views_b, 64, L.vv_b, B.b_vv);
  mat(sd->vconf_w, 64, 155, L.vv, 128, B.vv, 128, 64, 0, NS);             vec(sd->vconf_b, 64, L.vv_b + 64, B.b_vv + 64);
  mat(sd->crgb_w, 3, 128, L.base, 4, B.base, 16, 0, 0, NS);               vec(sd->crgb_b, 3, L.base_b, B.b_base);
  mat(sd->alpha1_w, 1, 128, L.base, 4, B.base, 16, 3, 0, NS);             vec(sd->alpha1_b, 1, L.base_b + 3, B.b_base + 3);
  mat(sd->rgb_w, 3, 64, L.adapt, 4, B.adapt, 16, 0, 0, NS);               vec(sd->rgb_b, 3, L.adapt_b, B.b_adapt);
  mat(sd->alpha_w, 1, 64, L.adapt, 4, B.adapt, 16, 3, 64, NS);            vec(sd->alpha_b, 1, L.adapt_b + 3, B.b_adapt + 3);
  e = cudaGetLastError();
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_pack_weights");
  if (dst_x3) return pack_weights_x3(sd, reinterpret_cast<unsigned char *>(dst_x3), st);
  return UCN_OK;
}
