// ucn_train.cu -- K3b: fp32 training primitives of the UC-NeRF MLP (forward with saved activations + backward).
//
// replaces (reference): the autograd graph of network/models.py:138-184 BaseAdapt_Renderer.forward as train.py:147-190
// runs it (nn.Linear forward / grad-input / grad-weight, the FiLM multiply + ReLU, h * feats_bias, the
// uncertainty-weighted blend) -- SURVEY.md 8a row a13, config 4.  The layer sequence is driven by
// uc-nerf_b200/train_path.py; every tensor is row-major fp32 with an explicit leading dimension, weights and their
// gradients stay in the reference's own [out][in] layout, so parameter .grad buffers are written in place.
//
// First cut, correctness before speed: one generic SIMT fp32 GEMM (64x64 tile, 4x4 micro-tile, strided operands,
// optional split of the reduction dimension with atomic accumulation) serves the three nn.Linear products.  The
// training batch is 2 048 rays (258 K samples, ~0.2 TFLOP per step), two orders of magnitude below a rendered frame.
#include "ucn_common.cuh"

namespace {

constexpr int BT = 64, PT = 16;     // output tile, reduction step

// C[i][j] (+)= sum_p A[i*sai + p*sap] * B[p*sbp + j*sbj]  (+ bias[j]),  optional ReLU, optional atomic accumulate
enum { G_ACCUM = 1, G_RELU = 2, G_ATOMIC = 4 };
__global__ void __launch_bounds__(256)
gemm_kernel(const float *__restrict__ A, long long sai, long long sap, const float *__restrict__ B, long long sbp,
            long long sbj, float *__restrict__ C, long long ldc, const float *__restrict__ bias, int I, int J,
            long long P, long long p_chunk, int flags) {
  __shared__ float As[PT][BT + 1], Bs[PT][BT + 1];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int i0 = blockIdx.x * BT, j0 = blockIdx.y * BT;
  const long long p_begin = (long long)blockIdx.z * p_chunk, p_end = min(P, p_begin + p_chunk);
  float acc[4][4] = {};
  for (long long p0 = p_begin; p0 < p_end; p0 += PT) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int e = tid + 256 * r;
      int ai, ap, bj, bp;
      if (sap == 1) { ap = e % PT; ai = e / PT; } else { ai = e % BT; ap = e / BT; }
      if (sbp == 1) { bp = e % PT; bj = e / PT; } else { bj = e % BT; bp = e / BT; }
      As[ap][ai] = (i0 + ai < I && p0 + ap < p_end) ? A[(long long)(i0 + ai) * sai + (p0 + ap) * sap] : 0.0f;
      Bs[bp][bj] = (j0 + bj < J && p0 + bp < p_end) ? B[(p0 + bp) * sbp + (long long)(j0 + bj) * sbj] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int p = 0; p < PT; ++p) {
      float a[4], b[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) { a[r] = As[p][ty + 16 * r]; b[r] = Bs[p][tx + 16 * r]; }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(a[r], b[c], acc[r][c]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = i0 + ty + 16 * r;
    if (i >= I) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int j = j0 + tx + 16 * c;
      if (j >= J) continue;
      float v = acc[r][c];
      float *dst = C + (long long)i * ldc + j;
      if (flags & G_ATOMIC) { atomicAdd(dst, v); continue; }
      if (bias) v += bias[j];
      if (flags & G_ACCUM) v += *dst;
      if (flags & G_RELU) v = fmaxf(v, 0.0f);
      *dst = v;
    }
  }
}

// db[n] += sum_m dY[m][n]
__global__ void __launch_bounds__(256)
colsum_kernel(const float *__restrict__ dY, long long ldy, long long M, int Nn, long long m_chunk, float *db) {
  const int n = blockIdx.x * 32 + (threadIdx.x & 31), sub = threadIdx.x >> 5;
  __shared__ float part[8][33];
  float s = 0.f;
  const long long m0 = (long long)blockIdx.y * m_chunk, m1 = min(M, m0 + m_chunk);
  if (n < Nn)
    for (long long m = m0 + sub; m < m1; m += 8) s += dY[m * ldy + n];
  part[sub][threadIdx.x & 31] = s;
  __syncthreads();
  if (sub == 0 && n < Nn) {
    for (int k = 1; k < 8; ++k) s += part[k][threadIdx.x];
    atomicAdd(db + n, s);
  }
}

// [x | sin(2^l x) l<L | cos(2^l x) l<L], each block xyz-contiguous (network/models.py:50-54); same arithmetic as the
// fp32 forward kernel (ucn_mlp_fp32.cu)
__device__ __forceinline__ void embed(const float x[3], int L, float *out) {
  out[0] = x[0]; out[1] = x[1]; out[2] = x[2];
  for (int l = 0; l < L; ++l) {
    const float fq = (float)(1 << l);
    for (int c = 0; c < 3; ++c) {
      float s, co;
      sincosf(x[c] * fq, &s, &co);
      out[3 + 3 * l + c] = s;
      out[3 + 3 * L + 3 * l + c] = co;
    }
  }
}
__global__ void prep_kernel(const float *__restrict__ ndc, const float *__restrict__ angle, int S, long long N,
                            float *__restrict__ ph, long long ldph, float *__restrict__ ftv, long long ldftv) {
  const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= N) return;
  const float x[3] = {ndc[3 * q], ndc[3 * q + 1], ndc[3 * q + 2]};
  const long long r = q / S;
  const float a[3] = {angle[3 * r], angle[3 * r + 1], angle[3 * r + 2]};
  embed(x, 10, ph + q * ldph);              // PE(ndc)  -> ph[:, 0:63]
  embed(a, 4, ftv + q * ldftv + 128);       // PE(dir)  -> ftv[:, 128:155]
}

__global__ void film_relu_fwd_kernel(const float *__restrict__ pre, const float *__restrict__ mod, float *__restrict__ h,
                                     long long ldh, long long M) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M * 128) return;
  h[(e >> 7) * ldh + (e & 127)] = fmaxf(pre[e] * mod[e], 0.0f);
}
__global__ void film_relu_bwd_kernel(const float *__restrict__ dh, long long lddh, const float *__restrict__ pre,
                                     const float *__restrict__ mod, const float *__restrict__ h, long long ldh,
                                     float *__restrict__ dpre, float *__restrict__ dmod, int accumulate, long long M) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M * 128) return;
  const long long m = e >> 7;
  const int c = (int)(e & 127);
  const float g = h[m * ldh + c] > 0.0f ? dh[m * lddh + c] : 0.0f;
  dpre[e] = g * mod[e];
  const float dm = g * pre[e];
  dmod[e] = accumulate ? dmod[e] + dm : dm;
}
__global__ void mul_fwd_kernel(const float *__restrict__ a, long long lda, const float *__restrict__ b,
                               float *__restrict__ o, long long M) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M * 128) return;
  o[e] = a[(e >> 7) * lda + (e & 127)] * b[e];
}
// o = a * b:  da (+)= do * b,  db = do * a
__global__ void mul_bwd_kernel(const float *__restrict__ d_o, const float *__restrict__ a, long long lda,
                               const float *__restrict__ b, float *__restrict__ da, int accumulate,
                               float *__restrict__ db, long long M) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M * 128) return;
  const float g = d_o[e];
  const float v = g * b[e];
  da[e] = accumulate ? da[e] + v : v;
  db[e] = g * a[(e >> 7) * lda + (e & 127)];
}
__global__ void relu_bwd_kernel(const float *__restrict__ dy, const float *__restrict__ y, float *__restrict__ dx,
                                long long n) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e < n) dx[e] = y[e] > 0.0f ? dy[e] : 0.0f;
}
// network/models.py:149,177-180
__global__ void blend_fwd_kernel(const float4 *__restrict__ base, const float4 *__restrict__ adapt,
                                 const float *__restrict__ conf, long long ldconf, float4 *__restrict__ raw, long long M) {
  const long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  const float u = 1.0f - conf[m * ldconf], cf = 1.0f - u;
  const float4 b = base[m], a = adapt[m];
  float4 o;
  o.x = 1.0f / (1.0f + expf(-(b.x * cf + a.x * u)));
  o.y = 1.0f / (1.0f + expf(-(b.y * cf + a.y * u)));
  o.z = 1.0f / (1.0f + expf(-(b.z * cf + a.z * u)));
  o.w = fmaxf(a.w * cf + b.w * u, 0.0f);
  raw[m] = o;
}
__global__ void blend_bwd_kernel(const float4 *__restrict__ d_raw, const float4 *__restrict__ raw,
                                 const float4 *__restrict__ base, const float4 *__restrict__ adapt,
                                 const float *__restrict__ conf, long long ldconf, float4 *__restrict__ d_base,
                                 float4 *__restrict__ d_adapt, float *__restrict__ d_conf, long long lddconf, long long M) {
  const long long m = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  const float u = 1.0f - conf[m * ldconf], cf = 1.0f - u;
  const float4 g = d_raw[m], o = raw[m], b = base[m], a = adapt[m];
  const float dx = g.x * o.x * (1.0f - o.x), dy = g.y * o.y * (1.0f - o.y), dz = g.z * o.z * (1.0f - o.z);
  const float dw = o.w > 0.0f ? g.w : 0.0f;
  d_base[m] = make_float4(dx * cf, dy * cf, dz * cf, dw * u);
  d_adapt[m] = make_float4(dx * u, dy * u, dz * u, dw * cf);
  const float du = dx * (a.x - b.x) + dy * (a.y - b.y) + dz * (a.z - b.z) + dw * (b.w - a.w);
  d_conf[m * lddconf] = -du;                 // u = 1 - confidence
}

inline unsigned nblk(long long n, int t) { return (unsigned)((n + t - 1) / t); }

int launch_gemm(const float *A, long long sai, long long sap, const float *B, long long sbp, long long sbj, float *C,
                long long ldc, const float *bias, long long I, long long J, long long P, int flags, int splits,
                void *stream) {
  if (I <= 0 || J <= 0) return UCN_OK;
  long long p_chunk = P;
  if (splits > 1) {
    p_chunk = ((P + splits - 1) / splits + PT - 1) / PT * PT;
    splits = (int)((P + p_chunk - 1) / p_chunk);
  } else {
    splits = 1;
  }
  dim3 grid(nblk(I, BT), nblk(J, BT), (unsigned)splits);
  UCN_CHECK_ARG(grid.y < 65536u, "ucn_linear: output too wide");
  gemm_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, sai, sap, B, sbp, sbj, C, ldc, bias, (int)I, (int)J, P, p_chunk, flags);
  UCN_LAUNCH_CHECK("gemm_kernel");
  return UCN_OK;
}

}  // namespace

// Y[m][n] (+)= sum_k X[m*ldx + k] W[n*ldw + k] + b[n]     (nn.Linear forward; b may be NULL; flags: 1 accumulate, 2 ReLU)
extern "C" int ucn_linear_fwd(const float *X, int64_t ldx, const float *W, int64_t ldw, const float *b, int64_t M, int K,
                              int N, float *Y, int64_t ldy, int flags, void *stream) {
  UCN_CHECK_ARG(X && W && Y, "ucn_linear_fwd: null argument");
  UCN_CHECK_ARG((flags & ~3) == 0, "ucn_linear_fwd: flags");
  return launch_gemm(X, ldx, 1, W, 1, ldw, Y, ldy, b, M, N, K, flags, 1, stream);
}
// dX[m][k] (+)= sum_n dY[m*ldy + n] W[n*ldw + k]           (nn.Linear grad-input; flags: 1 accumulate)
extern "C" int ucn_linear_bwd_input(const float *dY, int64_t ldy, const float *W, int64_t ldw, int64_t M, int K, int N,
                                    float *dX, int64_t ldx, int flags, void *stream) {
  UCN_CHECK_ARG(dY && W && dX, "ucn_linear_bwd_input: null argument");
  UCN_CHECK_ARG((flags & ~1) == 0, "ucn_linear_bwd_input: flags");
  return launch_gemm(dY, ldy, 1, W, ldw, 1, dX, ldx, nullptr, M, K, N, flags, 1, stream);
}
// dW[n*lddw + k] += sum_m dY[m*ldy + n] X[m*ldx + k],  db[n] += sum_m dY[m*ldy + n]   (grad-weight; db may be NULL)
extern "C" int ucn_linear_bwd_weight(const float *dY, int64_t ldy, const float *X, int64_t ldx, int64_t M, int K, int N,
                                     float *dW, int64_t lddw, float *db, void *stream) {
  UCN_CHECK_ARG(dY && X && dW, "ucn_linear_bwd_weight: null argument");
  if (M <= 0) return UCN_OK;
  const long long tiles = (long long)((N + BT - 1) / BT) * ((K + BT - 1) / BT);
  int splits = (int)((2 * 148 + tiles - 1) / tiles);
  const long long max_splits = (M + 4 * PT - 1) / (4 * PT);
  if (splits > max_splits) splits = (int)max_splits;
  const int rc = launch_gemm(dY, 1, ldy, X, ldx, 1, dW, lddw, nullptr, N, K, M, G_ATOMIC, splits < 1 ? 1 : splits, stream);
  if (rc != UCN_OK || !db) return rc;
  const long long chunks = (M + 4095) / 4096;
  colsum_kernel<<<dim3(nblk(N, 32), (unsigned)chunks), 256, 0, (cudaStream_t)stream>>>(dY, ldy, M, N, 4096, db);
  UCN_LAUNCH_CHECK("colsum_kernel");
  return UCN_OK;
}
// ph[:, 0:63] = PE_10(ndc), ftv[:, 128:155] = PE_4(angle[ray])      (network/models.py:50-54; renderer.py:88-100)
extern "C" int ucn_train_prep(const float *ndc, const float *angle, int S, int64_t N, float *ph, int64_t ldph, float *ftv,
                              int64_t ldftv, void *stream) {
  UCN_CHECK_ARG(ndc && angle && ph && ftv && S > 0, "ucn_train_prep: bad argument");
  UCN_CHECK_ARG(ldph >= 63 && ldftv >= 155, "ucn_train_prep: leading dimensions");
  if (N <= 0) return UCN_OK;
  prep_kernel<<<nblk(N, 128), 128, 0, (cudaStream_t)stream>>>(ndc, angle, S, N, ph, ldph, ftv, ldftv);
  UCN_LAUNCH_CHECK("prep_kernel");
  return UCN_OK;
}
// h = relu(pre * mod)                                              (network/models.py:154-155)
extern "C" int ucn_film_relu_fwd(const float *pre, const float *mod, float *h, int64_t ldh, int64_t M, void *stream) {
  UCN_CHECK_ARG(pre && mod && h, "ucn_film_relu_fwd: null argument");
  if (M <= 0) return UCN_OK;
  film_relu_fwd_kernel<<<nblk(M * 128, 256), 256, 0, (cudaStream_t)stream>>>(pre, mod, h, ldh, M);
  UCN_LAUNCH_CHECK("film_relu_fwd_kernel");
  return UCN_OK;
}
extern "C" int ucn_film_relu_bwd(const float *dh, int64_t lddh, const float *pre, const float *mod, const float *h,
                                 int64_t ldh, float *dpre, float *dmod, int accumulate, int64_t M, void *stream) {
  UCN_CHECK_ARG(dh && pre && mod && h && dpre && dmod, "ucn_film_relu_bwd: null argument");
  if (M <= 0) return UCN_OK;
  film_relu_bwd_kernel<<<nblk(M * 128, 256), 256, 0, (cudaStream_t)stream>>>(dh, lddh, pre, mod, h, ldh, dpre, dmod, accumulate, M);
  UCN_LAUNCH_CHECK("film_relu_bwd_kernel");
  return UCN_OK;
}
// o = a * b  (h * feats_bias, network/models.py:164) and its backward
extern "C" int ucn_mul_fwd(const float *a, int64_t lda, const float *b, float *o, int64_t M, void *stream) {
  UCN_CHECK_ARG(a && b && o, "ucn_mul_fwd: null argument");
  if (M <= 0) return UCN_OK;
  mul_fwd_kernel<<<nblk(M * 128, 256), 256, 0, (cudaStream_t)stream>>>(a, lda, b, o, M);
  UCN_LAUNCH_CHECK("mul_fwd_kernel");
  return UCN_OK;
}
extern "C" int ucn_mul_bwd(const float *d_o, const float *a, int64_t lda, const float *b, float *da, int accumulate,
                           float *db, int64_t M, void *stream) {
  UCN_CHECK_ARG(d_o && a && b && da && db, "ucn_mul_bwd: null argument");
  if (M <= 0) return UCN_OK;
  mul_bwd_kernel<<<nblk(M * 128, 256), 256, 0, (cudaStream_t)stream>>>(d_o, a, lda, b, da, accumulate, db, M);
  UCN_LAUNCH_CHECK("mul_bwd_kernel");
  return UCN_OK;
}
extern "C" int ucn_relu_bwd(const float *dy, const float *y, float *dx, int64_t n, void *stream) {
  UCN_CHECK_ARG(dy && y && dx, "ucn_relu_bwd: null argument");
  if (n <= 0) return UCN_OK;
  relu_bwd_kernel<<<nblk(n, 256), 256, 0, (cudaStream_t)stream>>>(dy, y, dx, n);
  UCN_LAUNCH_CHECK("relu_bwd_kernel");
  return UCN_OK;
}
// raw = [sigmoid(base_rgb (1-u) + adapt_rgb u), relu(adapt_s (1-u) + base_s u)],  u = 1 - conf   (models.py:149,177-180)
extern "C" int ucn_blend_fwd(const float *base4, const float *adapt4, const float *conf, int64_t ldconf, float *raw,
                             int64_t M, void *stream) {
  UCN_CHECK_ARG(base4 && adapt4 && conf && raw, "ucn_blend_fwd: null argument");
  if (M <= 0) return UCN_OK;
  blend_fwd_kernel<<<nblk(M, 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4 *>(base4), reinterpret_cast<const float4 *>(adapt4), conf, ldconf,
      reinterpret_cast<float4 *>(raw), M);
  UCN_LAUNCH_CHECK("blend_fwd_kernel");
  return UCN_OK;
}
extern "C" int ucn_blend_bwd(const float *d_raw, const float *raw, const float *base4, const float *adapt4,
                             const float *conf, int64_t ldconf, float *d_base4, float *d_adapt4, float *d_conf,
                             int64_t lddconf, int64_t M, void *stream) {
  UCN_CHECK_ARG(d_raw && raw && base4 && adapt4 && conf && d_base4 && d_adapt4 && d_conf, "ucn_blend_bwd: null argument");
  if (M <= 0) return UCN_OK;
  blend_bwd_kernel<<<nblk(M, 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4 *>(d_raw), reinterpret_cast<const float4 *>(raw),
      reinterpret_cast<const float4 *>(base4), reinterpret_cast<const float4 *>(adapt4), conf, ldconf,
      reinterpret_cast<float4 *>(d_base4), reinterpret_cast<float4 *>(d_adapt4), d_conf, lddconf, M);
  UCN_LAUNCH_CHECK("blend_bwd_kernel");
  return UCN_OK;
}
