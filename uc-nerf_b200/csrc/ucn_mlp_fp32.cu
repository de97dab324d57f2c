// ucn_mlp_fp32.cu -- K3 (parity mode): positional encoding + uncertainty-conditioned MLP in fp32 SIMT.
//
// replaces (reference): network/models.py:50-54 Embedder.embed, network/renderer.py:78-106
//   run_network_mvs (+batchify), network/models.py:138-184 BaseAdapt_Renderer.forward.
//
// This is the 1e-4 gate path (same arithmetic class as the reference: fp32 FMA, TF32 off).  The
// production path is the tcgen05 kernel in ucn_mlp_tc.cu.  One 256-thread block owns a tile of 64
// samples; activations stay in shared memory, each thread keeps an 8x4 output micro-tile (and its FiLM
// modulation / feature-bias values) in registers across all layers; weights ([K][128], L2-resident)
// stream through the read-only path.
#include "ucn_common.cuh"
#include "ucn_weights.cuh"

namespace {

constexpr int TM = 64;          // samples per tile
constexpr int LDA = 192;        // act : [pe 0..62 | 0 | h 64..191]
constexpr int LDB = 160;        // act2: [hf / feature 0..127 | pe27 128..154 | 0 ...]

struct Smem {
  float act[TM * LDA];
  float act2[TM * LDB];
  float base[TM * 4];
  float u[TM];
};

// acc[i][c] += sum_k A[(ty*8+i)*lda + k] * Wt[k*128 + tx*4 + c]
template <int K>
__device__ __forceinline__ void gemm_tile(float (&acc)[8][4], const float *__restrict__ A, int lda,
                                          const float *__restrict__ Wt, int tx, int ty) {
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    float4 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4 *>(A + (ty * 8 + i) * lda + k);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float4 w = ucn::ldg4(Wt + (size_t)(k + kk) * 128 + tx * 4);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float av = kk == 0 ? a[i].x : (kk == 1 ? a[i].y : (kk == 2 ? a[i].z : a[i].w));
        acc[i][0] = fmaf(av, w.x, acc[i][0]); acc[i][1] = fmaf(av, w.y, acc[i][1]);
        acc[i][2] = fmaf(av, w.z, acc[i][2]); acc[i][3] = fmaf(av, w.w, acc[i][3]);
      }
    }
  }
}

__device__ __forceinline__ void gemm_dyn(float (&acc)[8][4], const float *__restrict__ A, int lda,
                                         const float *__restrict__ Wt, int K, int tx, int ty) {
  for (int k = 0; k < K; k += 4) {
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float4 w = ucn::ldg4(Wt + (size_t)(k + kk) * 128 + tx * 4);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float av = A[(ty * 8 + i) * lda + k + kk];
        acc[i][0] = fmaf(av, w.x, acc[i][0]); acc[i][1] = fmaf(av, w.y, acc[i][1]);
        acc[i][2] = fmaf(av, w.z, acc[i][2]); acc[i][3] = fmaf(av, w.w, acc[i][3]);
      }
    }
  }
}

__device__ __forceinline__ void init_bias(float (&acc)[8][4], const float *__restrict__ b, int tx) {
  const float4 bv = ucn::ldg4(b + tx * 4);
#pragma unroll
  for (int i = 0; i < 8; ++i) { acc[i][0] = bv.x; acc[i][1] = bv.y; acc[i][2] = bv.z; acc[i][3] = bv.w; }
}

// [x | sin(2^l x) l<L | cos(2^l x) l<L], each block xyz-contiguous (network/models.py:50-54)
__device__ __forceinline__ void embed(const float x[3], int L, float *out) {
  out[0] = x[0]; out[1] = x[1]; out[2] = x[2];
  for (int l = 0; l < L; ++l) {
    const float fq = (float)(1 << l);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float s, co;
      sincosf(x[c] * fq, &s, &co);
      out[3 + 3 * l + c] = s;
      out[3 + 3 * L + 3 * l + c] = co;
    }
  }
}

__global__ void __launch_bounds__(256, 1)
mlp_fp32_kernel(const float *__restrict__ W, int V, const float *__restrict__ ndc_pts, const float *__restrict__ feat,
                const float *__restrict__ angle, int64_t N, int S, float *__restrict__ raw) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem &sm = *reinterpret_cast<Smem *>(smem_raw);
  const ucn::F32Layout L(V);
  const int F = 24 + 12 * V + 1, LDF = (F + 3) & ~3;
  float *featS = reinterpret_cast<float *>(smem_raw + sizeof(Smem));       // [TM][LDF]
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;

  for (int64_t q0 = (int64_t)blockIdx.x * TM; q0 < N; q0 += (int64_t)gridDim.x * TM) {
    const int rows = (int)min((int64_t)TM, N - q0);
    // ---- stage inputs ------------------------------------------------------------------------
    for (int e = tid; e < TM * LDF; e += 256) {
      const int i = e / LDF, c = e - i * LDF;
      featS[e] = (i < rows && c < F) ? feat[(q0 + i) * F + c] : 0.0f;
    }
    if (tid < TM) {
      const int i = tid;
      float x[3] = {0.f, 0.f, 0.f}, a[3] = {0.f, 0.f, 0.f};
      if (i < rows) {
        const int64_t q = q0 + i;
        x[0] = ndc_pts[3 * q]; x[1] = ndc_pts[3 * q + 1]; x[2] = ndc_pts[3 * q + 2];
        const int64_t r = q / S;
        a[0] = angle[3 * r]; a[1] = angle[3 * r + 1]; a[2] = angle[3 * r + 2];
      }
      embed(x, 10, sm.act + i * LDA);
      sm.act[i * LDA + 63] = 0.0f;
      embed(a, 4, sm.act2 + i * LDB + 128);
      for (int c = 155; c < LDB; ++c) sm.act2[i * LDB + c] = 0.0f;
    }
    __syncthreads();
    if (tid < TM) sm.u[tid] = 1.0f - featS[tid * LDF + F - 1];            // models.py:149

    // ---- FiLM modulation and feature bias (models.py:150-151), kept in registers -----------------
    float mod[8][4], fbv[8][4], acc[8][4];
    init_bias(mod, W + L.film_b, tx);
    gemm_dyn(mod, featS, LDF, W + L.film, L.Kfilm, tx, ty);
    init_bias(fbv, W + L.fb_b, tx);
    gemm_dyn(fbv, featS + L.Kfilm, LDF, W + L.fb, L.Kfb, tx, ty);

    // ---- 6 modulated layers, skip concat after layer 4 (models.py:153-157) -------------------------
#pragma unroll 1
    for (int l = 0; l < 6; ++l) {
      init_bias(acc, W + L.pts_b[l], tx);
      if (l == 0) gemm_tile<64>(acc, sm.act, LDA, W + L.pts[0], tx, ty);
      else if (l == 5) gemm_tile<192>(acc, sm.act, LDA, W + L.pts[5], tx, ty);
      else gemm_tile<128>(acc, sm.act + 64, LDA, W + L.pts[l], tx, ty);
      __syncthreads();                                     // everyone finished reading h
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float4 h;
        h.x = fmaxf(acc[i][0] * mod[i][0], 0.0f); h.y = fmaxf(acc[i][1] * mod[i][1], 0.0f);
        h.z = fmaxf(acc[i][2] * mod[i][2], 0.0f); h.w = fmaxf(acc[i][3] * mod[i][3], 0.0f);
        *reinterpret_cast<float4 *>(sm.act + (ty * 8 + i) * LDA + 64 + tx * 4) = h;
        if (l == 5) {                                      // h * feats_bias, models.py:164
          float4 g = make_float4(h.x * fbv[i][0], h.y * fbv[i][1], h.z * fbv[i][2], h.w * fbv[i][3]);
          *reinterpret_cast<float4 *>(sm.act2 + (ty * 8 + i) * LDB + tx * 4) = g;
        }
      }
      __syncthreads();
    }
    // ---- base heads: confi_rgb_linear | alpha_linear_1 (models.py:161-162) ------------------------
    {
      const int i = tid >> 2, c = tid & 3;
      float a = W[L.base_b + c];
      const float *h = sm.act + i * LDA + 64;
      for (int k = 0; k < 128; ++k) a = fmaf(h[k], W[L.base + k * 4 + c], a);
      sm.base[i * 4 + c] = a;
    }
    // ---- feature_linear (models.py:164) ---------------------------------------------------------
    init_bias(acc, W + L.feat_b, tx);
    gemm_tile<128>(acc, sm.act2, LDB, W + L.feat, tx, ty);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i)
      *reinterpret_cast<float4 *>(sm.act2 + (ty * 8 + i) * LDB + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    __syncthreads();
    // ---- views_linears | view_confi_linears on [feature, PE(dir)] (models.py:166-173) -----------
    init_bias(acc, W + L.vv_b, tx);
    gemm_tile<156>(acc, sm.act2, LDB, W + L.vv, tx, ty);
#pragma unroll
    for (int i = 0; i < 8; ++i)
      *reinterpret_cast<float4 *>(sm.act + (ty * 8 + i) * LDA + 64 + tx * 4) =
          make_float4(fmaxf(acc[i][0], 0.f), fmaxf(acc[i][1], 0.f), fmaxf(acc[i][2], 0.f), fmaxf(acc[i][3], 0.f));
    __syncthreads();
    // ---- adapt heads + uncertainty-aware blend (models.py:170,174,177-180) -----------------------
    {
      const int i = tid >> 2, c = tid & 3;
      float a = W[L.adapt_b + c];
      const float *h = sm.act + i * LDA + 64;
      for (int k = 0; k < 128; ++k) a = fmaf(h[k], W[L.adapt + k * 4 + c], a);
      const float un = sm.u[i], cf = 1.0f - un, b = sm.base[i * 4 + c];
      float o;
      if (c < 3) {
        const float x = b * cf + a * un;
        o = 1.0f / (1.0f + expf(-x));
      } else {
        o = fmaxf(a * cf + b * un, 0.0f);
      }
      if (i < rows) raw[(q0 + i) * 4 + c] = o;
    }
    __syncthreads();
  }
}

}  // namespace

int ucn_mlp_forward_fp32(const ucn_weights *w, const float *ndc_pts, const float *feat, const float *angle, int R,
                         int S, float *raw, void *stream) {
  UCN_CHECK_ARG(w->f32, "ucn_mlp_forward: fp32 weight pack missing");
  const int V = w->V, F = 24 + 12 * V + 1, LDF = (F + 3) & ~3;
  const int64_t N = (int64_t)R * S;
  const size_t smem = sizeof(Smem) + (size_t)TM * LDF * sizeof(float);
  static int attr[kUcnMaxDevices];
  {
    cudaError_t e = ucn_ensure_dyn_smem(mlp_fp32_kernel, 200 * 1024, attr);
    if (e != cudaSuccess) return ucn_cuda_status(e, "mlp_fp32 smem attribute");
  }
  const int sms = ucn_sm_count();
  const int64_t tiles = (N + TM - 1) / TM;
  const unsigned blocks = (unsigned)(tiles < sms ? tiles : sms);
  mlp_fp32_kernel<<<blocks, 256, smem, (cudaStream_t)stream>>>(w->f32, V, ndc_pts, feat, angle, N, S, raw);
  UCN_LAUNCH_CHECK("mlp_fp32_kernel");
  return UCN_OK;
}
