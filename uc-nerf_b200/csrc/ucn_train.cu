// ucn_train.cu -- K3b: fp32 training primitives of the UC-NeRF MLP (forward with saved activations + backward).
//
// replaces (reference): the autograd graph of network/models.py:138-184 BaseAdapt_Renderer.forward as train.py:147-190
// runs it (nn.Linear forward / grad-input / grad-weight, the FiLM multiply + ReLU, h * feats_bias, the
// uncertainty-weighted blend) -- SURVEY.md 8a row a13, config 4.  The layer sequence is driven by
// uc-nerf_b200/train_path.py; every tensor is row-major fp32 with an explicit leading dimension, weights and their
// gradients stay in the reference's own [out][in] layout, so parameter .grad buffers are written in place.
//
// The three nn.Linear products (forward, grad-input, grad-weight) are one strided GEMM  C[i][j] = sum_p A(i,p) B(p,j)
// with two implementations: gemm_tc_kernel (tcgen05, split-bf16 operands converted on the fly, fp32 accumulation in
// TMEM) for every product that fills a 128-wide tile, and gemm_kernel (fp32 CUDA cores) for the 3- and 1-wide heads.
// The training batch is 2 048 rays (258 K samples, ~0.2 TFLOP per step), two orders of magnitude below a rendered frame.
#include <mutex>
#include <vector>

#include "ucn_common.cuh"
#include "ucn_tc.cuh"

using namespace ucn_tc;

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


// ---- the same product on the tensor cores ---------------------------------------------------------------------------
// C tile 128 x 128 per CTA, fp32 operands converted on the fly to THREE bf16 parts (a = a1 + a2 + a3, 24 significant
// bits) and multiplied with the six tcgen05.mma whose terms are >= 2^-24 relative (a1b1, a1b2, a2b1, a1b3, a2b2, a3b1),
// fp32 accumulation in TMEM: fp32-class products on the tensor cores.  The two-part split of the inference kernel
// (~16 bits, 5e-6 per product) is not enough here: its 1.7e-4 forward error flips ReLU masks of near-zero units and a
// flipped mask changes a gradient sum by a whole term (measured 2.6e-3 of a weight gradient on 3 072 samples).  The
// tensor pipe has the headroom: this kernel is bound by operand staging, not by MMA issue.  The 256 threads
// stage a [128 x 32] slab of each operand per iteration straight from global memory (any strides: the three nn.Linear
// products differ only in which index is contiguous) into the UMMA no-swizzle K-major core-matrix layout
// (element (row, k) at (k/8)*2048 + row*16 + (k%8)*2 bytes), two stages deep: the six MMAs of slab t run while slab
// t+1 is being converted.  Epilogue: TMEM -> registers -> bias / accumulate / ReLU -> global, or fp32 atomics when the
// reduction dimension is split across CTAs (grad-weight).
constexpr int TC_SLAB = 32;                       // reduction elements per stage (two K=16 MMA steps)
constexpr uint32_t TC_CHUNK = 2048;               // one K-chunk (8 bf16) x 128 rows
constexpr int TC_PARTS = 3;                       // bf16 parts per fp32 operand (8 + 8 + 8 significant bits)
constexpr uint32_t TC_PART = 4 * TC_CHUNK;        // one part of one operand slab: 4 K-chunks = 8 KB
constexpr uint32_t TC_OPER = TC_PARTS * TC_PART;  // all parts of one operand slab = 24 KB
constexpr uint32_t TC_STAGE = 2 * TC_OPER;        // A + B
constexpr uint32_t TC_SMEM = 2 * TC_STAGE + 64;

// (a, b) -> three packed bf16 pairs: rounded values, rounded first residuals, rounded second residuals
__device__ __forceinline__ void split3(float a, float b, uint32_t (&w)[3]) {
  w[0] = pack2(a, b);
  const float ra = a - bf_lo(w[0]), rb = b - bf_hi(w[0]);
  w[1] = pack2(ra, rb);
  w[2] = pack2(ra - bf_lo(w[1]), rb - bf_hi(w[1]));
}

__global__ void __launch_bounds__(256, 2)
gemm_tc_kernel(const float *__restrict__ A, long long sai, long long sap, const float *__restrict__ B, long long sbp,
               long long sbj, float *__restrict__ C, long long ldc, const float *__restrict__ bias, int I, int J,
               long long P, long long p_chunk, int flags, int a_vec, int b_vec, const unsigned char *__restrict__ bsplit) {
  extern __shared__ __align__(1024) unsigned char tc_smem[];
  const uint32_t sbase = smem_u32(tc_smem);
  const uint32_t bar_done = sbase + 2 * TC_STAGE, tmem_slot = bar_done + 16, bar_b = bar_done + 32;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int j0 = blockIdx.y * 128, n_tile_i = (I + 127) / 128;
  const long long p_begin = (long long)blockIdx.z * p_chunk, p_end = min(P, p_begin + p_chunk);
  if (tid == 0) {
    mbar_init(bar_done, 1); mbar_init(bar_done + 8, 1);
    mbar_init(bar_b, 1); mbar_init(bar_b + 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(tc_smem + 2 * TC_STAGE + 16);
  const uint32_t idesc = umma_idesc_bf16(128);
  const int row = tid & 127, half = tid >> 7;                 // this thread stages k = 16*half .. 16*half+15 of its row
  const bool b_ok = j0 + row < J;
  const int n_iter = (int)((p_end - p_begin + TC_SLAB - 1) / TC_SLAB);
  // Persistent over the row tiles of the output (grid.x CTAs stride over them): TMEM allocation, barrier set-up and
  // CTA launch are paid once per CTA instead of once per tile (an empty tile pass of the one-tile-per-CTA form cost
  // 41 of the 145 us of an M = 258 048, K = 128 forward product).  `git` counts slabs across tiles: the two stages
  // and their barriers keep alternating through tile boundaries.
  // Operand slabs [128 rows x 32 p].  Two staging patterns per operand:
  //   vec    (p contiguous, 16-byte aligned rows): lanes sweep p, one float4 per lane -> four full 128-byte lines per
  //          warp instruction;
  //   scalar (rows contiguous, or unaligned): thread = row, 16 p values each -- coalesced when the row index is the
  //          contiguous one (grad-weight operands, grad-input weights), L1-cached otherwise (small weight matrices).
  // The loads of the NEXT slab (the next tile's first slab at a tile boundary) are issued as soon as the current slab has
  // been converted out of the same registers: they are in flight across the barrier, the MMA issue, the epilogue and
  // the wait for the stage.
  float va[16], vb[16];
  float4 fa[4], fb[4];
  auto load_slab = [&](int ti, int it) {
    const int i0 = ti * 128;
    const bool a_ok = i0 + row < I;
    const long long ps = p_begin + (long long)it * TC_SLAB, p0 = ps + 16 * half;
    if (a_vec) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int f = tid + 256 * q, r = f >> 3, c4 = f & 7;
        const long long pp = ps + 4 * c4;
        fa[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i0 + r < I && pp < p_end) {
          fa[q] = *reinterpret_cast<const float4 *>(A + (long long)(i0 + r) * sai + pp);
          if (pp + 1 >= p_end) fa[q].y = 0.f;
          if (pp + 2 >= p_end) fa[q].z = 0.f;
          if (pp + 3 >= p_end) fa[q].w = 0.f;
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 16; ++k) va[k] = (a_ok && p0 + k < p_end) ? A[(long long)(i0 + row) * sai + (p0 + k) * sap] : 0.0f;
    }
    if (bsplit) {
      // B (the layer's weight matrix) was split once for the whole launch (split_b_kernel): its slab is a plain copy
    } else if (b_vec) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int f = tid + 256 * q, r = f >> 3, c4 = f & 7;
        const long long pp = ps + 4 * c4;
        fb[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j0 + r < J && pp < p_end) {
          fb[q] = *reinterpret_cast<const float4 *>(B + (long long)(j0 + r) * sbj + pp);
          if (pp + 1 >= p_end) fb[q].y = 0.f;
          if (pp + 2 >= p_end) fb[q].z = 0.f;
          if (pp + 3 >= p_end) fb[q].w = 0.f;
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 16; ++k) vb[k] = (b_ok && p0 + k < p_end) ? B[(p0 + k) * sbp + (long long)(j0 + row) * sbj] : 0.0f;
    }
  };
  uint32_t git = 0;
  if ((int)blockIdx.x < n_tile_i) load_slab(blockIdx.x, 0);
  for (int ti = blockIdx.x; ti < n_tile_i; ti += gridDim.x) {
  const int i0 = ti * 128;
  for (int it = 0; it < n_iter; ++it, ++git) {
    const int st = git & 1;
    if (git >= 2) mbar_wait(bar_done + 8 * st, ((git >> 1) - 1) & 1);     // MMAs of slab git-2 have read this stage
    const uint32_t stage = sbase + st * TC_STAGE;
    if (bsplit && tid == 0) {
      mbar_expect_tx(bar_b + 8 * st, TC_OPER);
      bulk_g2s(stage + TC_OPER, bsplit + ((size_t)blockIdx.y * n_iter + it) * TC_OPER, TC_OPER, bar_b + 8 * st);
    }
#pragma unroll
    for (int op = 0; op < 2; ++op) {
      if (op == 1 && bsplit) break;
      const uint32_t base = stage + op * TC_OPER;
      if (op == 0 ? a_vec : b_vec) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int f = tid + 256 * q, r = f >> 3, c4 = f & 7;
          const float4 v = op == 0 ? fa[q] : fb[q];
          uint32_t w0[3], w1[3];
          split3(v.x, v.y, w0);
          split3(v.z, v.w, w1);
          const uint32_t addr = base + (c4 >> 1) * TC_CHUNK + r * 16 + (c4 & 1) * 8;
#pragma unroll
          for (int pt = 0; pt < TC_PARTS; ++pt)
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr + pt * TC_PART), "r"(w0[pt]), "r"(w1[pt]) : "memory");
        }
      } else {
        const float *v = op == 0 ? va : vb;
        const uint32_t sa = base + row * 16 + (2 * half) * TC_CHUNK;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t w[4][3];
#pragma unroll
          for (int q = 0; q < 4; ++q) split3(v[8 * c + 2 * q], v[8 * c + 2 * q + 1], w[q]);
#pragma unroll
          for (int pt = 0; pt < TC_PARTS; ++pt) sts128(sa + c * TC_CHUNK + pt * TC_PART, w[0][pt], w[1][pt], w[2][pt], w[3][pt]);
        }
      }
    }
    {
      int nti = ti, nit = it + 1;
      if (nit == n_iter) { nti = ti + (int)gridDim.x; nit = 0; }
      if (nti < n_tile_i) load_slab(nti, nit);
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      if (bsplit) mbar_wait(bar_b + 8 * st, (git >> 1) & 1);
      tc_fence_after();
      const uint32_t a0 = sbase + st * TC_STAGE, b0 = a0 + TC_OPER;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        uint64_t da[TC_PARTS], db[TC_PARTS];
#pragma unroll
        for (int pt = 0; pt < TC_PARTS; ++pt) {
          da[pt] = umma_desc(a0 + pt * TC_PART + u * 2 * TC_CHUNK, TC_CHUNK, 128);
          db[pt] = umma_desc(b0 + pt * TC_PART + u * 2 * TC_CHUNK, TC_CHUNK, 128);
        }
        umma_ss(tmem_base, da[0], db[0], idesc, (it > 0 || u > 0) ? 1u : 0u);
        umma_ss(tmem_base, da[0], db[1], idesc, 1u);
        umma_ss(tmem_base, da[1], db[0], idesc, 1u);
        umma_ss(tmem_base, da[0], db[2], idesc, 1u);
        umma_ss(tmem_base, da[1], db[1], idesc, 1u);
        umma_ss(tmem_base, da[2], db[0], idesc, 1u);
      }
      umma_commit(bar_done + 8 * st);
    }
  }
  {
    const uint32_t last = git - 1;
    mbar_wait(bar_done + 8 * (last & 1), (last >> 1) & 1);
    tc_fence_after();
  }
  // ---- epilogue: warp w reads TMEM lanes 32*(w%4).., columns 64*(w/4).. (thread = row), transposes 32 x 32 blocks
  // through shared memory (the operand stages are free: every MMA has completed) and writes each output row segment
  // as one coalesced 128-byte store / atomic ---------------------------------------------------------------------------
  {
    const int q = warp & 3, hf = warp >> 2;
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + 64 * hf;
    float *tile = reinterpret_cast<float *>(tc_smem) + warp * (32 * 33);
#pragma unroll 1
    for (int c0 = 0; c0 < 64; c0 += 32) {
      uint32_t acc[32];
      tmem_ld32(taddr + c0, acc);
      tmem_wait_ld();
#pragma unroll
      for (int c = 0; c < 32; ++c) tile[lane * 33 + c] = __uint_as_float(acc[c]);
      __syncwarp();
      const int j = j0 + 64 * hf + c0 + lane;
      const float bj = (bias && j < J && !(flags & G_ATOMIC)) ? bias[j] : 0.0f;
      if (j < J) {
#pragma unroll 4
        for (int r = 0; r < 32; ++r) {
          const int i = i0 + 32 * q + r;
          if (i >= I) break;
          float v = tile[r * 33 + lane];
          float *dst = C + (long long)i * ldc + j;
          if (flags & G_ATOMIC) { atomicAdd(dst, v); continue; }
          v += bj;
          if (flags & G_ACCUM) v += *dst;
          if (flags & G_RELU) v = fmaxf(v, 0.0f);
          *dst = v;
        }
      }
      __syncwarp();
    }
  }
  // the next tile overwrites the accumulator and the shared-memory tiles the epilogue has just read (its B slab
  // arrives through the async proxy: order it after these generic-proxy accesses)
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  }
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u) : "memory");
  }
}

// The weight operand of a forward / grad-input product is the same for every 128-row tile: it is split ONCE per
// launch into the three bf16 parts, already in the shared-memory image of a B slab ([part][K-chunk][row][8]), so
// that the GEMM's CTAs fetch it with one bulk copy per slab instead of each re-converting it (half of the staging
// instructions of those CTAs).  block = (slab, 128-column tile of the output), thread = (row, 16 reduction elements).
__global__ void __launch_bounds__(256)
split_b_kernel(const float *__restrict__ B, long long sbp, long long sbj, int J, long long P, int n_iter,
               unsigned char *__restrict__ out) {
  const int it = blockIdx.x, row = threadIdx.x & 127, half = threadIdx.x >> 7;
  const long long j = (long long)blockIdx.y * 128 + row, p0 = (long long)it * TC_SLAB + 16 * half;
  float v[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) v[k] = (j < J && p0 + k < P) ? B[(p0 + k) * sbp + j * sbj] : 0.0f;
  unsigned char *dst = out + ((size_t)blockIdx.y * n_iter + it) * TC_OPER + row * 16 + (2 * half) * TC_CHUNK;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    uint32_t w[4][3];
#pragma unroll
    for (int q = 0; q < 4; ++q) split3(v[8 * c + 2 * q], v[8 * c + 2 * q + 1], w[q]);
#pragma unroll
    for (int pt = 0; pt < TC_PARTS; ++pt)
      *reinterpret_cast<uint4 *>(dst + c * TC_CHUNK + pt * TC_PART) = make_uint4(w[0][pt], w[1][pt], w[2][pt], w[3][pt]);
  }
}

// ---- the streaming products (forward, grad-input) as a warp-specialised pipeline --------------------------------------
// gemm_tc_kernel does everything with the same 256 threads: load a slab, convert it, barrier, one thread issues the
// MMAs, epilogue -- the phases of a tile serialise inside a CTA and only the second resident CTA overlaps them (ncu r01:
// 1.5 TB/s, 24 % warps active, long-scoreboard 4.2 per issue: too few bytes in flight).  For the two products whose A
// operand is a [samples x K] activation matrix with 16-byte-aligned rows and whose B operand is a pre-split weight matrix
// (nn.Linear forward and grad-input: two thirds of the step's products) this kernel gives every phase its own warps,
// one persistent CTA per SM:
//   warps 0-1     A loaders: cp.async (16 B per lane and instruction, four full 128-byte lines per warp instruction) of
//                 raw fp32 slabs [128 rows x 32 p] into a 4-stage ring (64 KB in flight per SM, no registers)
//   warp 2        B loader: one bulk copy per slab of the pre-split weight slab (24 KB, L2-resident) into a 3-stage ring
//   warp 3        MMA issuer: six tcgen05.mma per K16 step (the >= 2^-24 terms of the three-part split), accumulators
//                 double-buffered in TMEM (2 x 128 columns)
//   warps 4-11    converters: raw fp32 slab -> three bf16 parts in the UMMA core-matrix layout (2-stage operand ring)
//   warps 12-19   epilogue of tile t while tile t+1 is loaded / converted / multiplied: TMEM -> registers ->
//                 32 x 32 transposes through shared memory -> bias / accumulate / ReLU -> coalesced 128-byte row stores
//                 (two warps per TMEM lane quarter, 64 columns each)
constexpr int WS_THREADS = 640;
constexpr int WS_RAW_STAGES = 4, WS_AOP_STAGES = 2, WS_B_STAGES = 3;
constexpr uint32_t WS_RAW = 128 * TC_SLAB * 4;                           // 16 KB raw fp32 slab, row pitch 128 B
constexpr uint32_t WS_OFF_AOP = WS_RAW_STAGES * WS_RAW;
constexpr uint32_t WS_OFF_B = WS_OFF_AOP + WS_AOP_STAGES * TC_OPER;
constexpr uint32_t WS_OFF_TILE = WS_OFF_B + WS_B_STAGES * TC_OPER;
constexpr uint32_t WS_TILE = 32 * 36 * 4;                                  // per epilogue warp: 32 rows, pitch 36 floats
constexpr uint32_t WS_OFF_BAR = WS_OFF_TILE + 8 * WS_TILE;
constexpr uint32_t WS_SMEM = WS_OFF_BAR + 256;

__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void *src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

template <bool VEC_OUT>
__global__ void __launch_bounds__(WS_THREADS, 1)
gemm_ws_kernel(const float *__restrict__ A, long long sai, float *__restrict__ C, long long ldc,
               const float *__restrict__ bias, int I, int J, int P, int flags, const unsigned char *__restrict__ bsplit) {
  extern __shared__ __align__(1024) unsigned char ws_smem[];
  const uint32_t sbase = smem_u32(ws_smem);
  const uint32_t bars = sbase + WS_OFF_BAR;
  // barriers (8 bytes each)
  const uint32_t raw_full = bars, raw_empty = raw_full + 8 * WS_RAW_STAGES, aop_full = raw_empty + 8 * WS_RAW_STAGES,
                 aop_empty = aop_full + 8 * WS_AOP_STAGES, b_full = aop_empty + 8 * WS_AOP_STAGES,
                 b_empty = b_full + 8 * WS_B_STAGES, acc_full = b_empty + 8 * WS_B_STAGES, acc_empty = acc_full + 16,
                 tmem_slot = acc_empty + 16;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int j0 = blockIdx.y * 128, n_tile_i = (I + 127) / 128;
  const int n_iter = (P + TC_SLAB - 1) / TC_SLAB;
  if (tid == 0) {
    for (int i = 0; i < WS_RAW_STAGES; ++i) { mbar_init(raw_full + 8 * i, 64); mbar_init(raw_empty + 8 * i, 8); }
    for (int i = 0; i < WS_AOP_STAGES; ++i) { mbar_init(aop_full + 8 * i, 8); mbar_init(aop_empty + 8 * i, 1); }
    for (int i = 0; i < WS_B_STAGES; ++i) { mbar_init(b_full + 8 * i, 1); mbar_init(b_empty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(acc_full + 8 * i, 1); mbar_init(acc_empty + 8 * i, 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 3) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(ws_smem + WS_OFF_BAR + (tmem_slot - bars));

  if (warp < 2) {
    // ---- A loaders (warp w: rows 64w .. 64w+63 of every slab) ----------------------------------------------------------------------------------------------------
    uint32_t g = 0;
    for (int ti = blockIdx.x; ti < n_tile_i; ti += gridDim.x) {
      const int i0 = ti * 128;
      for (int it = 0; it < n_iter; ++it, ++g) {
        const uint32_t rs = g % WS_RAW_STAGES;
        mbar_wait(raw_empty + 8 * rs, ((g / WS_RAW_STAGES) & 1) ^ 1);
        const uint32_t dst = sbase + rs * WS_RAW;
        const int ps = it * TC_SLAB;
#pragma unroll 8
        for (int q = 0; q < 16; ++q) {
          const int f = lane + 32 * q + 512 * warp, r = f >> 3, c4 = f & 7;
          const int pp = ps + 4 * c4;
          int nb = (P - pp) * 4;
          nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
          if (i0 + r >= I) nb = 0;
          const float *src = nb ? A + (long long)(i0 + r) * sai + pp : A;
          cp_async16_zfill(dst + r * 128 + ((c4 ^ (r & 7)) << 4), src, (uint32_t)nb);   // XOR swizzle: see the converters
        }
        cp_async_mbar_arrive_noinc(raw_full + 8 * rs);
      }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  } else if (warp == 2) {
    // ---- B loader ----------------------------------------------------------------------------------------------------
    if (lane == 0) {
      uint32_t g = 0;
      for (int ti = blockIdx.x; ti < n_tile_i; ti += gridDim.x)
        for (int it = 0; it < n_iter; ++it, ++g) {
          const uint32_t bs = g % WS_B_STAGES;
          mbar_wait(b_empty + 8 * bs, ((g / WS_B_STAGES) & 1) ^ 1);
          mbar_expect_tx(b_full + 8 * bs, TC_OPER);
          bulk_g2s(sbase + WS_OFF_B + bs * TC_OPER, bsplit + ((size_t)blockIdx.y * n_iter + it) * TC_OPER, TC_OPER,
                   b_full + 8 * bs);
        }
    }
    __syncwarp();
  } else if (warp == 3) {
    // ---- MMA issuer --------------------------------------------------------------------------------------------------
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128);
      uint32_t g = 0, t = 0;
      for (int ti = blockIdx.x; ti < n_tile_i; ti += gridDim.x, ++t) {
        const uint32_t buf = t & 1;
        mbar_wait(acc_empty + 8 * buf, ((t >> 1) & 1) ^ 1);            // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d = tmem_base + 128 * buf;
        for (int it = 0; it < n_iter; ++it, ++g) {
          const uint32_t as = g % WS_AOP_STAGES, bs = g % WS_B_STAGES;
          mbar_wait(aop_full + 8 * as, (g / WS_AOP_STAGES) & 1);
          mbar_wait(b_full + 8 * bs, (g / WS_B_STAGES) & 1);
          tc_fence_after();
          const uint32_t a0 = sbase + WS_OFF_AOP + as * TC_OPER, b0 = sbase + WS_OFF_B + bs * TC_OPER;
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            uint64_t da[TC_PARTS], db[TC_PARTS];
#pragma unroll
            for (int pt = 0; pt < TC_PARTS; ++pt) {
              da[pt] = umma_desc(a0 + pt * TC_PART + u * 2 * TC_CHUNK, TC_CHUNK, 128);
              db[pt] = umma_desc(b0 + pt * TC_PART + u * 2 * TC_CHUNK, TC_CHUNK, 128);
            }
            umma_ss(d, da[0], db[0], idesc, (it > 0 || u > 0) ? 1u : 0u);
            umma_ss(d, da[0], db[1], idesc, 1u);
            umma_ss(d, da[1], db[0], idesc, 1u);
            umma_ss(d, da[0], db[2], idesc, 1u);
            umma_ss(d, da[1], db[1], idesc, 1u);
            umma_ss(d, da[2], db[0], idesc, 1u);
          }
          umma_commit(aop_empty + 8 * as);
          umma_commit(b_empty + 8 * bs);
        }
        umma_commit(acc_full + 8 * buf);
      }
    }
    __syncwarp();
  } else if (warp < 12) {
    // ---- converters: raw fp32 slab -> three bf16 parts in the UMMA K-major core-matrix layout -----------------------------
    const int ct = tid - 128;
    uint32_t g = 0;
    for (int ti = blockIdx.x; ti < n_tile_i; ti += gridDim.x)
      for (int it = 0; it < n_iter; ++it, ++g) {
        const uint32_t rs = g % WS_RAW_STAGES, as = g % WS_AOP_STAGES;
        mbar_wait(raw_full + 8 * rs, (g / WS_RAW_STAGES) & 1);
        mbar_wait(aop_empty + 8 * as, ((g / WS_AOP_STAGES) & 1) ^ 1);     // the MMAs of slab g-2 have read this stage
        const uint32_t raw = sbase + rs * WS_RAW, base = sbase + WS_OFF_AOP + as * TC_OPER;
        // lane -> (row, 16-byte column): 8 consecutive lanes take 8 consecutive rows of ONE column, the next 8 the column's
        // pair, lanes 16-31 the next 8 rows.  With the loader's XOR swizzle a quarter-warp's 128-bit loads hit eight
        // different 16-byte bank groups, and a half-warp's 64-bit stores fill one 128-byte row of banks of the operand
        // chunk (8 rows x 16 B): both directions conflict-free (the row-major mapping stored with 4-way conflicts, which
        // made the shared-memory pipe -- shared with the MMAs' operand reads -- the limiter).
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = (ct & 7) | (((ct >> 4) & 1) << 3) | ((ct >> 5) << 4), c4 = ((ct >> 3) & 1) | (q << 1);
          const uint4 v = lds128(raw + r * 128 + ((c4 ^ (r & 7)) << 4));
          uint32_t w0[3], w1[3];
          split3(__uint_as_float(v.x), __uint_as_float(v.y), w0);
          split3(__uint_as_float(v.z), __uint_as_float(v.w), w1);
          const uint32_t addr = base + (c4 >> 1) * TC_CHUNK + r * 16 + (c4 & 1) * 8;
#pragma unroll
          for (int pt = 0; pt < TC_PARTS; ++pt)
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr + pt * TC_PART), "r"(w0[pt]), "r"(w1[pt]) : "memory");
        }
        fence_async_smem();                                            // generic-proxy stores -> the MMA's async-proxy reads
        __syncwarp();
        if (lane == 0) { mbar_arrive(aop_full + 8 * as); mbar_arrive(raw_empty + 8 * rs); }
      }
  } else {
    // ---- epilogue warps: warp w owns TMEM lanes 32*(w%4).. (rows), all 128 columns of the tile ----------------------------
    const int q = warp & 3;
    const int ch = (warp - 12) >> 2;                                 // this warp's column half: chunks 64*ch, 64*ch + 32
    float *tile = reinterpret_cast<float *>(ws_smem + WS_OFF_TILE + (warp - 12) * WS_TILE);
    const uint32_t tile_a = sbase + WS_OFF_TILE + (warp - 12) * WS_TILE;
    const bool accum = flags & G_ACCUM, relu = flags & G_RELU;
    float4 bias4[2];                                                 // this lane's four columns of its two chunks (VEC_OUT)
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      const int j = j0 + 64 * ch + 32 * cc + 4 * (lane & 7);
      bias4[cc] = (VEC_OUT && bias && j < J) ? *reinterpret_cast<const float4 *>(bias + j) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    uint32_t t = 0;
    for (int ti = blockIdx.x; ti < n_tile_i; ti += gridDim.x, ++t) {
      const int i0 = ti * 128;
      const uint32_t buf = t & 1;
      mbar_wait_backoff(acc_full + 8 * buf, (t >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + 128 * buf;
      const int nrow = min(32, I - (i0 + 32 * q));                    // rows of this warp's quarter inside the matrix
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c0 = 64 * ch + 32 * cc;
        uint32_t acc[32];
        tmem_ld32(taddr + c0, acc);
        tmem_wait_ld();
        if (VEC_OUT) {
          // 32 x 32 transpose through shared memory in 16-byte pieces (row pitch 144 B: conflict-free both ways), so that a
          // warp instruction writes four full 128-byte lines: 8 STS.128 + 8 LDS.128 + 8 STG.128 per thread and chunk (the
          // scalar form below costs 2.4 K instructions per tile and warp and was the kernel's bottleneck)
#pragma unroll
          for (int k = 0; k < 8; ++k) sts128(tile_a + lane * 144 + k * 16, acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
          __syncwarp();
          const int cq = lane & 7, rq = lane >> 3;
          const int j = j0 + c0 + 4 * cq;                             // J % 4 == 0: the four columns are inside or outside together
          if (j < J) {
            const float4 b4 = bias4[cc];
            float *dst = C + (long long)(i0 + 32 * q + rq) * ldc + j;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
              const uint4 w = lds128(tile_a + (4 * b + rq) * 144 + cq * 16);
              float4 v = make_float4(__uint_as_float(w.x) + b4.x, __uint_as_float(w.y) + b4.y, __uint_as_float(w.z) + b4.z,
                                     __uint_as_float(w.w) + b4.w);
              if (4 * b + rq < nrow) {
                float4 *d4 = reinterpret_cast<float4 *>(dst + (long long)(4 * b) * ldc);
                if (accum) { const float4 o = *d4; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
                if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
                *d4 = v;
              }
            }
          }
          __syncwarp();
        } else {
#pragma unroll
          for (int c = 0; c < 32; ++c) tile[lane * 36 + c] = __uint_as_float(acc[c]);
          __syncwarp();
          const int j = j0 + c0 + lane;
          const float bj = (bias && j < J) ? bias[j] : 0.0f;
          if (j < J) {
            float *dst = C + (long long)(i0 + 32 * q) * ldc + j;
#pragma unroll 1
            for (int r0 = 0; r0 < 32; r0 += 8) {
              float v[8];
#pragma unroll
              for (int k = 0; k < 8; ++k) v[k] = tile[(r0 + k) * 36 + lane] + bj;
              if (accum) {
#pragma unroll
                for (int k = 0; k < 8; ++k) if (r0 + k < nrow) v[k] += dst[(long long)(r0 + k) * ldc];
              }
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                if (relu) v[k] = fmaxf(v[k], 0.0f);
                if (r0 + k < nrow) dst[(long long)(r0 + k) * ldc] = v[k];
              }
            }
          }
          __syncwarp();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty + 8 * buf);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u) : "memory");
  }
}

// ---- the grad-weight product as the same kind of pipeline ----------------------------------------------------------------
// dW[i][j] += sum_p dY[p][i] X[p][j]: both operands are activation matrices whose CONTIGUOUS index is the output index, not
// the reduction index.  gemm_tc_kernel stages them thread = row with scalar loads and converts both in every CTA: its time
// does not depend on K (105 us at K = 64 and K = 128, M = 258 048) -- conversion- and latency-bound, not HBM-bound.  Here the
// raw [32 samples x 128] fp32 slabs go straight into shared memory by cp.async and are converted in their natural order into
// MN-MAJOR operands (UMMA no-swizzle MN-major: 8 consecutive output indices are the 16 contiguous bytes of a core-matrix row,
// the 8 reduction steps of a core matrix are 16 bytes apart; element (mn, k) at (k/8)*LBO + (mn/8)*SBO + (k%8)*16 + (mn%8)*2
// with SBO = 128, LBO = 2048): no transposition anywhere.  One persistent CTA per SM and output tile takes a contiguous range
// of the reduction, accumulates the 128 x 128 tile in TMEM and adds it to dW with coalesced fp32 atomics at the end.
//   warp 0        dY loader (cp.async 16 B)          warps 1-2   X loader (16 B, or 4 B when X's rows are not 16-byte aligned)
//   warp 3        MMA issuer                          warps 4-19  converters (8: dY, 8: X), then the epilogue
constexpr int WW_STAGES = 3;
constexpr uint32_t WW_RAW = 32 * 128 * 4;                                 // one operand's raw slab: 32 samples x 128 floats
constexpr uint32_t WW_OFF_OP = WW_STAGES * 2 * WW_RAW;
constexpr uint32_t WW_OFF_BAR = WW_OFF_OP + 2 * 2 * TC_OPER;
constexpr uint32_t WW_SMEM = WW_OFF_BAR + 256;

__device__ __forceinline__ void cp_async4_zfill(uint32_t dst, const void *src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ uint32_t umma_idesc_bf16_mn(int N) {          // as umma_idesc_bf16, A and B MN-major
  return umma_idesc_bf16(N) | (1u << 15) | (1u << 16);
}

template <bool X_VEC>
__global__ void __launch_bounds__(WS_THREADS, 1)
gemm_wsw_kernel(const float *__restrict__ dY, long long ldy, const float *__restrict__ X, long long ldx,
                float *__restrict__ C, long long ldc, int I, int J, long long P, int slabs_per_cta) {
  extern __shared__ __align__(1024) unsigned char ww_smem[];
  const uint32_t sbase = smem_u32(ww_smem);
  const uint32_t bars = sbase + WW_OFF_BAR;
  const uint32_t rawa_full = bars, rawb_full = rawa_full + 8 * WW_STAGES, raw_empty = rawb_full + 8 * WW_STAGES,
                 op_full = raw_empty + 8 * WW_STAGES, op_empty = op_full + 16, acc_full = op_empty + 16, tmem_slot = acc_full + 8;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_tj = (J + 127) / 128;
  const int i0 = ((int)blockIdx.y / n_tj) * 128, j0 = ((int)blockIdx.y % n_tj) * 128;
  const long long n_slab_all = (P + TC_SLAB - 1) / TC_SLAB;
  const long long s_begin = (long long)blockIdx.x * slabs_per_cta;
  const int n_slab = (int)max(0ll, min((long long)slabs_per_cta, n_slab_all - s_begin));
  if (tid == 0) {
    for (int i = 0; i < WW_STAGES; ++i) {
      mbar_init(rawa_full + 8 * i, 32); mbar_init(rawb_full + 8 * i, 64); mbar_init(raw_empty + 8 * i, 16);
    }
    for (int i = 0; i < 2; ++i) { mbar_init(op_full + 8 * i, 16); mbar_init(op_empty + 8 * i, 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 3) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(ww_smem + WW_OFF_BAR + (tmem_slot - bars));

  if (warp == 0) {
    // ---- dY loader: 32 samples x 128 output rows i0.. of dY, 16 bytes per lane and instruction ------------------------------
    for (int g = 0; g < n_slab; ++g) {
      const uint32_t rs = g % WW_STAGES;
      mbar_wait(raw_empty + 8 * rs, ((g / WW_STAGES) & 1) ^ 1);
      const uint32_t dst = sbase + rs * 2 * WW_RAW;
      const long long p0 = (s_begin + g) * TC_SLAB;
#pragma unroll 8
      for (int q = 0; q < 32; ++q) {
        const int f = lane + 32 * q, pl = f >> 5, c4 = f & 31;
        int nb = (I - i0 - 4 * c4) * 4;
        nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
        if (p0 + pl >= P) nb = 0;
        const float *src = nb ? dY + (p0 + pl) * ldy + i0 + 4 * c4 : dY;
        cp_async16_zfill(dst + pl * 512 + ((c4 ^ (pl & 7)) << 4), src, (uint32_t)nb);
      }
      cp_async_mbar_arrive_noinc(rawa_full + 8 * rs);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  } else if (warp < 3) {
    // ---- X loaders (warp 1: samples 0..15 of a slab, warp 2: 16..31) -----------------------------------------------------------
    const int wh = warp - 1;
    for (int g = 0; g < n_slab; ++g) {
      const uint32_t rs = g % WW_STAGES;
      mbar_wait(raw_empty + 8 * rs, ((g / WW_STAGES) & 1) ^ 1);
      const uint32_t dst = sbase + rs * 2 * WW_RAW + WW_RAW;
      const long long p0 = (s_begin + g) * TC_SLAB;
      if (X_VEC) {
#pragma unroll 8
        for (int q = 0; q < 16; ++q) {
          const int f = lane + 32 * q + 512 * wh, pl = f >> 5, c4 = f & 31;
          int nb = (J - j0 - 4 * c4) * 4;
          nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
          if (p0 + pl >= P) nb = 0;
          const float *src = nb ? X + (p0 + pl) * ldx + j0 + 4 * c4 : X;
          cp_async16_zfill(dst + pl * 512 + ((c4 ^ (pl & 7)) << 4), src, (uint32_t)nb);
        }
      } else {
#pragma unroll 8
        for (int q = 0; q < 64; ++q) {
          const int e = lane + 32 * q + 2048 * wh, pl = e >> 7, j = e & 127;
          const bool ok = p0 + pl < P && j0 + j < J;
          const float *src = ok ? X + (p0 + pl) * ldx + j0 + j : X;
          cp_async4_zfill(dst + pl * 512 + ((((j >> 2) ^ (pl & 7))) << 4) + (j & 3) * 4, src, ok ? 4u : 0u);
        }
      }
      cp_async_mbar_arrive_noinc(rawb_full + 8 * rs);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  } else if (warp == 3) {
    // ---- MMA issuer ---------------------------------------------------------------------------------------------------------------
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16_mn(128);
      for (int g = 0; g < n_slab; ++g) {
        const uint32_t os = g & 1;
        mbar_wait(op_full + 8 * os, (g >> 1) & 1);
        tc_fence_after();
        const uint32_t a0 = sbase + WW_OFF_OP + os * 2 * TC_OPER, b0 = a0 + TC_OPER;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          uint64_t da[TC_PARTS], db[TC_PARTS];
#pragma unroll
          for (int pt = 0; pt < TC_PARTS; ++pt) {
            da[pt] = umma_desc(a0 + pt * TC_PART + u * 2 * TC_CHUNK, TC_CHUNK, 128);
            db[pt] = umma_desc(b0 + pt * TC_PART + u * 2 * TC_CHUNK, TC_CHUNK, 128);
          }
          umma_ss(tmem_base, da[0], db[0], idesc, (g > 0 || u > 0) ? 1u : 0u);
          umma_ss(tmem_base, da[0], db[1], idesc, 1u);
          umma_ss(tmem_base, da[1], db[0], idesc, 1u);
          umma_ss(tmem_base, da[0], db[2], idesc, 1u);
          umma_ss(tmem_base, da[1], db[1], idesc, 1u);
          umma_ss(tmem_base, da[2], db[0], idesc, 1u);
        }
        umma_commit(op_empty + 8 * os);
      }
      umma_commit(acc_full);
    }
    __syncwarp();
  } else {
    // ---- converters: warps 4-11 dY, 12-19 X; lane -> (sample, 16-byte column) as in gemm_ws_kernel: conflict-free both ways ----
    const int ct = tid - 128, opnd = ct >> 8, c = ct & 255;
    for (int g = 0; g < n_slab; ++g) {
      const uint32_t rs = g % WW_STAGES, os = g & 1;
      mbar_wait((opnd ? rawb_full : rawa_full) + 8 * rs, (g / WW_STAGES) & 1);
      mbar_wait(op_empty + 8 * os, ((g >> 1) & 1) ^ 1);
      const uint32_t raw = sbase + rs * 2 * WW_RAW + opnd * WW_RAW;
      const uint32_t base = sbase + WW_OFF_OP + os * 2 * TC_OPER + opnd * TC_OPER;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int pl = (c & 7) | (q << 3), c4 = ((c >> 3) & 1) | (((c >> 4) & 15) << 1);
        const uint4 v = lds128(raw + pl * 512 + ((c4 ^ (pl & 7)) << 4));
        uint32_t w0[3], w1[3];
        split3(__uint_as_float(v.x), __uint_as_float(v.y), w0);
        split3(__uint_as_float(v.z), __uint_as_float(v.w), w1);
        const uint32_t addr = base + q * TC_CHUNK + (c4 >> 1) * 128 + (pl & 7) * 16 + (c4 & 1) * 8;
#pragma unroll
        for (int pt = 0; pt < TC_PARTS; ++pt)
          asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr + pt * TC_PART), "r"(w0[pt]), "r"(w1[pt]) : "memory");
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) { mbar_arrive(op_full + 8 * os); mbar_arrive(raw_empty + 8 * rs); }
    }
    // ---- epilogue: warp (4 + 4b + q) owns TMEM lanes 32q.. and columns 32b..: transpose through the (now idle) raw ring and
    // add the 32 x 32 block to dW with one coalesced 128-byte atomic row per instruction ------------------------------------------
    if (n_slab > 0) {
      const int q = warp & 3, b = (warp - 4) >> 2;
      mbar_wait_backoff(acc_full, 0);
      tc_fence_after();
      uint32_t acc[32];
      tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + 32 * b, acc);
      tmem_wait_ld();
      float *tile = reinterpret_cast<float *>(ww_smem) + (warp - 4) * (32 * 33);
#pragma unroll
      for (int k = 0; k < 32; ++k) tile[lane * 33 + k] = __uint_as_float(acc[k]);
      __syncwarp();
      const int j = j0 + 32 * b + lane;
      if (j < J) {
        const int nrow = min(32, I - (i0 + 32 * q));
        float *dst = C + (long long)(i0 + 32 * q) * ldc + j;
#pragma unroll 8
        for (int r = 0; r < 32; ++r)
          if (r < nrow) atomicAdd(dst + (long long)r * ldc, tile[r * 33 + lane]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u) : "memory");
  }
}

// ---- the 3- and 1-wide heads (confi_rgb / alpha_1 / rgb / alpha, network/models.py:161-162,170,174) ------------------
// A 64- or 128-wide GEMM tile would be >95 % padding there; these three kernels stream the wide operand once, coalesced.
// forward: warp per row, lanes over k, shuffle reduction.  Y[m][n] (+)= sum_k X[m][k] W[n][k] + b[n], N <= 4
__global__ void __launch_bounds__(256)
skinny_fwd_kernel(const float *__restrict__ X, long long ldx, const float *__restrict__ W, long long ldw,
                  const float *__restrict__ b, long long M, int K, int N, float *__restrict__ Y, long long ldy, int flags) {
  const int lane = threadIdx.x & 31;
  const long long m = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (m >= M) return;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k = lane; k < K; k += 32) {
    const float x = X[m * ldx + k];
#pragma unroll
    for (int n = 0; n < 4; ++n)
      if (n < N) acc[n] = fmaf(x, W[n * ldw + k], acc[n]);
  }
#pragma unroll
  for (int n = 0; n < 4; ++n)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], o);
  if (lane < N) {
    float v = lane == 0 ? acc[0] : (lane == 1 ? acc[1] : (lane == 2 ? acc[2] : acc[3]));
    if (b) v += b[lane];
    float *dst = Y + m * ldy + lane;
    if (flags & G_ACCUM) v += *dst;
    if (flags & G_RELU) v = fmaxf(v, 0.0f);
    *dst = v;
  }
}
// grad-input: dX[m][k] (+)= sum_{n<N} dY[m][n] W[n][k], N <= 4 (thread per output element)
__global__ void __launch_bounds__(256)
skinny_bwd_input_kernel(const float *__restrict__ dY, long long ldy, const float *__restrict__ W, long long ldw,
                        long long M, int K, int N, float *__restrict__ dX, long long ldx, int flags) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M * K) return;
  const long long m = e / K;
  const int k = (int)(e - m * K);
  float v = 0.f;
  for (int n = 0; n < N; ++n) v = fmaf(dY[m * ldy + n], W[n * ldw + k], v);
  float *dst = dX + m * ldx + k;
  *dst = (flags & G_ACCUM) ? *dst + v : v;
}
// grad-weight: dW[n][k] += sum_m dY[m][n] X[m][k], N <= 4: thread = k, block = one chunk of rows, atomics at the end
__global__ void __launch_bounds__(128)
skinny_bwd_weight_kernel(const float *__restrict__ dY, long long ldy, const float *__restrict__ X, long long ldx,
                         long long M, int K, int N, long long m_chunk, float *dW, long long lddw) {
  const int k = blockIdx.y * 128 + threadIdx.x;
  if (k >= K) return;
  const long long m0 = (long long)blockIdx.x * m_chunk, m1 = min(M, m0 + m_chunk);
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  long long m = m0;
  for (; m + 8 <= m1; m += 8) {                                   // eight independent row loads in flight per thread
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = X[(m + i) * ldx + k];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int n = 0; n < 4; ++n)
        if (n < N) acc[n] = fmaf(dY[(m + i) * ldy + n], x[i], acc[n]);
  }
  for (; m < m1; ++m) {
    const float x = X[m * ldx + k];
#pragma unroll
    for (int n = 0; n < 4; ++n)
      if (n < N) acc[n] = fmaf(dY[m * ldy + n], x, acc[n]);
  }
  for (int n = 0; n < N; ++n) atomicAdd(dW + n * lddw + k, acc[n]);
}

int g_gemm_mode = 0;     // 0: tensor cores where the shape fills a tile, 1: SIMT fp32 only

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
// thread = sample computes its 63 + 27 values into shared memory (odd pitch: conflict-free), then the block writes
// the rows out with consecutive lanes on consecutive columns (a per-thread row write would touch one 32-byte sector
// per lane and store instruction)
constexpr int PREP_T = 128, PREP_LD = 91;
__global__ void __launch_bounds__(PREP_T)
prep_kernel(const float *__restrict__ ndc, const float *__restrict__ angle, int S, long long N,
            float *__restrict__ ph, long long ldph, float *__restrict__ ftv, long long ldftv) {
  __shared__ float sm[PREP_T * PREP_LD];
  const long long q0 = (long long)blockIdx.x * PREP_T, q = q0 + threadIdx.x;
  if (q < N) {
    const float x[3] = {ndc[3 * q], ndc[3 * q + 1], ndc[3 * q + 2]};
    const long long r = q / S;
    const float a[3] = {angle[3 * r], angle[3 * r + 1], angle[3 * r + 2]};
    embed(x, 10, sm + threadIdx.x * PREP_LD);              // PE(ndc)  -> ph[:, 0:63]
    embed(a, 4, sm + threadIdx.x * PREP_LD + 63);          // PE(dir)  -> ftv[:, 128:155]
  }
  __syncthreads();
  const int rows = (int)min((long long)PREP_T, N - q0);
  for (int e = threadIdx.x; e < rows * 63; e += PREP_T) {
    const int r = e / 63, c = e - 63 * r;
    ph[(q0 + r) * ldph + c] = sm[r * PREP_LD + c];
  }
  for (int e = threadIdx.x; e < rows * 27; e += PREP_T) {
    const int r = e / 27, c = e - 27 * r;
    ftv[(q0 + r) * ldftv + 128 + c] = sm[r * PREP_LD + 63 + c];
  }
}

__global__ void film_relu_fwd_kernel(const float *__restrict__ pre, const float *__restrict__ mod, float *__restrict__ h,
                                     long long ldh, long long M) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M * 128) return;
  h[(e >> 7) * ldh + (e & 127)] = fmaxf(pre[e] * mod[e], 0.0f);
}
// thread = column, block = 64 rows: coalesced, and the bias gradient of the layer (column sums of dpre) falls out of
// the same pass (one atomic per thread)
__global__ void __launch_bounds__(128)
film_relu_bwd_kernel(const float *__restrict__ dh, long long lddh, const float *__restrict__ pre,
                     const float *__restrict__ mod, float *__restrict__ dpre, float *__restrict__ dmod, int accumulate,
                     float *db, long long M) {
  const int c = threadIdx.x;
  const long long m0 = (long long)blockIdx.x * 64, m1 = min(M, m0 + 64);
  float sum = 0.f;
  for (long long m = m0; m < m1; ++m) {
    const long long e = m * 128 + c;
    const float p = pre[e], md = mod[e];
    const float g = p * md > 0.0f ? dh[m * lddh + c] : 0.0f;      // == h > 0 (h = relu(pre * mod)): h is not re-read
    const float dp = g * md;
    dpre[e] = dp;
    sum += dp;
    const float dm = g * p;
    dmod[e] = accumulate ? dmod[e] + dm : dm;
  }
  if (db) atomicAdd(db + c, sum);
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

// ---- 16-byte variants of the streaming kernels above ------------------------------------------------------------------
// One float4 per thread and per operand (4x the bytes in flight of the scalar kernels, which reach 35-60 % of the HBM
// roofline at M = 258 048); the launchers use them whenever every pointer is 16-byte aligned and every pitch a multiple
// of four floats, and fall back to the scalar kernels otherwise (h of pts layer 4 lives at column 63 of [PE63 | h]:
// its loads / stores stay scalar through the ALIGNED template flag).
template <bool ALIGNED> __device__ __forceinline__ float4 ld4(const float *p) {
  if (ALIGNED) return *reinterpret_cast<const float4 *>(p);
  return make_float4(p[0], p[1], p[2], p[3]);
}
template <bool ALIGNED> __device__ __forceinline__ void st4(float *p, const float4 v) {
  if (ALIGNED) { *reinterpret_cast<float4 *>(p) = v; return; }
  p[0] = v.x; p[1] = v.y; p[2] = v.z; p[3] = v.w;
}
inline bool al16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <bool H_AL>
__global__ void __launch_bounds__(256)
film_relu_fwd_v4_kernel(const float *__restrict__ pre, const float *__restrict__ mod, float *__restrict__ h, long long ldh,
                        long long M) {
  const long long e4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e4 >= M * 32) return;
  const float4 p = ld4<true>(pre + 4 * e4), md = ld4<true>(mod + 4 * e4);
  st4<H_AL>(h + (e4 >> 5) * ldh + 4 * (e4 & 31),
            make_float4(fmaxf(p.x * md.x, 0.0f), fmaxf(p.y * md.y, 0.0f), fmaxf(p.z * md.z, 0.0f), fmaxf(p.w * md.w, 0.0f)));
}
// 128 threads = 4 rows x 32 column groups per pass, 64 rows per block; column sums through shared memory
template <bool DH_AL>
__global__ void __launch_bounds__(128)
film_relu_bwd_v4_kernel(const float *__restrict__ dh, long long lddh, const float *__restrict__ pre,
                        const float *__restrict__ mod, float *__restrict__ dpre, float *__restrict__ dmod, int accumulate,
                        float *db, long long M) {
  const int cg = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const long long m0 = (long long)blockIdx.x * 64;
  float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int i = 0; i < 16; ++i) {
    const long long m = m0 + 4 * i + rg;
    if (m >= M) break;
    const long long e = m * 128 + 4 * cg;
    const float4 p = ld4<true>(pre + e), md = ld4<true>(mod + e), g0 = ld4<DH_AL>(dh + m * lddh + 4 * cg);
    float4 dm = accumulate ? ld4<true>(dmod + e) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 g = make_float4(p.x * md.x > 0.0f ? g0.x : 0.0f, p.y * md.y > 0.0f ? g0.y : 0.0f,
                                 p.z * md.z > 0.0f ? g0.z : 0.0f, p.w * md.w > 0.0f ? g0.w : 0.0f);
    const float4 dp = make_float4(g.x * md.x, g.y * md.y, g.z * md.z, g.w * md.w);
    st4<true>(dpre + e, dp);
    sum.x += dp.x; sum.y += dp.y; sum.z += dp.z; sum.w += dp.w;
    dm.x += g.x * p.x; dm.y += g.y * p.y; dm.z += g.z * p.z; dm.w += g.w * p.w;
    st4<true>(dmod + e, dm);
  }
  if (db) {
    __shared__ float4 part[4][32];
    part[rg][cg] = sum;
    __syncthreads();
    if (rg == 0) {
#pragma unroll
      for (int k = 1; k < 4; ++k) { const float4 o = part[k][cg]; sum.x += o.x; sum.y += o.y; sum.z += o.z; sum.w += o.w; }
      atomicAdd(db + 4 * cg, sum.x); atomicAdd(db + 4 * cg + 1, sum.y);
      atomicAdd(db + 4 * cg + 2, sum.z); atomicAdd(db + 4 * cg + 3, sum.w);
    }
  }
}
template <bool A_AL>
__global__ void __launch_bounds__(256)
mul_fwd_v4_kernel(const float *__restrict__ a, long long lda, const float *__restrict__ b, float *__restrict__ o, long long M) {
  const long long e4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e4 >= M * 32) return;
  const float4 x = ld4<A_AL>(a + (e4 >> 5) * lda + 4 * (e4 & 31)), y = ld4<true>(b + 4 * e4);
  st4<true>(o + 4 * e4, make_float4(x.x * y.x, x.y * y.y, x.z * y.z, x.w * y.w));
}
template <bool A_AL>
__global__ void __launch_bounds__(256)
mul_bwd_v4_kernel(const float *__restrict__ d_o, const float *__restrict__ a, long long lda, const float *__restrict__ b,
                  float *__restrict__ da, int accumulate, float *__restrict__ db, long long M) {
  const long long e4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e4 >= M * 32) return;
  const float4 g = ld4<true>(d_o + 4 * e4), y = ld4<true>(b + 4 * e4), x = ld4<A_AL>(a + (e4 >> 5) * lda + 4 * (e4 & 31));
  float4 v = accumulate ? ld4<true>(da + 4 * e4) : make_float4(0.f, 0.f, 0.f, 0.f);
  v.x += g.x * y.x; v.y += g.y * y.y; v.z += g.z * y.z; v.w += g.w * y.w;
  st4<true>(da + 4 * e4, v);
  st4<true>(db + 4 * e4, make_float4(g.x * x.x, g.y * x.y, g.z * x.z, g.w * x.w));
}
__global__ void __launch_bounds__(256)
relu_bwd_v4_kernel(const float *__restrict__ dy, const float *__restrict__ y, float *__restrict__ dx, long long n4) {
  const long long e4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e4 >= n4) return;
  const float4 g = ld4<true>(dy + 4 * e4), v = ld4<true>(y + 4 * e4);
  st4<true>(dx + 4 * e4, make_float4(v.x > 0.0f ? g.x : 0.0f, v.y > 0.0f ? g.y : 0.0f, v.z > 0.0f ? g.z : 0.0f, v.w > 0.0f ? g.w : 0.0f));
}
// heads, K a multiple of 128: forward = warp per row, one float4 per lane and 128 columns
__global__ void __launch_bounds__(256)
skinny_fwd_v4_kernel(const float *__restrict__ X, long long ldx, const float *__restrict__ W, long long ldw,
                     const float *__restrict__ b, long long M, int K, int N, float *__restrict__ Y, long long ldy, int flags) {
  const int lane = threadIdx.x & 31;
  const long long m = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (m >= M) return;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k = 4 * lane; k < K; k += 128) {
    const float4 x = ld4<true>(X + m * ldx + k);
#pragma unroll
    for (int n = 0; n < 4; ++n)
      if (n < N) {
        const float4 w = ld4<true>(W + n * ldw + k);
        acc[n] = fmaf(x.x, w.x, fmaf(x.y, w.y, fmaf(x.z, w.z, fmaf(x.w, w.w, acc[n]))));
      }
  }
#pragma unroll
  for (int n = 0; n < 4; ++n)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], o);
  if (lane < N) {
    float v = lane == 0 ? acc[0] : (lane == 1 ? acc[1] : (lane == 2 ? acc[2] : acc[3]));
    if (b) v += b[lane];
    float *dst = Y + m * ldy + lane;
    if (flags & G_ACCUM) v += *dst;
    if (flags & G_RELU) v = fmaxf(v, 0.0f);
    *dst = v;
  }
}
// grad-input of the heads: thread = four consecutive k of one row
__global__ void __launch_bounds__(256)
skinny_bwd_input_v4_kernel(const float *__restrict__ dY, long long ldy, const float *__restrict__ W, long long ldw,
                           long long M, int K, int N, float *__restrict__ dX, long long ldx, int flags) {
  const long long e4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int K4 = K >> 2;
  if (e4 >= M * K4) return;
  const long long m = e4 / K4;
  const int k = 4 * (int)(e4 - m * K4);
  float *dst = dX + m * ldx + k;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int n = 0; n < N; ++n) {
    const float g = dY[m * ldy + n];
    const float4 w = ld4<true>(W + n * ldw + k);
    v.x = fmaf(g, w.x, v.x); v.y = fmaf(g, w.y, v.y); v.z = fmaf(g, w.z, v.z); v.w = fmaf(g, w.w, v.w);
  }
  if (flags & G_ACCUM) { const float4 o = ld4<true>(dst); v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
  st4<true>(dst, v);
}

inline unsigned nblk(long long n, int t) { return (unsigned)((n + t - 1) / t); }

// scratch for the pre-split weight operand, one buffer per (device, stream): launches on one stream are ordered, so
// the next product may overwrite it
constexpr size_t kSplitBytes = 1u << 20;
unsigned char *split_scratch(void *stream) {
  struct Ent { int dev; void *stream; unsigned char *buf; };
  static std::mutex mu;
  static std::vector<Ent> tab;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(mu);
  for (const Ent &e : tab) if (e.dev == dev && e.stream == stream) return e.buf;
  if (tab.size() >= 16) return nullptr;                     // many streams: the remaining ones convert in the CTAs
  void *p = nullptr;
  if (cudaMalloc(&p, kSplitBytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  tab.push_back({dev, stream, static_cast<unsigned char *>(p)});
  return static_cast<unsigned char *>(p);
}

int g_gemm_dbg = 0;   // timing switches for tools/ubench_gemm.py: 16 no pre-split B, 32 one tile per CTA (results unchanged)
int launch_gemm(const float *A, long long sai, long long sap, const float *B, long long sbp, long long sbj, float *C,
                long long ldc, const float *bias, long long I, long long J, long long P, int flags, int splits,
                void *stream) {
  if (I <= 0 || J <= 0) return UCN_OK;
  // heads (3 / 1 outputs) and tiny reductions stay on the CUDA cores: a 128 x 128 tensor-core tile would be padding
  const bool tc = g_gemm_mode == 0 && I >= 64 && J >= 16 && P >= 16;
  const int tile = tc ? 128 : BT, pstep = tc ? TC_SLAB : PT;
  long long p_chunk = P;
  if (splits > 1) {
    p_chunk = ((P + splits - 1) / splits + pstep - 1) / pstep * pstep;
    splits = (int)((P + p_chunk - 1) / p_chunk);
  } else {
    splits = 1;
  }
  dim3 grid(nblk(I, tile), nblk(J, tile), (unsigned)splits);
  UCN_CHECK_ARG(grid.y < 65536u, "ucn_linear: output too wide");
  if (tc) {
    static int attr[kUcnMaxDevices];
    {
      cudaError_t e = ucn_ensure_dyn_smem(gemm_tc_kernel, (int)TC_SMEM, attr);
      if (e != cudaSuccess) return ucn_cuda_status(e, "gemm_tc smem attribute");
    }
    // float4 staging needs p contiguous, 16-byte aligned rows, and the last (partial) float4 of a row inside its pitch
    const long long P4 = (P + 3) & ~3ll;
    const int a_vec = sap == 1 && splits == 1 && (reinterpret_cast<uintptr_t>(A) & 15) == 0 && (sai & 3) == 0 && P4 <= sai;
    const int b_vec = sbp == 1 && splits == 1 && (reinterpret_cast<uintptr_t>(B) & 15) == 0 && (sbj & 3) == 0 && P4 <= sbj;
    // B is a weight matrix shared by many row tiles (forward, grad-input): split it once (split_b_kernel)
    const int n_iter = (int)((P + TC_SLAB - 1) / TC_SLAB);
    unsigned char *bsplit = nullptr;
    if (splits == 1 && grid.x >= 8 && (size_t)grid.y * n_iter * TC_OPER <= kSplitBytes) bsplit = split_scratch(stream);
    if (bsplit) {
      split_b_kernel<<<dim3((unsigned)n_iter, grid.y), 256, 0, (cudaStream_t)stream>>>(B, sbp, sbj, (int)J, P, n_iter, bsplit);
      UCN_LAUNCH_CHECK("split_b_kernel");
    }
    // forward / grad-input with aligned activation rows: the warp-specialised pipeline (one persistent CTA per SM)
    if (bsplit && a_vec && !(flags & G_ATOMIC) && !(g_gemm_dbg & 64) && P < (1ll << 30)) {
      static int attr_ws[kUcnMaxDevices];
      static int attr_ws2[kUcnMaxDevices];
      // 16-byte output stores need aligned rows of C (grad-input of the [PE63 | h] layer writes rows of 191 floats: scalar form)
      const bool vec_out = (reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc & 3) == 0 && (J & 3) == 0 &&
                           (!bias || (reinterpret_cast<uintptr_t>(bias) & 15) == 0);
      cudaError_t e = vec_out ? ucn_ensure_dyn_smem(gemm_ws_kernel<true>, (int)WS_SMEM, attr_ws)
                              : ucn_ensure_dyn_smem(gemm_ws_kernel<false>, (int)WS_SMEM, attr_ws2);
      if (e != cudaSuccess) return ucn_cuda_status(e, "gemm_ws smem attribute");
      const unsigned cap = (unsigned)ucn_sm_count() / grid.y;
      dim3 g2(cap >= 1 && grid.x > cap ? cap : grid.x, grid.y, 1);
      if (vec_out)
        gemm_ws_kernel<true><<<g2, WS_THREADS, WS_SMEM, (cudaStream_t)stream>>>(A, sai, C, ldc, bias, (int)I, (int)J, (int)P, flags, bsplit);
      else
        gemm_ws_kernel<false><<<g2, WS_THREADS, WS_SMEM, (cudaStream_t)stream>>>(A, sai, C, ldc, bias, (int)I, (int)J, (int)P, flags, bsplit);
      UCN_LAUNCH_CHECK("gemm_ws_kernel");
      return UCN_OK;
    }
    if (!(g_gemm_dbg & 32)) {
      const int sms = ucn_sm_count();
      const unsigned cap = (unsigned)(2 * sms) / grid.y / grid.z;             // two CTAs per SM are resident
      if (cap >= 1 && grid.x > cap) grid.x = cap;
    }
    gemm_tc_kernel<<<grid, 256, TC_SMEM, (cudaStream_t)stream>>>(A, sai, sap, B, sbp, sbj, C, ldc, bias, (int)I, (int)J, P, p_chunk, flags, a_vec, b_vec, (g_gemm_dbg & 16) ? nullptr : bsplit);
    UCN_LAUNCH_CHECK("gemm_tc_kernel");
    return UCN_OK;
  }
  gemm_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, sai, sap, B, sbp, sbj, C, ldc, bias, (int)I, (int)J, P, p_chunk, flags);
  UCN_LAUNCH_CHECK("gemm_kernel");
  return UCN_OK;
}

}  // namespace

// debug aid (A/B timing of the staging variants of gemm_tc_kernel; every variant computes the same product)
extern "C" int ucn_debug_gemm_flags(int f) { g_gemm_dbg = f; return 0; }
// 0: nn.Linear products on tcgen05 (split-bf16, fp32 accumulate) where the shape fills a tile; 1: fp32 CUDA cores only
extern "C" int ucn_train_gemm_mode(int mode) {
  UCN_CHECK_ARG(mode == 0 || mode == 1, "ucn_train_gemm_mode: 0 (tensor cores) or 1 (SIMT fp32)");
  g_gemm_mode = mode;
  return UCN_OK;
}
// Y[m][n] (+)= sum_k X[m*ldx + k] W[n*ldw + k] + b[n]     (nn.Linear forward; b may be NULL; flags: 1 accumulate, 2 ReLU)
extern "C" int ucn_linear_fwd(const float *X, int64_t ldx, const float *W, int64_t ldw, const float *b, int64_t M, int K,
                              int N, float *Y, int64_t ldy, int flags, void *stream) {
  UCN_CHECK_ARG(X && W && Y, "ucn_linear_fwd: null argument");
  UCN_CHECK_ARG((flags & ~3) == 0, "ucn_linear_fwd: flags");
  if (N <= 4 && M > 0) {
    if (K % 128 == 0 && al16(X) && al16(W) && ldx % 4 == 0 && ldw % 4 == 0)
      skinny_fwd_v4_kernel<<<nblk(M, 8), 256, 0, (cudaStream_t)stream>>>(X, ldx, W, ldw, b, M, K, N, Y, ldy, flags);
    else
      skinny_fwd_kernel<<<nblk(M, 8), 256, 0, (cudaStream_t)stream>>>(X, ldx, W, ldw, b, M, K, N, Y, ldy, flags);
    UCN_LAUNCH_CHECK("skinny_fwd_kernel");
    return UCN_OK;
  }
  return launch_gemm(X, ldx, 1, W, 1, ldw, Y, ldy, b, M, N, K, flags, 1, stream);
}
// dX[m][k] (+)= sum_n dY[m*ldy + n] W[n*ldw + k]           (nn.Linear grad-input; flags: 1 accumulate)
extern "C" int ucn_linear_bwd_input(const float *dY, int64_t ldy, const float *W, int64_t ldw, int64_t M, int K, int N,
                                    float *dX, int64_t ldx, int flags, void *stream) {
  UCN_CHECK_ARG(dY && W && dX, "ucn_linear_bwd_input: null argument");
  UCN_CHECK_ARG((flags & ~1) == 0, "ucn_linear_bwd_input: flags");
  if (N <= 4 && M > 0) {
    if (K % 4 == 0 && al16(W) && al16(dX) && ldw % 4 == 0 && ldx % 4 == 0)
      skinny_bwd_input_v4_kernel<<<nblk(M * (K / 4), 256), 256, 0, (cudaStream_t)stream>>>(dY, ldy, W, ldw, M, K, N, dX, ldx, flags);
    else
      skinny_bwd_input_kernel<<<nblk(M * K, 256), 256, 0, (cudaStream_t)stream>>>(dY, ldy, W, ldw, M, K, N, dX, ldx, flags);
    UCN_LAUNCH_CHECK("skinny_bwd_input_kernel");
    return UCN_OK;
  }
  return launch_gemm(dY, ldy, 1, W, ldw, 1, dX, ldx, nullptr, M, K, N, flags, 1, stream);
}
// dW[n*lddw + k] += sum_m dY[m*ldy + n] X[m*ldx + k],  db[n] += sum_m dY[m*ldy + n]   (grad-weight; db may be NULL)
extern "C" int ucn_linear_bwd_weight(const float *dY, int64_t ldy, const float *X, int64_t ldx, int64_t M, int K, int N,
                                     float *dW, int64_t lddw, float *db, void *stream) {
  UCN_CHECK_ARG(dY && X && dW, "ucn_linear_bwd_weight: null argument");
  if (M <= 0) return UCN_OK;
  if (N <= 4) {
    skinny_bwd_weight_kernel<<<dim3(nblk(M, 256), nblk(K, 128)), 128, 0, (cudaStream_t)stream>>>(dY, ldy, X, ldx, M, K, N, 256, dW, lddw);
    UCN_LAUNCH_CHECK("skinny_bwd_weight_kernel");
    if (db) {
      colsum_kernel<<<dim3(nblk(N, 32), nblk(M, 512)), 256, 0, (cudaStream_t)stream>>>(dY, ldy, M, N, 512, db);
      UCN_LAUNCH_CHECK("colsum_kernel");
    }
    return UCN_OK;
  }
  const bool tc = g_gemm_mode == 0 && N >= 64 && K >= 16 && M >= 16;
  const int tile = tc ? 128 : BT;
  const long long tiles = (long long)((N + tile - 1) / tile) * ((K + tile - 1) / tile);
  int splits = (int)((2 * ucn_sm_count() + tiles - 1) / tiles);
  const long long max_splits = (M + 255) / 256;
  if (splits > max_splits) splits = (int)max_splits;
  if (g_gemm_dbg >> 12) splits = g_gemm_dbg >> 12;    // timing switch (tools/ubench_gemm.py)
  int rc;
  // long reductions over 16-byte-aligned dY rows: the MN-major pipeline (X may have any row pitch: 4-byte copies then)
  // (measured: with 4-byte copies of an unaligned X the pipeline is SLOWER than gemm_tc_kernel, 279 vs 195 us at K = 191 --
  // those products, K2's 61-float feature rows in this path, stay on the old kernel unless the debug flag 256 forces them here)
  const bool x_vec = (reinterpret_cast<uintptr_t>(X) & 15) == 0 && (ldx & 3) == 0;
  if (tc && M >= 4096 && !(g_gemm_dbg & 128) && (x_vec || (g_gemm_dbg & 256)) && (reinterpret_cast<uintptr_t>(dY) & 15) == 0 && (ldy & 3) == 0) {
    static int attr_w[kUcnMaxDevices], attr_w2[kUcnMaxDevices];
    cudaError_t e = x_vec ? ucn_ensure_dyn_smem(gemm_wsw_kernel<true>, (int)WW_SMEM, attr_w)
                          : ucn_ensure_dyn_smem(gemm_wsw_kernel<false>, (int)WW_SMEM, attr_w2);
    if (e != cudaSuccess) return ucn_cuda_status(e, "gemm_wsw smem attribute");
    const long long n_slab = (M + TC_SLAB - 1) / TC_SLAB;
    long long ctas = ucn_sm_count() / tiles;
    if (ctas < 1) ctas = 1;
    const int per = (int)((n_slab + ctas - 1) / ctas);
    dim3 grid((unsigned)((n_slab + per - 1) / per), (unsigned)tiles);
    if (x_vec)
      gemm_wsw_kernel<true><<<grid, WS_THREADS, WW_SMEM, (cudaStream_t)stream>>>(dY, ldy, X, ldx, dW, lddw, N, K, M, per);
    else
      gemm_wsw_kernel<false><<<grid, WS_THREADS, WW_SMEM, (cudaStream_t)stream>>>(dY, ldy, X, ldx, dW, lddw, N, K, M, per);
    UCN_LAUNCH_CHECK("gemm_wsw_kernel");
    rc = UCN_OK;
  } else {
    rc = launch_gemm(dY, 1, ldy, X, ldx, 1, dW, lddw, nullptr, N, K, M, G_ATOMIC, splits < 1 ? 1 : splits, stream);
  }
  if (rc != UCN_OK || !db) return rc;
  const long long chunks = (M + 511) / 512;
  colsum_kernel<<<dim3(nblk(N, 32), (unsigned)chunks), 256, 0, (cudaStream_t)stream>>>(dY, ldy, M, N, 512, db);
  UCN_LAUNCH_CHECK("colsum_kernel");
  return UCN_OK;
}
// ph[:, 0:63] = PE_10(ndc), ftv[:, 128:155] = PE_4(angle[ray])      (network/models.py:50-54; renderer.py:88-100)
extern "C" int ucn_train_prep(const float *ndc, const float *angle, int S, int64_t N, float *ph, int64_t ldph, float *ftv,
                              int64_t ldftv, void *stream) {
  UCN_CHECK_ARG(ndc && angle && ph && ftv && S > 0, "ucn_train_prep: bad argument");
  UCN_CHECK_ARG(ldph >= 63 && ldftv >= 155, "ucn_train_prep: leading dimensions");
  if (N <= 0) return UCN_OK;
  prep_kernel<<<nblk(N, PREP_T), PREP_T, 0, (cudaStream_t)stream>>>(ndc, angle, S, N, ph, ldph, ftv, ldftv);
  UCN_LAUNCH_CHECK("prep_kernel");
  return UCN_OK;
}
// h = relu(pre * mod)                                              (network/models.py:154-155)
extern "C" int ucn_film_relu_fwd(const float *pre, const float *mod, float *h, int64_t ldh, int64_t M, void *stream) {
  UCN_CHECK_ARG(pre && mod && h, "ucn_film_relu_fwd: null argument");
  if (M <= 0) return UCN_OK;
  if (al16(pre) && al16(mod)) {
    if (al16(h) && ldh % 4 == 0)
      film_relu_fwd_v4_kernel<true><<<nblk(M * 32, 256), 256, 0, (cudaStream_t)stream>>>(pre, mod, h, ldh, M);
    else
      film_relu_fwd_v4_kernel<false><<<nblk(M * 32, 256), 256, 0, (cudaStream_t)stream>>>(pre, mod, h, ldh, M);
  } else {
    film_relu_fwd_kernel<<<nblk(M * 128, 256), 256, 0, (cudaStream_t)stream>>>(pre, mod, h, ldh, M);
  }
  UCN_LAUNCH_CHECK("film_relu_fwd_kernel");
  return UCN_OK;
}
extern "C" int ucn_film_relu_bwd(const float *dh, int64_t lddh, const float *pre, const float *mod, float *dpre,
                                 float *dmod, int accumulate, float *db, int64_t M, void *stream) {
  UCN_CHECK_ARG(dh && pre && mod && dpre && dmod, "ucn_film_relu_bwd: null argument");
  if (M <= 0) return UCN_OK;
  if (al16(pre) && al16(mod) && al16(dpre) && al16(dmod) && (!db || al16(db))) {
    if (al16(dh) && lddh % 4 == 0)
      film_relu_bwd_v4_kernel<true><<<nblk(M, 64), 128, 0, (cudaStream_t)stream>>>(dh, lddh, pre, mod, dpre, dmod, accumulate, db, M);
    else
      film_relu_bwd_v4_kernel<false><<<nblk(M, 64), 128, 0, (cudaStream_t)stream>>>(dh, lddh, pre, mod, dpre, dmod, accumulate, db, M);
  } else {
    film_relu_bwd_kernel<<<nblk(M, 64), 128, 0, (cudaStream_t)stream>>>(dh, lddh, pre, mod, dpre, dmod, accumulate, db, M);
  }
  UCN_LAUNCH_CHECK("film_relu_bwd_kernel");
  return UCN_OK;
}
// o = a * b  (h * feats_bias, network/models.py:164) and its backward
extern "C" int ucn_mul_fwd(const float *a, int64_t lda, const float *b, float *o, int64_t M, void *stream) {
  UCN_CHECK_ARG(a && b && o, "ucn_mul_fwd: null argument");
  if (M <= 0) return UCN_OK;
  if (al16(b) && al16(o)) {
    if (al16(a) && lda % 4 == 0) mul_fwd_v4_kernel<true><<<nblk(M * 32, 256), 256, 0, (cudaStream_t)stream>>>(a, lda, b, o, M);
    else mul_fwd_v4_kernel<false><<<nblk(M * 32, 256), 256, 0, (cudaStream_t)stream>>>(a, lda, b, o, M);
  } else {
    mul_fwd_kernel<<<nblk(M * 128, 256), 256, 0, (cudaStream_t)stream>>>(a, lda, b, o, M);
  }
  UCN_LAUNCH_CHECK("mul_fwd_kernel");
  return UCN_OK;
}
extern "C" int ucn_mul_bwd(const float *d_o, const float *a, int64_t lda, const float *b, float *da, int accumulate,
                           float *db, int64_t M, void *stream) {
  UCN_CHECK_ARG(d_o && a && b && da && db, "ucn_mul_bwd: null argument");
  if (M <= 0) return UCN_OK;
  if (al16(d_o) && al16(b) && al16(da) && al16(db)) {
    if (al16(a) && lda % 4 == 0)
      mul_bwd_v4_kernel<true><<<nblk(M * 32, 256), 256, 0, (cudaStream_t)stream>>>(d_o, a, lda, b, da, accumulate, db, M);
    else
      mul_bwd_v4_kernel<false><<<nblk(M * 32, 256), 256, 0, (cudaStream_t)stream>>>(d_o, a, lda, b, da, accumulate, db, M);
  } else {
    mul_bwd_kernel<<<nblk(M * 128, 256), 256, 0, (cudaStream_t)stream>>>(d_o, a, lda, b, da, accumulate, db, M);
  }
  UCN_LAUNCH_CHECK("mul_bwd_kernel");
  return UCN_OK;
}
extern "C" int ucn_relu_bwd(const float *dy, const float *y, float *dx, int64_t n, void *stream) {
  UCN_CHECK_ARG(dy && y && dx, "ucn_relu_bwd: null argument");
  if (n <= 0) return UCN_OK;
  if (n % 4 == 0 && al16(dy) && al16(y) && al16(dx))
    relu_bwd_v4_kernel<<<nblk(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(dy, y, dx, n / 4);
  else
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
