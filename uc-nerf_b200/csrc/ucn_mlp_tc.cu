// ucn_mlp_tc.cu -- K3 (production mode): positional encoding + uncertainty-conditioned MLP on tcgen05
// tensor cores (bf16 operands, fp32 accumulators in TMEM), one persistent CTA per SM.
//
// replaces (reference): network/models.py:50-54 Embedder.embed, network/renderer.py:78-106
//   run_network_mvs (+batchify), network/models.py:138-184 BaseAdapt_Renderer.forward -- 13 cuBLAS sgemm
//   launches + ~40 elementwise kernels per 1024-ray chunk in the reference.
//
// Structure (per CTA, 384 threads):
//   warp 0      weight producer: streams the 21 weight K-blocks of the network (~290 KB bf16, L2 resident)
//               through a 7-stage shared-memory ring with cp.async.bulk + mbarrier complete_tx
//   warp 1      MMA issuer: one thread issues every tcgen05.mma (M=128, N=128 or 16, K=16)
//   warp 2      TMEM allocator (512 columns)
//   warps 4-7   epilogue warpgroup of tile slot 0   (thread = sample row = TMEM lane)
//   warps 8-11  epilogue warpgroup of tile slot 1
// Two 128-sample tiles ("slots") are in flight and share every weight block: while the tensor core runs
// round r of one slot, the other slot's warpgroup does the epilogue of its round (TMEM -> registers ->
// bias/FiLM/ReLU -> bf16 -> shared memory A operand of the next round).  Each slot owns 256 TMEM columns:
// [mod | acc].  A operands use the UMMA no-swizzle K-major core-matrix layout, element (row,k) at
// ((k/8)*128 + row)*16 + (k%8)*2 bytes: LBO = 2048 B between K-chunks, SBO = 128 B between 8-row groups.
// Slot memory = 24 K-chunks: P region (chunks 0-7: PE(ndc), later fb inputs, later PE(view dir)) and
// H region (chunks 8-23: film inputs, then the hidden state, updated in place by every epilogue).
//
// Rounds per tile (MMA -> epilogue):
//   R0 film(->mod), L0   R1-R4 L1-L4   R5 L5 (K=192 = [PE|h], the skip concat)   R6 fb(->mod), base heads
//   R7 feature_linear    R8 views|view_confi on [feature|PE(dir)] (K=160)          R9 adapt heads -> blend
//
// Roofline: tensor-bound.  Algorithmic FLOPs/sample = 2*(138112+1536V); issued MMA
// FLOPs/sample (padded) = 2*128*(48+64+512+192+32+128+160) + 2*16*256.
#include <cuda_bf16.h>

#include "ucn_common.cuh"
#include "ucn_weights.cuh"

namespace {

constexpr int kThreads = 384;
constexpr int NST = 7;
constexpr uint32_t CHUNK = 2048;            // bytes of one K-chunk (8 bf16) x 128 rows
constexpr uint32_t STAGE = 8 * CHUNK;       // 16 KB weight ring stage
constexpr uint32_t SLOT = 24 * CHUNK;       // 48 KB per tile slot
constexpr int NBLK = 21;
constexpr int NROUND = 10;
__device__ __constant__ int kRoundBegin[NROUND + 1] = {0, 2, 4, 6, 8, 10, 13, 15, 17, 20, 21};

struct SmemLayout {
  static constexpr uint32_t slots = 0;
  static constexpr uint32_t ring = 2 * SLOT;
  static constexpr uint32_t bias = ring + NST * STAGE;          // 8*128 + 8 floats
  static constexpr uint32_t bars = bias + (8 * 128 + 8) * 4;    // in_ready[2], acc_full[2], w_full[NST], w_empty[NST]
  static constexpr uint32_t tmem_ptr = bars + (4 + 2 * NST) * 8;
  static constexpr uint32_t total = tmem_ptr + 16;
};

struct Blk { uint32_t goff, bytes; int nch, a0, N, dcol, fresh; };

__device__ __forceinline__ Blk block_info(int b, const ucn::BF16Layout &L) {
  Blk k;
  k.N = 128; k.nch = 8; k.dcol = 128; k.fresh = 1; k.a0 = 8; k.goff = 0;
  if (b == 0) { k.goff = (uint32_t)L.film; k.nch = L.Kfilm / 8; k.a0 = 8; k.dcol = 0; }
  else if (b == 1) { k.goff = (uint32_t)L.pts[0]; k.a0 = 0; }
  else if (b < 10) { const int l = 1 + (b - 2) / 2, half = (b - 2) & 1; k.goff = (uint32_t)L.pts[l] + half * STAGE; k.a0 = 8 + 8 * half; k.fresh = !half; }
  else if (b < 13) { const int j = b - 10; k.goff = (uint32_t)L.pts[5] + j * STAGE; k.a0 = 8 * j; k.fresh = j == 0; }
  else if (b == 13) { k.goff = (uint32_t)L.fb; k.nch = L.Kfb / 8; k.a0 = 0; k.dcol = 0; }
  else if (b == 14) { k.goff = (uint32_t)L.base; k.nch = 16; k.N = 16; }
  else if (b < 17) { const int j = b - 15; k.goff = (uint32_t)L.feat + j * STAGE; k.a0 = 8 + 8 * j; k.fresh = j == 0; }
  else if (b < 19) { const int j = b - 17; k.goff = (uint32_t)L.vv + j * STAGE; k.a0 = 8 + 8 * j; k.dcol = 0; k.fresh = j == 0; }
  else if (b == 19) { k.goff = (uint32_t)L.vv + 2 * STAGE; k.nch = 4; k.a0 = 0; k.dcol = 0; k.fresh = 0; }
  else { k.goff = (uint32_t)L.adapt; k.nch = 16; k.N = 16; }
  k.bytes = (uint32_t)k.nch * k.N * 16;
  return k;
}

// ---- PTX wrappers -----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
// K-major, no swizzle: LBO = bytes between the two K-chunks of one K=16 step, SBO = bytes between 8-row groups
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
__device__ __forceinline__ uint32_t umma_idesc(int N) {      // bf16 x bf16 -> fp32, A and B K-major, M = 128
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
               "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
               "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                 "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                 "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
               : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&r)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&v);
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

// store 8 consecutive K values (one chunk row) of this thread's row
__device__ __forceinline__ void store_chunk(uint32_t row_addr, int chunk, const float *v) {
  sts128(row_addr + chunk * CHUNK, pack2(v[0], v[1]), pack2(v[2], v[3]), pack2(v[4], v[5]), pack2(v[6], v[7]));
}

// Epilogue of a 128-wide round: out = act((acc + bias) * mod) -> bf16 -> H region (in place), 32 columns at a time.
//   MODE 0: relu((acc + b) * mod)      pts layers        (models.py:154-155)
//   MODE 1: acc + b                    feature_linear    (models.py:164)
//   MODE 2: relu(acc + b)              views / view_confi hidden (models.py:168-173)
//   MODE 3: h(smem) * acc              h * feats_bias    (models.py:164), bias already folded into the MMA
template <int MODE>
__device__ __forceinline__ void epi128(uint32_t t_src, uint32_t t_mod, const float *__restrict__ bias, uint32_t h_row) {
#pragma unroll 1
  for (int c0 = 0; c0 < 128; c0 += 32) {
    uint32_t a[32], m[32];
    tmem_ld32(t_src + c0, a);
    if (MODE == 0) tmem_ld32(t_mod + c0, m);
    tmem_wait_ld();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v[8];
      uint4 old;
      if (MODE == 3) old = lds128(h_row + (c0 / 8 + j) * CHUNK);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float x = __uint_as_float(a[8 * j + i]);
        if (MODE == 0) v[i] = fmaxf((x + bias[c0 + 8 * j + i]) * __uint_as_float(m[8 * j + i]), 0.0f);
        if (MODE == 1) v[i] = x + bias[c0 + 8 * j + i];
        if (MODE == 2) v[i] = fmaxf(x + bias[c0 + 8 * j + i], 0.0f);
        if (MODE == 3) {
          const uint32_t w = i < 2 ? old.x : (i < 4 ? old.y : (i < 6 ? old.z : old.w));
          v[i] = ((i & 1) ? bf_hi(w) : bf_lo(w)) * x;
        }
      }
      store_chunk(h_row, c0 / 8 + j, v);
    }
  }
}

__global__ void __launch_bounds__(kThreads, 1)
mlp_tc_kernel(const unsigned char *__restrict__ wpack, int V, const float *__restrict__ ndc_pts,
              const float *__restrict__ feat, const float *__restrict__ angle, long long N, int S,
              float *__restrict__ raw) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const ucn::BF16Layout L(V);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar_in = sbase + SmemLayout::bars, bar_acc = bar_in + 16, bar_wfull = bar_in + 32,
                 bar_wempty = bar_wfull + NST * 8;
  float *bias_s = reinterpret_cast<float *>(smem + SmemLayout::bias);
  volatile uint32_t *tmem_ptr_s = reinterpret_cast<volatile uint32_t *>(smem + SmemLayout::tmem_ptr);
  const int F = 24 + 12 * V + 1, Kfilm_real = 24 + 4 * V, Kfb_real = 8 * V;

  // ---- one-time setup ---------------------------------------------------------------------------
  {
    const float *bsrc = reinterpret_cast<const float *>(wpack + L.bias);
    for (int i = threadIdx.x; i < 8 * 128 + 8; i += kThreads) {
      float v;
      if (i < 6 * 128) v = bsrc[L.b_pts[0] + i];                       // b_pts[0..5] are contiguous
      else if (i < 7 * 128) v = bsrc[L.b_feat + (i - 6 * 128)];
      else if (i < 8 * 128) v = bsrc[L.b_vv + (i - 7 * 128)];
      else if (i < 8 * 128 + 4) v = bsrc[L.b_base + (i - 8 * 128)];
      else v = bsrc[L.b_adapt + (i - 8 * 128 - 4)];
      bias_s[i] = v;
    }
  }
  if (threadIdx.x == 0) {
    mbar_init(bar_in, 128); mbar_init(bar_in + 8, 128);
    mbar_init(bar_acc, 1); mbar_init(bar_acc + 8, 1);
    for (int i = 0; i < NST; ++i) { mbar_init(bar_wfull + 8 * i, 1); mbar_init(bar_wempty + 8 * i, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + SmemLayout::tmem_ptr), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_s;

  const long long ntiles = (N + 127) / 128;
  const long long npairs = (ntiles + 1) / 2;

  if (warp == 0) {
    // ================= weight producer ===========================================================
    if (lane == 0) {
      uint32_t cnt = 0;
      for (long long p = blockIdx.x; p < npairs; p += gridDim.x) {
        for (int b = 0; b < NBLK; ++b, ++cnt) {
          const Blk k = block_info(b, L);
          const uint32_t st = cnt % NST, ph = (cnt / NST) & 1;
          mbar_wait(bar_wempty + 8 * st, ph ^ 1);
          mbar_expect_tx(bar_wfull + 8 * st, k.bytes);
          bulk_g2s(sbase + SmemLayout::ring + st * STAGE, wpack + k.goff, k.bytes, bar_wfull + 8 * st);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================================================================
    if (lane == 0) {
      uint32_t cnt = 0, ph_in[2] = {0, 0};
      for (long long p = blockIdx.x; p < npairs; p += gridDim.x) {
        const int nslots = (2 * p + 1 < ntiles) ? 2 : 1;
        for (int r = 0; r < NROUND; ++r) {
          const int b0 = kRoundBegin[r], b1 = kRoundBegin[r + 1];
          for (int s = 0; s < nslots; ++s) {
            mbar_wait(bar_in + 8 * s, ph_in[s]);
            ph_in[s] ^= 1;
            tc_fence_after();
            const uint32_t slot_addr = sbase + SmemLayout::slots + s * SLOT;
            for (int b = b0; b < b1; ++b) {
              const Blk k = block_info(b, L);
              const uint32_t c = cnt + (b - b0), st = c % NST;
              if (s == 0) { mbar_wait(bar_wfull + 8 * st, (c / NST) & 1); tc_fence_after(); }
              const uint32_t w_addr = sbase + SmemLayout::ring + st * STAGE;
              const uint32_t d = tmem_base + s * 256 + k.dcol;
              const uint32_t idesc = umma_idesc(k.N);
              for (int ks = 0; ks < k.nch / 2; ++ks) {
                const uint64_t ad = umma_desc(slot_addr + (k.a0 + 2 * ks) * CHUNK, CHUNK, 128);
                const uint64_t bd = umma_desc(w_addr + (2 * ks) * k.N * 16, k.N * 16, 128);
                umma_bf16(d, ad, bd, idesc, (k.fresh && ks == 0) ? 0u : 1u);
              }
            }
            umma_commit(bar_acc + 8 * s);
          }
          for (int b = b0; b < b1; ++b) umma_commit(bar_wempty + 8 * ((cnt + (b - b0)) % NST));
          cnt += b1 - b0;
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ================= epilogue warpgroups =========================================================
    const int s = (warp - 4) >> 2;                               // tile slot
    const int row = ((warp & 3) << 5) | lane;                    // sample row == TMEM lane
    const uint32_t slot_addr = sbase + SmemLayout::slots + s * SLOT;
    const uint32_t p_row = slot_addr + row * 16;                 // chunk c of this row: p_row + c*CHUNK
    const uint32_t h_row = p_row + 8 * CHUNK;
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) << 5) << 16) + s * 256;
    const uint32_t t_mod = t_lane, t_acc = t_lane + 128;
    uint32_t ph_acc = 0;
    for (long long p = blockIdx.x; p < npairs; p += gridDim.x) {
      const long long tile = 2 * p + s;
      if (tile >= ntiles) break;
      const long long q = tile * 128 + row;
      const bool valid = q < N;
      const float *frow = feat + (valid ? q : 0) * F;
      // ---- stage inputs: PE(ndc) -> P region, [film inputs | 1] -> H region chunks 0.. ----------------
      float u;
      {
        float x[3] = {0.f, 0.f, 0.f};
        if (valid) { x[0] = ndc_pts[3 * q]; x[1] = ndc_pts[3 * q + 1]; x[2] = ndc_pts[3 * q + 2]; }
        float pe[64];
        pe[0] = x[0]; pe[1] = x[1]; pe[2] = x[2]; pe[63] = 0.0f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float sn, cs;
          sincosf(x[c], &sn, &cs);
#pragma unroll
          for (int l = 0; l < 10; ++l) {                         // double-angle recurrence, fp32
            pe[3 + 3 * l + c] = sn;
            pe[33 + 3 * l + c] = cs;
            const float s2 = 2.0f * sn * cs, c2 = 1.0f - 2.0f * sn * sn;
            sn = s2; cs = c2;
          }
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) store_chunk(p_row, c, pe + 8 * c);
        for (int c = 0; c < L.Kfilm / 8; ++c) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int k = 8 * c + i;
            v[i] = k < Kfilm_real ? (valid ? __ldg(frow + k) : 0.0f) : (k == Kfilm_real ? 1.0f : 0.0f);
          }
          store_chunk(h_row, c, v);
        }
        u = 1.0f - (valid ? __ldg(frow + F - 1) : 0.0f);         // models.py:149
      }
      tc_fence_before();
      fence_async_smem();
      mbar_arrive(bar_in + 8 * s);

      float base4[4];
#pragma unroll 1
      for (int r = 0; r < NROUND; ++r) {
        mbar_wait(bar_acc + 8 * s, ph_acc);
        ph_acc ^= 1;
        tc_fence_after();
        if (r < 5) {
          epi128<0>(t_acc, t_mod, bias_s + 128 * r, h_row);
        } else if (r == 5) {
          epi128<0>(t_acc, t_mod, bias_s + 128 * 5, h_row);
          for (int c = 0; c < L.Kfb / 8; ++c) {                  // [img feats | 1] -> P region (models.py:151)
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int k = 8 * c + i;
              v[i] = k < Kfb_real ? (valid ? __ldg(frow + Kfilm_real + k) : 0.0f) : (k == Kfb_real ? 1.0f : 0.0f);
            }
            store_chunk(p_row, c, v);
          }
        } else if (r == 6) {
          uint32_t b[4];
          tmem_ld4(t_acc, b);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 4; ++i) base4[i] = __uint_as_float(b[i]) + bias_s[8 * 128 + i];
          epi128<3>(t_mod, 0, nullptr, h_row);
          float a[3] = {0.f, 0.f, 0.f};                          // PE(view dir) -> P chunks 0..3
          if (valid) { const long long ray = q / S; a[0] = angle[3 * ray]; a[1] = angle[3 * ray + 1]; a[2] = angle[3 * ray + 2]; }
          float pe[32];
          pe[0] = a[0]; pe[1] = a[1]; pe[2] = a[2];
#pragma unroll
          for (int i = 27; i < 32; ++i) pe[i] = 0.0f;
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            float sn, cs;
            sincosf(a[c], &sn, &cs);
#pragma unroll
            for (int l = 0; l < 4; ++l) {
              pe[3 + 3 * l + c] = sn;
              pe[15 + 3 * l + c] = cs;
              const float s2 = 2.0f * sn * cs, c2 = 1.0f - 2.0f * sn * sn;
              sn = s2; cs = c2;
            }
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) store_chunk(p_row, c, pe + 8 * c);
        } else if (r == 7) {
          epi128<1>(t_acc, 0, bias_s + 128 * 6, h_row);
        } else if (r == 8) {
          epi128<2>(t_mod, 0, bias_s + 128 * 7, h_row);
        } else {
          uint32_t b[4];
          tmem_ld4(t_acc, b);
          tmem_wait_ld();
          const float cf = 1.0f - u;                             // models.py:177-180
          float4 o;
          float ad[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) ad[i] = __uint_as_float(b[i]) + bias_s[8 * 128 + 4 + i];
          o.x = 1.0f / (1.0f + __expf(-(base4[0] * cf + ad[0] * u)));
          o.y = 1.0f / (1.0f + __expf(-(base4[1] * cf + ad[1] * u)));
          o.z = 1.0f / (1.0f + __expf(-(base4[2] * cf + ad[2] * u)));
          o.w = fmaxf(ad[3] * cf + base4[3] * u, 0.0f);
          if (valid) reinterpret_cast<float4 *>(raw)[q] = o;
        }
        if (r < NROUND - 1) {
          tc_fence_before();
          fence_async_smem();
          mbar_arrive(bar_in + 8 * s);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace

int ucn_mlp_forward_tc(const ucn_weights *w, const float *ndc_pts, const float *feat, const float *angle, int R, int S,
                       float *raw, void *stream) {
  UCN_CHECK_ARG(w->bf16, "ucn_mlp_forward: bf16 weight pack missing");
  UCN_CHECK_ARG(w->V <= 7, "ucn_mlp_forward: BF16_TC mode supports at most 7 source views");
  const long long N = (long long)R * S;
  static int attr[kUcnMaxDevices];
  {
    cudaError_t e = ucn_ensure_dyn_smem(mlp_tc_kernel, (int)SmemLayout::total, attr);
    if (e != cudaSuccess) return ucn_cuda_status(e, "mlp_tc smem attribute");
  }
  const int sms = ucn_sm_count();
  const long long npairs = ((N + 127) / 128 + 1) / 2;
  const unsigned blocks = (unsigned)(npairs < sms ? npairs : sms);
  mlp_tc_kernel<<<blocks, kThreads, SmemLayout::total, (cudaStream_t)stream>>>(
      reinterpret_cast<const unsigned char *>(w->bf16), w->V, ndc_pts, feat, angle, N, S, raw);
  UCN_LAUNCH_CHECK("mlp_tc_kernel");
  return UCN_OK;
}
