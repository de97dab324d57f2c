// ucn_mlp_x3.cu -- K3, production mode: positional encoding + uncertainty-conditioned MLP on tcgen05 tensor
// cores with SPLIT-bf16 operands (hi + lo, three MMAs per K step, fp32 accumulators in TMEM).
//
// replaces (reference): network/models.py:50-54 Embedder.embed, network/renderer.py:78-106 run_network_mvs
//   (+batchify), network/models.py:138-184 BaseAdapt_Renderer.forward.
//
// Why split operands: plain bf16 x bf16 through six FiLM-modulated layers gives ~2e-2 worst-case RGB error
// on random-init weights (measured, and reproduced by tools/precision_study.py); the parity gate is 1e-3.
// a*w ~= a_hi*w_hi + a_lo*w_hi + a_hi*w_lo carries ~16 significant bits per operand (measured 6e-5 RGB).
//
// One persistent CTA per SM, 640 threads:
//   warp 0      weight producer: streams the split weight pack (~550 KB, L2 resident) through a shared-memory
//               ring of 16 KB stages with cp.async.bulk + mbarrier complete_tx
//   warp 1      MMA issuer (one thread issues every tcgen05.mma, M=128 N=128 K=16)
//   warp 2      TMEM allocator (512 columns)
//   warps 4-19  four epilogue warpgroups: tile slot s = g/2, column half hf = g%2 (thread = sample row = TMEM lane)
// Two 128-sample tiles ("slots") are in flight and share every weight stage; while the tensor core runs a round
// of one slot, the other slot's two warpgroups run their epilogue (TMEM -> registers -> bias/FiLM/ReLU ->
// hi/lo bf16 -> shared-memory A operand of the next round).  Each slot owns 256 TMEM columns [mod | acc] and
// 64 KB of shared memory: H_hi | H_lo, 16 K-chunks each in the UMMA no-swizzle K-major core-matrix layout
// (element (row,k) at (k/8)*2048 + row*16 + (k%8)*2 bytes).
//
// Rounds per tile (MMA -> epilogue):
//   R0  film(H[8..]) -> mod, L0(PE in H[0..7]) -> acc      relu((acc+b)*mod) -> H
//   R1-R4  L1..L4                                          same
//   R5  L5, hidden part (K=128)                            warpgroup 0 re-creates PE in H[0..7]
//   R6  L5, PE part (K=64, accumulates)                    relu((acc+b)*mod) -> H; base heads on CUDA cores;
//                                                          [img feats | 1] hi/lo -> TMEM (acc columns 64..127)
//   R7  fb (A operand from TMEM) -> mod                    h * fb -> H
//   R8  feature_linear -> acc                              acc + b -> H; per-ray view-direction term -> smem
//   R9  views|view_confi hidden (K=128) -> acc             relu(acc + b + per-ray term); adapt heads on CUDA
//                                                          cores; uncertainty blend -> raw
// The 27 view-direction inputs of views/view_confi are constant along a ray, so their contribution is computed
// once per ray on the CUDA cores in fp32 (no K padding on the tensor core); the 4-wide heads are fp32 dot
// products in the epilogue (an N=16 MMA costs as much as N=128: measured 66.8 cycles either way).
//
// Roofline: tensor-bound.  Algorithmic FLOPs/sample = 2*(138112+1536V); issued tensor FLOPs per
// sample = 3 * 2*128*(Kfilm + 64 + 4*128 + 128 + 64 + Kfb + 128 + 128).
#include <string.h>

#include "ucn_common.cuh"
#include "ucn_tc.cuh"
#include "ucn_weights.cuh"

using namespace ucn_tc;

namespace {

constexpr int kThreads = 640;
constexpr uint32_t CHUNK = 2048;             // one K-chunk (8 bf16) x 128 rows
constexpr uint32_t HALF = 16 * CHUNK;        // H_hi -> H_lo distance (32 KB)
constexpr uint32_t SLOT = 2 * HALF;          // 64 KB per tile slot
constexpr uint32_t STAGE = 2 * ucn::X3Layout::kUnit;   // 16 KB ring stage = two K16 units (hi+lo each)
constexpr int NROUND = 10;
constexpr int MAXBLK = 48;

// optional timeline trace of CTA 0 (debug aid for profiling the round pipeline; off unless a buffer is set).
// Each tracer (0 = MMA thread, 1..4 = warpgroup leaders, 5 = producer) owns a private region: no atomics.
__device__ long long *g_trace = nullptr;
__device__ int g_dbg = 0;      // experiment switches for tools/trace_x3.py (timing experiments only; results may be wrong)
constexpr int kTraceCap = 2048;
struct Tracer {
  long long *base; int n;
  __device__ __forceinline__ Tracer(int who) : base(nullptr), n(0) {
    if (g_trace && blockIdx.x == 0) base = g_trace + (size_t)who * kTraceCap * 2;
  }
  __device__ __forceinline__ void ev(int tag, int r, int who, long long payload = 0) {
    if (base && n < kTraceCap) {
      base[2 * n] = (long long)tag | ((long long)r << 8) | ((long long)who << 16) | (payload << 24);
      base[2 * n + 1] = clock64();
      ++n;
    }
  }
};

enum { BF_FRESH = 1, BF_MOD = 2, BF_TS = 4 };
struct Blk { uint32_t goff; uint8_t units, a0, flags, pad; };   // a0: first A K-chunk (SS) / first K16 unit (TS)
struct Plan {
  Blk blk[MAXBLK];
  uint8_t round_begin[NROUND + 1];
  uint8_t nblk, nst, rpt, pad;
  uint32_t off_f32, off_exch, off_rayb, off_bars, smem_total;   // shared-memory carve-up after the ring
};

// Epilogue of one 64-column half of a 128-wide round, 32 columns at a time.
//   MODE 0: relu((acc + b) * mod) -> H          pts layers                          (models.py:154-155)
//   MODE 1: acc + b -> H                        feature_linear                      (models.py:164)
//   MODE 2: relu(acc + b + ray term), no store  views / view_confi hidden           (models.py:168-173)
//   MODE 3: h(smem) * mod -> H                  h * feats_bias, bias folded in MMA  (models.py:164)
// HEADS: accumulate the 4-wide head  head4 += v[c] * w4[c]  (fp32, CUDA cores) for this thread's 64 columns.
template <int MODE, bool HEADS>
__device__ __forceinline__ void epi64(uint32_t t_acc, uint32_t t_mod, int col0, const float *__restrict__ bias,
                                      const float *__restrict__ rayb, uint32_t h_row, const float *__restrict__ w4,
                                      float (&head)[4]) {
#pragma unroll 1
  for (int c0 = col0; c0 < col0 + 64; c0 += 32) {
    uint32_t a[32], m[32];
    if (MODE != 3) tmem_ld32(t_acc + c0, a);
    if (MODE == 0 || MODE == 3) tmem_ld32(t_mod + c0, m);
    tmem_wait_ld();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c0 + 8 * j;
      float v[8];
      if (MODE == 3) {
        const uint4 oh = lds128(h_row + (c >> 3) * CHUNK), ol = lds128(h_row + HALF + (c >> 3) * CHUNK);
        const uint32_t wh[4] = {oh.x, oh.y, oh.z, oh.w}, wl[4] = {ol.x, ol.y, ol.z, ol.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          v[2 * i] = (bf_lo(wh[i]) + bf_lo(wl[i])) * __uint_as_float(m[8 * j + 2 * i]);
          v[2 * i + 1] = (bf_hi(wh[i]) + bf_hi(wl[i])) * __uint_as_float(m[8 * j + 2 * i + 1]);
        }
      } else {
        const float4 b0 = *reinterpret_cast<const float4 *>(bias + c), b1 = *reinterpret_cast<const float4 *>(bias + c + 4);
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        float rb[8];
        if (MODE == 2) {
          const float4 r0 = *reinterpret_cast<const float4 *>(rayb + c), r1 = *reinterpret_cast<const float4 *>(rayb + c + 4);
          rb[0] = r0.x; rb[1] = r0.y; rb[2] = r0.z; rb[3] = r0.w; rb[4] = r1.x; rb[5] = r1.y; rb[6] = r1.z; rb[7] = r1.w;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float x = __uint_as_float(a[8 * j + i]) + bb[i];
          if (MODE == 0) v[i] = fmaxf(x * __uint_as_float(m[8 * j + i]), 0.0f);
          if (MODE == 1) v[i] = x;
          if (MODE == 2) v[i] = fmaxf(x + rb[i], 0.0f);
        }
      }
      if (HEADS) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 w = *reinterpret_cast<const float4 *>(w4 + 4 * (c + i));
          head[0] = fmaf(v[i], w.x, head[0]); head[1] = fmaf(v[i], w.y, head[1]);
          head[2] = fmaf(v[i], w.z, head[2]); head[3] = fmaf(v[i], w.w, head[3]);
        }
      }
      if (MODE != 2) {
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
        sts128(h_row + (c >> 3) * CHUNK, hi[0], hi[1], hi[2], hi[3]);
        sts128(h_row + HALF + (c >> 3) * CHUNK, lo[0], lo[1], lo[2], lo[3]);
      }
    }
  }
}

// write 8 consecutive K values (one chunk) of this thread's row as hi/lo bf16
__device__ __forceinline__ void store_chunk_x3(uint32_t row_addr, int chunk, const float *v) {
  uint32_t hi[4], lo[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
  sts128(row_addr + chunk * CHUNK, hi[0], hi[1], hi[2], hi[3]);
  sts128(row_addr + HALF + chunk * CHUNK, lo[0], lo[1], lo[2], lo[3]);
}

// PE(x) with L frequencies, reference channel order [x | sin(2^l x) (l-major, xyz-minor) | cos ...]
// (network/models.py:37-54); fp32 sincos at frequencies 2^0 and 2^5, double-angle recurrence in between.
template <int L>
__device__ __forceinline__ void pos_enc(const float (&x)[3], float *pe) {
  pe[0] = x[0]; pe[1] = x[1]; pe[2] = x[2];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float sn, cs;
    sincosf(x[c], &sn, &cs);
#pragma unroll
    for (int l = 0; l < L; ++l) {
      if (l == 5) sincosf(32.0f * x[c], &sn, &cs);      // re-seed: keeps the recurrence error below 2^5 ulp
      pe[3 + 3 * l + c] = sn;
      pe[3 + 3 * L + 3 * l + c] = cs;
      const float s2 = 2.0f * sn * cs, c2 = 1.0f - 2.0f * sn * sn;
      sn = s2; cs = c2;
    }
  }
}


// "touch" load: pulls a line into L1/L2 ahead of its real use (result discarded).  Plain loads issued while the MMAs
// saturate shared memory take 3-4 K cycles to come back (trace: tools/trace_x3.py), so every global input of an
// epilogue is touched one round early, by warps that are idle then.
__device__ __forceinline__ void touch(const float *p) {
  float t;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(t) : "l"(p));
}

__global__ void __launch_bounds__(kThreads, 1)
mlp_x3_kernel(const __grid_constant__ Plan plan, const unsigned char *__restrict__ wpack, int V,
              const float *__restrict__ ndc_pts, const float *__restrict__ feat, long long f_tile_stride,
              long long f_k_stride, long long f_row_stride, const float *__restrict__ angle, long long N, int S,
              float *__restrict__ raw) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const ucn::X3Layout L(V);
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform by construction
  const int nst = plan.nst, rpt = plan.rpt, nblk = plan.nblk;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t ring = sbase + 2 * SLOT;
  const uint32_t bar_in = pin(sbase + plan.off_bars), bar_acc = pin(bar_in + 16), bar_wfull = pin(bar_in + 32),
                 bar_wempty = pin(bar_wfull + 8 * nst), tmem_ptr_addr = bar_wempty + 8 * nst;
  float *f32s = reinterpret_cast<float *>(smem + plan.off_f32);
  float4 *exch_s = reinterpret_cast<float4 *>(smem + plan.off_exch);
  float *rayb_s = reinterpret_cast<float *>(smem + plan.off_rayb);
  __shared__ uint32_t blk_s[MAXBLK], goff_s[MAXBLK];
  __shared__ int rb_s[NROUND + 1];
  const int F = 24 + 12 * V + 1, Kfilm_real = 24 + 4 * V, Kfb_real = 8 * V;
  const float *f32g = reinterpret_cast<const float *>(wpack + L.f32);

  // ---- one-time setup ---------------------------------------------------------------------------
  for (int i = threadIdx.x; i < ucn::X3Layout::kSmemFloats; i += kThreads) f32s[i] = f32g[i];
  if (threadIdx.x < plan.nblk) {
    const Blk k = plan.blk[threadIdx.x];
    blk_s[threadIdx.x] = (uint32_t)k.units | ((uint32_t)k.a0 << 8) | ((uint32_t)k.flags << 16);
    goff_s[threadIdx.x] = k.goff;
  }
  if (threadIdx.x <= NROUND) rb_s[threadIdx.x] = plan.round_begin[threadIdx.x];
  if (threadIdx.x == 0) {
    mbar_init(bar_in, 8); mbar_init(bar_in + 8, 8);
    mbar_init(bar_acc, 1); mbar_init(bar_acc + 8, 1);
    for (int i = 0; i < nst; ++i) { mbar_init(bar_wfull + 8 * i, 1); mbar_init(bar_wempty + 8 * i, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) tmem_alloc512(tmem_ptr_addr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t *>(smem + (tmem_ptr_addr - sbase)), 0);

  const long long ntiles = (N + 127) / 128;
  const long long npairs = (ntiles + 1) / 2;

  if (warp == 0) {
    // ================= weight producer ===========================================================
    if (lane == 0) {
      // A round's weight blocks are loaded once per tile pair and used by both slots (slot 1 runs about half a
      // period behind slot 0); a stage is released when the last slot's MMAs on it have completed.
      uint32_t st = 0, wph = 1;
      for (long long p = blockIdx.x; p < npairs; p += gridDim.x) {
#pragma unroll 1
        for (int b = 0; b < nblk; ++b) {
          const uint32_t bytes = (blk_s[b] & 0xFF) * ucn::X3Layout::kUnit;
          mbar_wait(bar_wempty + 8 * st, wph);
          mbar_expect_tx(bar_wfull + 8 * st, bytes);
          bulk_g2s(ring + st * STAGE, wpack + goff_s[b], bytes, bar_wfull + 8 * st);
          if (++st == (uint32_t)nst) { st = 0; wph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================================================================
    // The whole warp runs this loop convergently (one elected lane issues each MMA), kept to a handful of
    // integer instructions per MMA:
    // every descriptor is "template + (address >> 4)" (LBO/SBO/version bits never change), the ring stage and
    // its phase are running counters, and the block table is read from shared memory.
    {
      const uint32_t idesc = umma_idesc_bf16(128);
      const uint64_t dt = umma_desc(0, CHUNK, 128);               // LBO = 2048 (A chunks and B tiles alike), SBO = 128
      const uint32_t dhi = (uint32_t)(dt >> 32), dlo = (uint32_t)dt;
      auto mk = [&](uint32_t lo) { return ((uint64_t)dhi << 32) | lo; };
      uint32_t st0 = 0, wph0 = 0, ph_in0 = 0, ph_in1 = 0;      // ring position of the current round's first block
      Tracer tr(0);
      if (lane != 0) tr.base = nullptr;
      const uint32_t ring_lo = dlo + (ring >> 4);
      for (long long p = blockIdx.x; p < npairs; p += gridDim.x) {
        const int nslots = (2 * p + 1 < ntiles) ? 2 : 1;
#pragma unroll 1
        for (int r = 0; r < NROUND; ++r) {
          const int b0 = plan.round_begin[r], b1 = plan.round_begin[r + 1];
#pragma unroll 1
          for (int s = 0; s < nslots; ++s) {
            if (s == 0) { mbar_wait(bar_in, ph_in0); ph_in0 ^= 1; } else { mbar_wait(bar_in + 8, ph_in1); ph_in1 ^= 1; }
            tc_fence_after();
            tr.ev(1, r, s);
            long long wwait = 0;
            const uint32_t a_lo = dlo + ((sbase + s * SLOT) >> 4);
            const uint32_t t_slot = tmem_base + s * 256;
            uint32_t st = st0, wph = wph0;
#pragma unroll 1
            for (int b = b0; b < b1; ++b) {
              const uint32_t units = plan.blk[b].units, a0 = plan.blk[b].a0, flags = plan.blk[b].flags;
              const long long w0 = tr.base ? clock64() : 0;
              if (s == 0) mbar_wait(bar_wfull + 8 * st, wph);       // slot 1 reuses the stage slot 0 waited for
              if (tr.base) wwait += clock64() - w0;
              // (no tcgen05 fence here: the stage was filled by the bulk-copy engine, not by tcgen05 ops, and a
              //  fence per block stalls the MMA pipeline -- measured 91 vs 67 cycles per MMA)
              const uint32_t b_lo = ring_lo + st * (STAGE >> 4);
              const uint32_t d = t_slot + ((flags & BF_MOD) ? 0 : 128);
              const uint32_t first = (flags & BF_FRESH) ? 0u : 1u;
              if (flags & BF_TS) {
                const uint32_t ah = t_slot + 128 + 64 + 8 * a0;
                umma_ts_w(d, ah, mk(b_lo), idesc, first);
                umma_ts_w(d, ah + 32, mk(b_lo), idesc, 1u);
                umma_ts_w(d, ah, mk(b_lo + 256), idesc, 1u);
                if (units > 1) {
                  umma_ts_w(d, ah + 8, mk(b_lo + 512), idesc, 1u);
                  umma_ts_w(d, ah + 40, mk(b_lo + 512), idesc, 1u);
                  umma_ts_w(d, ah + 8, mk(b_lo + 768), idesc, 1u);
                }
              } else {
                const uint32_t ah = a_lo + a0 * (CHUNK >> 4);       // lo part: + HALF >> 4
                umma_ss_w(d, mk(ah), mk(b_lo), idesc, first);
                umma_ss_w(d, mk(ah + (HALF >> 4)), mk(b_lo), idesc, 1u);
                umma_ss_w(d, mk(ah), mk(b_lo + 256), idesc, 1u);
                if (units > 1) {
                  umma_ss_w(d, mk(ah + 256), mk(b_lo + 512), idesc, 1u);
                  umma_ss_w(d, mk(ah + 256 + (HALF >> 4)), mk(b_lo + 512), idesc, 1u);
                  umma_ss_w(d, mk(ah + 256), mk(b_lo + 768), idesc, 1u);
                }
              }
              tr.ev(5, r, s);
              if (s == nslots - 1) umma_commit_w(bar_wempty + 8 * st);   // free once the last slot's MMAs have read it
              if (++st == (uint32_t)nst) { st = 0; wph ^= 1; }
            }
            umma_commit_w(bar_acc + 8 * s);
            if (s == nslots - 1) { st0 = st; wph0 = wph; }
            tr.ev(2, r, s, wwait);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ================= epilogue warpgroups =========================================================
    const int g = (warp - 4) >> 2, s = g >> 1, hf = g & 1;
    const int row = ((warp & 3) << 5) | lane;                    // sample row == TMEM lane
    const uint32_t h_row = sbase + s * SLOT + row * 16;          // chunk c of this row: h_row + c*CHUNK (+HALF: lo)
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) << 5) << 16) + s * 256;
    const uint32_t t_mod = t_lane, t_acc = t_lane + 128;
    const int col0 = 64 * hf, bar_id = 1 + s;
    const uint32_t my_acc = pin(bar_acc + 8 * s), my_in = pin(bar_in + 8 * s);
    const float *bias_s = f32s + ucn::X3Layout::b_pts;
    uint32_t ph_acc = 0;
    Tracer tr(((warp & 3) == 0 && lane == 0) ? 1 + g : 6);
    if (!((warp & 3) == 0 && lane == 0)) tr.base = nullptr;
    float head[4], base4[4] = {0.f, 0.f, 0.f, 0.f}, x[3] = {0.f, 0.f, 0.f}, u = 0.f;
    const int dbg = g_dbg;
    auto arrive_in = [&]() {
      tc_fence_before();
      if (!(dbg & 1)) fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(my_in);
    };
    for (long long p = blockIdx.x; p < npairs; p += gridDim.x) {
      const long long tile = 2 * p + s;
      if (tile >= ntiles) break;
      // 32-bit sample / ray arithmetic (the launcher guarantees N < 2^31)
      const uint32_t q0 = (uint32_t)tile * 128u, q = q0 + row, Nu = (uint32_t)N, Su = (uint32_t)S;
      const bool valid = q < Nu;
      const uint32_t ray_first = q0 / Su;
      const uint32_t q_last = q0 + 127u < Nu ? q0 + 127u : Nu - 1u;
      const int nrays_tile = (int)(q_last / Su - ray_first) + 1;
      const int ri = valid ? (int)(q / Su - ray_first) : 0;
      const float *frow = feat + tile * f_tile_stride + (long long)row * f_row_stride;
      // ---- stage inputs: warpgroup 0 -> PE(ndc) in H[0..7]; warpgroup 1 -> [film inputs | 1] in H[8..] ------------
      if (hf == 0) {
        if (valid) { x[0] = ndc_pts[3 * q]; x[1] = ndc_pts[3 * q + 1]; x[2] = ndc_pts[3 * q + 2]; }
        else { x[0] = x[1] = x[2] = 0.f; }
        float pe[64];
        pos_enc<10>(x, pe);
        pe[63] = 0.0f;
#pragma unroll
        for (int c = 0; c < 8; ++c) store_chunk_x3(h_row, c, pe + 8 * c);
        u = 1.0f - (valid ? __ldg(frow + (F - 1) * f_k_stride) : 0.0f);          // models.py:149
        tr.ev(6, 6, g);
      } else {
        // all loads of a 32-channel group are issued before the first use: one memory round trip per group
#pragma unroll 1
        for (int g4 = 0; g4 < L.Kfilm / 8; g4 += 4) {
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int k = 8 * g4 + i;
            v[i] = (k < Kfilm_real && valid) ? __ldg(frow + k * f_k_stride) : (k == Kfilm_real ? 1.0f : 0.0f);
          }
#pragma unroll
          for (int c = 0; c < 4; ++c)
            if (g4 + c < L.Kfilm / 8) store_chunk_x3(h_row, 8 + g4 + c, v + 8 * c);
          tr.ev(6, 7, g);
        }
      }
      arrive_in();

#pragma unroll 1
      for (int r = 0; r < NROUND; ++r) {
        mbar_wait_backoff(my_acc, ph_acc);
        ph_acc ^= 1;
        tc_fence_after();
        tr.ev(3, r, g);
        if (dbg & 2) {
          // experiment: no epilogue work at all
        } else if (r < 5) {
          epi64<0, false>(t_acc, t_mod, col0, bias_s + 128 * r, nullptr, h_row, nullptr, head);
        } else if (r == 5) {
          if (hf == 0) {                                         // skip connection: PE back into H[0..7]
            float pe[64];
            pos_enc<10>(x, pe);
            pe[63] = 0.0f;
#pragma unroll
            for (int c = 0; c < 8; ++c) store_chunk_x3(h_row, c, pe + 8 * c);
          } else {
            // warpgroup 1 is idle in this round: it touches the [img feats] lines round 6 will read, and computes the
            // per-ray view-direction term of views|view_confi (column `row`) for every ray of the tile (read in round 9)
            if (valid)
              for (int k = 0; k < Kfb_real; ++k) touch(frow + (Kfilm_real + k) * f_k_stride);
            for (int i = 0; i < nrays_tile; ++i) {
              const uint32_t ray = ray_first + i;
              const float a[3] = {__ldg(angle + 3 * ray), __ldg(angle + 3 * ray + 1), __ldg(angle + 3 * ray + 2)};
              float pe[27];
              pos_enc<4>(a, pe);
              float acc = 0.f;
#pragma unroll
              for (int k = 0; k < 27; ++k) acc = fmaf(__ldg(f32g + ucn::X3Layout::vd + k * 128 + row), pe[k], acc);
              rayb_s[(s * rpt + i) * 128 + row] = acc;
            }
          }
        } else if (r == 6) {
          head[0] = head[1] = head[2] = head[3] = 0.f;
          epi64<0, true>(t_acc, t_mod, col0, bias_s + 128 * 5, nullptr, h_row, f32s + ucn::X3Layout::w_base, head);
          tr.ev(6, 1, g);
          // [img feats | 1] as split bf16 pairs -> TMEM acc columns 64.. (hi) and 96.. (lo): A operand of fb.  The two
          // warpgroups of the slot stage half of the words each (words 8*half .. 8*half+7 of every 16-word pass), once
          // warpgroup 1 has finished reading those accumulator columns (the named barrier below).
          auto stage_fb_half = [&](int half) {
#pragma unroll 1
            for (int h16 = 0; h16 < (L.Kfb + 31) / 32; ++h16) {     // 32 K values (16 hi + 16 lo words) per pass
              uint32_t hi[8], lo[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const int k = 32 * h16 + 16 * half + 2 * j;
                float a = 0.f, b = 0.f;
                if (k < Kfb_real) a = valid ? __ldg(frow + (Kfilm_real + k) * f_k_stride) : 0.f; else if (k == Kfb_real) a = 1.f;
                if (k + 1 < Kfb_real) b = valid ? __ldg(frow + (Kfilm_real + k + 1) * f_k_stride) : 0.f; else if (k + 1 == Kfb_real) b = 1.f;
                split2(a, b, hi[j], lo[j]);
              }
              tr.ev(6, 2, g);
              tmem_st8(t_acc + 64 + 16 * h16 + 8 * half, hi);
              tmem_st8(t_acc + 96 + 16 * h16 + 8 * half, lo);
            }
            tmem_wait_st();
          };
          if (hf == 1) {
            exch_s[s * 128 + row] = make_float4(head[0], head[1], head[2], head[3]);
            named_bar_arrive(bar_id, 256);
            stage_fb_half(0);
          } else {
            named_bar_sync(bar_id, 256);
            stage_fb_half(1);
            const float4 o = exch_s[s * 128 + row];
            const float *bb = f32s + ucn::X3Layout::b_base;
            base4[0] = head[0] + o.x + bb[0]; base4[1] = head[1] + o.y + bb[1];
            base4[2] = head[2] + o.z + bb[2]; base4[3] = head[3] + o.w + bb[3];
          }
        } else if (r == 7) {
          epi64<3, false>(t_acc, t_mod, col0, nullptr, nullptr, h_row, nullptr, head);
          {
            // touch the next tile's inputs (position, confidence, film channels: half per warpgroup)
            const long long nt = tile + 2ll * gridDim.x;
            const uint32_t nq = (uint32_t)nt * 128u + row;
            if (nt < ntiles && nq < Nu) {
              const float *nrow = feat + nt * f_tile_stride + (long long)row * f_row_stride;
              if (hf == 0) { touch(ndc_pts + 3 * nq); touch(ndc_pts + 3 * nq + 2); touch(nrow + (F - 1) * f_k_stride); }
              for (int k = hf; k < Kfilm_real; k += 2) touch(nrow + k * f_k_stride);
            }
          }
        } else if (r == 8) {
          epi64<1, false>(t_acc, t_mod, col0, f32s + ucn::X3Layout::b_feat, nullptr, h_row, nullptr, head);
          tr.ev(6, 3, g);
        } else {
          head[0] = head[1] = head[2] = head[3] = 0.f;
          epi64<2, true>(t_acc, t_mod, col0, f32s + ucn::X3Layout::b_vv, rayb_s + (s * rpt + ri) * 128, h_row,
                         f32s + ucn::X3Layout::w_adapt, head);
          tr.ev(6, 5, g);
          if (hf == 1) {
            exch_s[s * 128 + row] = make_float4(head[0], head[1], head[2], head[3]);
            named_bar_arrive(bar_id, 256);
          } else {
            named_bar_sync(bar_id, 256);
            const float4 o = exch_s[s * 128 + row];
            const float *bb = f32s + ucn::X3Layout::b_adapt;
            const float ad[4] = {head[0] + o.x + bb[0], head[1] + o.y + bb[1], head[2] + o.z + bb[2], head[3] + o.w + bb[3]};
            const float cf = 1.0f - u;                           // models.py:177-180
            float4 out;
            out.x = 1.0f / (1.0f + __expf(-(base4[0] * cf + ad[0] * u)));
            out.y = 1.0f / (1.0f + __expf(-(base4[1] * cf + ad[1] * u)));
            out.z = 1.0f / (1.0f + __expf(-(base4[2] * cf + ad[2] * u)));
            out.w = fmaxf(ad[3] * cf + base4[3] * u, 0.0f);
            if (valid) reinterpret_cast<float4 *>(raw)[q] = out;
          }
        }
        tr.ev(4, r, g);
        if (r < NROUND - 1) arrive_in();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc512(tmem_base);
  }
}

}  // namespace

// Builds the weight-block / round schedule and the shared-memory carve-up for V source views and S samples/ray.
static int make_plan(int V, int S, Plan &P) {
  const ucn::X3Layout L(V);
  memset(&P, 0, sizeof(P));
  int nb = 0, r = 0;
  auto add = [&](size_t mat_off, int units_total, int a0, int flags, bool ts) {
    for (int u0 = 0; u0 < units_total; u0 += 2) {
      Blk &b = P.blk[nb++];
      b.goff = (uint32_t)(mat_off + (size_t)u0 * ucn::X3Layout::kUnit);
      b.units = (uint8_t)(units_total - u0 >= 2 ? 2 : 1);
      b.a0 = (uint8_t)(ts ? a0 + u0 : a0 + 2 * u0);
      b.flags = (uint8_t)(flags | (ts ? BF_TS : 0) | (u0 == 0 ? 0 : 0));
      if (u0 != 0) b.flags &= ~BF_FRESH;
    }
  };
  P.round_begin[r++] = 0;
  add(L.film, L.Kfilm / 16, 8, BF_FRESH | BF_MOD, false);
  add(L.pts0, 4, 0, BF_FRESH, false);
  P.round_begin[r++] = (uint8_t)nb;                                  // R1..R4
  for (int i = 0; i < 4; ++i) { add(L.pts14[i], 8, 0, BF_FRESH, false); P.round_begin[r++] = (uint8_t)nb; }
  add(L.pts5h, 8, 0, BF_FRESH, false);  P.round_begin[r++] = (uint8_t)nb;   // R5
  add(L.pts5pe, 4, 0, 0, false);        P.round_begin[r++] = (uint8_t)nb;   // R6 (accumulates onto R5)
  add(L.fb, L.Kfb / 16, 0, BF_FRESH | BF_MOD, true);  P.round_begin[r++] = (uint8_t)nb;   // R7
  add(L.feat, 8, 0, BF_FRESH, false);   P.round_begin[r++] = (uint8_t)nb;   // R8
  add(L.vvh, 8, 0, BF_FRESH, false);    P.round_begin[r++] = (uint8_t)nb;   // R9
  if (r != NROUND + 1 || nb > MAXBLK) return -1;
  P.nblk = (uint8_t)nb;
  const int rpt = 127 / S + 2;                                       // rays a 128-sample tile can touch
  if (rpt > 255) return -1;
  P.rpt = (uint8_t)rpt;
  for (int nst = 5; nst >= 4; --nst) {
    uint32_t o = 2 * SLOT + nst * STAGE;
    P.off_f32 = o;  o += (ucn::X3Layout::kSmemFloats * 4 + 127) & ~127u;
    P.off_exch = o; o += 2 * 128 * 16;
    P.off_rayb = o; o += 2u * rpt * 128 * 4;
    P.off_bars = o; o += (4 + 2 * nst) * 8 + 16;
    P.smem_total = o;
    P.nst = (uint8_t)nst;
    if (o <= 232448u - 1024u) return 0;              // 1 KB of static shared memory (block tables)
  }
  return -2;
}

extern "C" int ucn_debug_set_flags(int flags) {
  cudaError_t e = cudaMemcpyToSymbol(g_dbg, &flags, sizeof(flags));
  return e == cudaSuccess ? UCN_OK : ucn_cuda_status(e, "ucn_debug_set_flags");
}
extern "C" int ucn_debug_set_trace(long long *buf) {
  cudaError_t e = cudaMemcpyToSymbol(g_trace, &buf, sizeof(buf));
  return e == cudaSuccess ? UCN_OK : ucn_cuda_status(e, "ucn_debug_set_trace");
}

int ucn_mlp_forward_x3(const ucn_weights *w, const float *ndc_pts, const float *feat, long long f_tile_stride,
                       long long f_k_stride, long long f_row_stride, const float *angle, int R, int S, float *raw,
                       void *stream) {
  UCN_CHECK_ARG(w->bf16x3, "ucn_mlp_forward: split-bf16 weight pack missing");
  UCN_CHECK_ARG(w->V <= 7, "ucn_mlp_forward: tensor-core modes support at most 7 source views");
  Plan P;
  const int rc = make_plan(w->V, S, P);
  if (rc == -2) return ucn_set_error(UCN_E_UNSUPPORTED, "ucn_mlp_forward: BF16X3_TC needs S >= 7 samples per ray (shared-memory budget)");
  if (rc != 0) return ucn_set_error(UCN_E_ARG, "ucn_mlp_forward: cannot build the x3 schedule");
  const long long N = (long long)R * S;
  UCN_CHECK_ARG(N < (1ll << 31), "ucn_mlp_forward: more than 2^31 samples in one call; split the ray range");
  static int attr[kUcnMaxDevices];
  {
    cudaError_t e = ucn_ensure_dyn_smem(mlp_x3_kernel, (int)P.smem_total, attr);
    if (e != cudaSuccess) return ucn_cuda_status(e, "mlp_x3 smem attribute");
  }
  const int sms = ucn_sm_count();
  const long long npairs = ((N + 127) / 128 + 1) / 2;
  const unsigned blocks = (unsigned)(npairs < sms ? npairs : sms);
  mlp_x3_kernel<<<blocks, kThreads, P.smem_total, (cudaStream_t)stream>>>(
      P, reinterpret_cast<const unsigned char *>(w->bf16x3), w->V, ndc_pts, feat, f_tile_stride, f_k_stride,
      f_row_stride, angle, N, S, raw);
  UCN_LAUNCH_CHECK("mlp_x3_kernel");
  return UCN_OK;
}
