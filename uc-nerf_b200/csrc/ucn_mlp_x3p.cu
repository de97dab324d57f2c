// ucn_mlp_x3p.cu -- K3, production mode, second generation ("x3p": split-bf16, TMEM-resident activations, pipelined).
//
// replaces (reference): network/models.py:50-54 Embedder.embed, network/renderer.py:78-106 run_network_mvs
//   (+batchify), network/models.py:138-184 BaseAdapt_Renderer.forward.  Same arithmetic as ucn_mlp_x3.cu
//   (a*w ~= a_hi*w_hi + a_lo*w_hi + a_hi*w_lo on bf16 hi/lo parts, fp32 accumulation in TMEM, three tcgen05.mma per
//   K16 step; heads, FiLM, bias, blend in fp32 on the CUDA cores); what changes is WHERE the activations live.
//
// Round 1's kernel kept the A operand (activations) in shared memory: every MMA read 8 KB of operands through the
// shared-memory pipe the epilogue's stores also use, the accumulator of a round was also its successor's D, so a
// slot's chain MMA -> wake -> epilogue -> hand-off -> MMA was strictly serial (tensor pipe 48 % busy).  Here:
//   * the A operand of every 128-wide layer lives in TENSOR MEMORY (tcgen05.mma "TS" form): the epilogue converts a
//     32-column chunk of the fp32 accumulator IN PLACE into the bf16 hi/lo operand of the next layer
//     (tcgen05.ld -> bias/FiLM/ReLU -> split -> tcgen05.st); an MMA reads only its 4 KB weight tile from shared memory;
//   * each tile slot owns two 128-column TMEM regions P and Q that alternate as (A of round r, D of round r): the
//     next round's MMAs on K-chunk c are issued as soon as the epilogue has converted chunk c -- the epilogue of
//     round r and the MMAs of round r+1 of the SAME slot overlap (four chunk barriers per slot);
//   * the FiLM multiplier `mod` (fp32, 128 x 128 per tile) moves from TMEM to shared memory (64 KB per slot), which is
//     what frees the TMEM columns for the P/Q pair; the same 64 KB stage the narrow shared-memory operands of a tile
//     (PE + film inputs before `mod` exists, image features after its last use).
//
// One persistent CTA per SM, 640 threads: warp 0 weight producer (cp.async.bulk ring, one load per tile PAIR),
// warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-19 four epilogue warpgroups (slot s = g/2; warpgroup hf = g%2
// owns the 32-column chunks {hf, hf+2} of every 128-wide round, so chunks complete in K order).
//
// Rounds per tile (D region; A operand):
//   R0  film -> Q, L0 -> P          (A: film inputs / PE from shared memory)   epilogue: Q -> mod (smem); relu((P+b)*mod) -> P
//   R1..R4  L1..L4  Q,P,Q,P         (A: other region, TMEM)                    relu((D+b)*mod) in place, chunked hand-off
//   R5  L5 hidden part -> Q         (A: P)                                     warpgroup 0 rebuilds PE in P[0..64); wg 1: per-ray terms
//   R6  L5 PE part  -> Q (accum)    (A: P[0..64))                              relu((Q+b)*mod) in place; base heads; image features -> smem
//   R7  fb -> P                     (A: image features, shared memory)         h5(Q) * fb(P) -> P
//   R8  feature_linear -> Q         (A: P)                                     Q + b in place, chunked; next tile's inputs -> smem
//   R9  views|view_confi -> P       (A: Q)                                     relu(P + b + per-ray term); adapt heads; blend -> raw
//
// Roofline: tensor-bound.  Algorithmic FLOPs/sample = 2*(138112+1536V); three MMAs are issued per
// algorithmic K16 step (no layer group tolerates fewer passes at the 1e-3 gate, profiles/r02_precision_*).
#include <string.h>

#include <mutex>
#include <vector>

#include "ucn_common.cuh"
#include "ucn_tc.cuh"
#include "ucn_weights.cuh"

using namespace ucn_tc;

namespace {

constexpr int kThreads = 640;
constexpr uint32_t CHUNK = 2048;             // one 8-wide K-chunk (or one 4-column fp32 group) x 128 rows
constexpr uint32_t MODB = 32 * CHUNK;        // 64 KB per slot: mod [32 column groups][128 rows][4 fp32]
constexpr uint32_t LOD = 8 * CHUNK;          // shared-memory A operands: lo part 8 chunks after the hi part
constexpr uint32_t STAGE = 2 * ucn::X3Layout::kUnit;   // 16 KB ring stage = two K16 units (hi+lo each)
constexpr int NROUND = 10;
constexpr int MAXBLK = 48;

__device__ long long *g_trace_p = nullptr;   // CTA-0 timeline (tools/trace_x3.py), off unless a buffer is set
__device__ int g_dbg_p = 0;                  // timing experiments (tools/trace_x3.py X3_DBG); 0 in production
constexpr int kTraceCap = 2048;
struct Tracer {
  long long *base; int n;
  __device__ __forceinline__ Tracer(int who) : base(nullptr), n(0) {
    if (g_trace_p && blockIdx.x == 0) base = g_trace_p + (size_t)who * kTraceCap * 2;
  }
  __device__ __forceinline__ void ev(int tag, int r, int who, long long payload = 0) {
    if (base && n < kTraceCap) {
      base[2 * n] = (long long)tag | ((long long)r << 8) | ((long long)who << 16) | (payload << 24);
      base[2 * n + 1] = clock64();
      ++n;
    }
  }
};

// one weight block = one ring stage (<= 2 K16 units) and the MMAs that consume it
enum { BF_FRESH = 1, BF_DQ = 2, BF_SS = 4 };      // first MMA overwrites D; D region is Q (else P); A from shared memory
// a0: SS first K-chunk in the slot's 64 KB / TS first unit; wait: go barriers (bit j = chunk j) that must have completed
// before this block's MMAs are issued
struct Blk { uint32_t goff; uint8_t units, a0, flags, wait; };
struct Plan {
  Blk blk[MAXBLK];
  uint8_t round_begin[NROUND + 1];
  uint8_t a_reg[NROUND];                     // TS rounds: A region (0 = P, 1 = Q)
  uint8_t nblk, nst, rpt, pad;
  uint32_t off_f32, off_exch, off_rayb, off_ubuf, off_bars, smem_total;
};

// (v0..v15) -> one K16 unit of the TMEM A operand: 8 words of bf16 hi pairs, then 8 words of lo pairs
__device__ __forceinline__ void split16(const float *v, uint32_t (&w)[16]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) split2(v[2 * i], v[2 * i + 1], w[i], w[8 + i]);
}
__device__ __forceinline__ void tmem_ld16w(uint32_t taddr, uint32_t (&r)[16]) { tmem_ld16(taddr, r); }

// write 8 consecutive K values (one chunk) of this thread's row as hi/lo bf16 into a shared-memory A operand
__device__ __forceinline__ void store_chunk_ss(uint32_t row_addr, int chunk, const float *v) {
  uint32_t hi[4], lo[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) split2(v[2 * i], v[2 * i + 1], hi[i], lo[i]);
  sts128(row_addr + chunk * CHUNK, hi[0], hi[1], hi[2], hi[3]);
  sts128(row_addr + LOD + chunk * CHUNK, lo[0], lo[1], lo[2], lo[3]);
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

// sm_100 packed fp32 arithmetic: two IEEE round-to-nearest operations per instruction (same results as the scalar forms)
__device__ __forceinline__ void add2_f32(float &x0, float &x1, float b0, float b1) {
  asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tadd.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
      : "+f"(x0), "+f"(x1) : "f"(b0), "f"(b1));
}
__device__ __forceinline__ void mul2_f32(float &x0, float &x1, float b0, float b1) {
  asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tmul.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
      : "+f"(x0), "+f"(x1) : "f"(b0), "f"(b1));
}
__device__ __forceinline__ void sub2_f32(float &x0, float &x1, float b0, float b1) {
  asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tsub.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
      : "+f"(x0), "+f"(x1) : "f"(b0), "f"(b1));
}
// (a0, a1) += v * (w0, w1): one FFMA2 (the scalar operand is broadcast, the weight pair comes from a uniform register)
__device__ __forceinline__ void fma2_bcast(float &a0, float &a1, float v, float w0, float w1) {
  asm("{\n\t.reg .b64 a, b, c;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %2};\n\tmov.b64 c, {%3, %4};\n\tfma.rn.f32x2 a, b, c, a;\n\tmov.b64 {%0, %1}, a;\n\t}"
      : "+f"(a0), "+f"(a1) : "f"(v), "f"(w0), "f"(w1));
}
// split2 with the two residual subtractions as one packed instruction
__device__ __forceinline__ void split2p(float a, float b, uint32_t &hi, uint32_t &lo) {
  hi = pack2(a, b);
  sub2_f32(a, b, bf_lo(hi), bf_hi(hi));
  lo = pack2(a, b);
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void touch(const float *p) {
  float t;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(t) : "l"(p));
}

// Epilogue of ONE 32-column chunk [c0, c0+32) of a 128-wide round.
//   MODE 0: relu((D + b) * mod) -> A operand in place     pts layers                        (models.py:154-155)
//   MODE 1: D + b -> A operand in place                   feature_linear                    (models.py:164)
//   MODE 2: relu(D + b + per-ray term), nothing stored    views / view_confi hidden         (models.py:168-173)
//   MODE 3: h(other region, A format) * D -> A in place   h * feats_bias (bias folded into the MMA)   (models.py:164)
// HEADS: head4 += v[c] * w4[c] (fp32) over this chunk's columns.
// C0 is a template parameter so that the head weights are immediate-offset constant-bank operands of the FFMAs (no
// load instruction: the 4-wide heads were 256 broadcast LDS.128 per thread and round through the shared-memory pipe).
// (C0 < 0: the chunk offset is the runtime argument `c0r` -- one instantiation per mode keeps the code in the I-cache.)
// HM: which of the four head outputs this chunk's columns feed (bit k = output k).  The adapt heads are block-structured --
// rgb reads columns 0..63 of the view hidden, alpha columns 64..127 (models.py:170,174) -- so half of their 512 products per row
// are multiplications by the zero blocks of [[W_rgb, 0], [0, W_alpha]]: skipping them is bit-identical (x * 0 + acc == acc).
template <int MODE, bool HEADS, int C0, int HM = 15>
__device__ __forceinline__ void epi32(uint32_t t_d, uint32_t t_other, const float *__restrict__ bias,
                                      const float *__restrict__ rayb, uint32_t mod_row, const float (&w4)[512],
                                      float (&head)[4], int c0r = 0) {
  static_assert(!HEADS || C0 >= 0, "head weights are immediate-offset constants: the chunk must be a template parameter");
  const int c0 = C0 >= 0 ? C0 : c0r;
  uint32_t a[32];
  tmem_ld32(t_d + c0, a);
  uint32_t hq[32];
  if (MODE == 3) tmem_ld32(t_other + c0, hq);            // h5 as [hi8 | lo8] x 2 units
  tmem_wait_ld();
#pragma unroll
  for (int u = 0; u < 2; ++u) {                          // one K16 unit of the next layer's operand at a time
    float v[16];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int c = c0 + 16 * u + 8 * j;
      if (MODE == 3) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t wh = hq[16 * u + 4 * j + i], wl = hq[16 * u + 8 + 4 * j + i];
          float h0 = bf_lo(wh), h1 = bf_hi(wh);
          add2_f32(h0, h1, bf_lo(wl), bf_hi(wl));
          mul2_f32(h0, h1, __uint_as_float(a[16 * u + 8 * j + 2 * i]), __uint_as_float(a[16 * u + 8 * j + 2 * i + 1]));
          v[8 * j + 2 * i] = h0; v[8 * j + 2 * i + 1] = h1;
        }
      } else {
        const float4 b0 = *reinterpret_cast<const float4 *>(bias + c), b1 = *reinterpret_cast<const float4 *>(bias + c + 4);
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        float mm[8], rb[8];
        if (MODE == 0) {
          const uint4 m0 = lds128(mod_row + (c >> 2) * CHUNK), m1 = lds128(mod_row + ((c >> 2) + 1) * CHUNK);
          mm[0] = __uint_as_float(m0.x); mm[1] = __uint_as_float(m0.y); mm[2] = __uint_as_float(m0.z); mm[3] = __uint_as_float(m0.w);
          mm[4] = __uint_as_float(m1.x); mm[5] = __uint_as_float(m1.y); mm[6] = __uint_as_float(m1.z); mm[7] = __uint_as_float(m1.w);
        }
        if (MODE == 2) {
          const float4 r0 = *reinterpret_cast<const float4 *>(rayb + c), r1 = *reinterpret_cast<const float4 *>(rayb + c + 4);
          rb[0] = r0.x; rb[1] = r0.y; rb[2] = r0.z; rb[3] = r0.w; rb[4] = r1.x; rb[5] = r1.y; rb[6] = r1.z; rb[7] = r1.w;
        }
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          float x0 = __uint_as_float(a[16 * u + 8 * j + i]), x1 = __uint_as_float(a[16 * u + 8 * j + i + 1]);
          add2_f32(x0, x1, bb[i], bb[i + 1]);
          if (MODE == 0) { mul2_f32(x0, x1, mm[i], mm[i + 1]); x0 = fmaxf(x0, 0.0f); x1 = fmaxf(x1, 0.0f); }
          if (MODE == 2) { add2_f32(x0, x1, rb[i], rb[i + 1]); x0 = fmaxf(x0, 0.0f); x1 = fmaxf(x1, 0.0f); }
          v[8 * j + i] = x0; v[8 * j + i + 1] = x1;
        }
      }
      if (HEADS) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if ((HM & 3) == 3) fma2_bcast(head[0], head[1], v[8 * j + i], w4[4 * (c + i)], w4[4 * (c + i) + 1]);
          else {
            if (HM & 1) head[0] = fmaf(v[8 * j + i], w4[4 * (c + i)], head[0]);
            if (HM & 2) head[1] = fmaf(v[8 * j + i], w4[4 * (c + i) + 1], head[1]);
          }
          if ((HM & 12) == 12) fma2_bcast(head[2], head[3], v[8 * j + i], w4[4 * (c + i) + 2], w4[4 * (c + i) + 3]);
          else {
            if (HM & 4) head[2] = fmaf(v[8 * j + i], w4[4 * (c + i) + 2], head[2]);
            if (HM & 8) head[3] = fmaf(v[8 * j + i], w4[4 * (c + i) + 3], head[3]);
          }
        }
      }
    }
    if (MODE != 2) {                                       // the unit's accumulator words are dead: reuse them for the operand
#pragma unroll
      for (int i = 0; i < 8; ++i) split2p(v[2 * i], v[2 * i + 1], a[16 * u + i], a[16 * u + 8 + i]);
    }
  }
  if (MODE != 2) tmem_st32(t_d + c0, a);                   // one 4 KB store per warp and chunk
}

// both 32-column chunks of a warpgroup (hf 0: columns 0..31 and 64..95, hf 1: 32..63 and 96..127), optionally handing each
// chunk to the MMA issuer as soon as it is converted
#define EPI_LOOP(MODE, TD, TO, BIAS, HANDOFF)                                                       \
  do {                                                                                              \
    _Pragma("unroll 1") for (int h2 = 0; h2 < 2; ++h2) {                                            \
      if (!(dbg & 1)) epi32<MODE, false, -1>(TD, TO, BIAS, nullptr, mod_row, hw.base, head, h2 ? cB : cA); \
      if (HANDOFF) arrive(h2 ? goB : goA);                                                          \
    }                                                                                               \
  } while (0)
#define EPI_PAIR(MODE, HEADS, TD, TO, BIAS, RB, W4, HANDOFF)                                        \
  do {                                                                                              \
    if (hf == 0) {                                                                                  \
      epi32<MODE, HEADS, 0>(TD, TO, BIAS, RB, mod_row, W4, head);  if (HANDOFF) arrive(goA);        \
      epi32<MODE, HEADS, 64>(TD, TO, BIAS, RB, mod_row, W4, head); if (HANDOFF) arrive(goB);        \
    } else {                                                                                        \
      epi32<MODE, HEADS, 32>(TD, TO, BIAS, RB, mod_row, W4, head); if (HANDOFF) arrive(goA);        \
      epi32<MODE, HEADS, 96>(TD, TO, BIAS, RB, mod_row, W4, head); if (HANDOFF) arrive(goB);        \
    }                                                                                               \
  } while (0)

struct HeadW { float base[512], adapt[512]; };   // [column][4]: confi_rgb|alpha_1 on h5; rgb (cols 0..63) | alpha (64..127) on the view hidden

__global__ void __launch_bounds__(kThreads, 1)
mlp_x3p_kernel(const __grid_constant__ Plan plan, const __grid_constant__ HeadW hw, const unsigned char *__restrict__ wpack, int V,
               const float *__restrict__ ndc_pts, const unsigned char *__restrict__ oper, long long rec_bytes,
               const float *__restrict__ rayterm, long long N, int S, float *__restrict__ raw,
               uint4 *__restrict__ pe_stash) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const ucn::X3Layout L(V);
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int nst = plan.nst, rpt = plan.rpt, nblk = plan.nblk;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t ring = sbase + 2 * MODB;
  // barriers: go[2][4] (count 4: the warps of the warpgroup that owns the chunk), acc[2] (count 1), wfull/wempty[nst];
  // stager hand-shake per slot: modfree (count 8: every epilogue warp after its last FiLM read of the tile), r7 (count 1:
  // fb's MMAs done, tcgen05.commit), staged (count 2: both stager warps)
  const uint32_t bar_go = pin(sbase + plan.off_bars), bar_acc = pin(bar_go + 64), bar_wfull = pin(bar_acc + 16),
                 bar_wempty = pin(bar_wfull + 8 * nst), bar_modfree = pin(bar_wempty + 8 * nst), bar_r7 = pin(bar_modfree + 16),
                 bar_staged = pin(bar_r7 + 16), tmem_ptr_addr = bar_staged + 16;
  float *ubuf_s = reinterpret_cast<float *>(smem + plan.off_ubuf);      // [slot][tile parity][row]: 1 - confidence
  float *f32s = reinterpret_cast<float *>(smem + plan.off_f32);
  float4 *exch_s = reinterpret_cast<float4 *>(smem + plan.off_exch);
  float *rayb_s = reinterpret_cast<float *>(smem + plan.off_rayb);
  __shared__ uint32_t blk_s[MAXBLK], goff_s[MAXBLK];
  // K-chunks of the film / img operands, and how many of them K2's operand tile stores (ucn_gather.cu): the rest -- the
  // chunk that is only the ones column (bias), and zero K padding -- are constants written here
  const int Kf8 = L.Kfilm / 8, Kb8 = L.Kfb / 8, Kfd = 3 + (V + 1) / 2, Kbd = V;
  const float *f32g = reinterpret_cast<const float *>(wpack + L.f32);

  // ---- one-time setup ---------------------------------------------------------------------------
  for (int i = threadIdx.x; i < ucn::X3Layout::kSmemFloats; i += kThreads) f32s[i] = f32g[i];
  if (threadIdx.x < plan.nblk) {
    const Blk k = plan.blk[threadIdx.x];
    blk_s[threadIdx.x] = (uint32_t)k.units | ((uint32_t)k.a0 << 8) | ((uint32_t)k.flags << 16);
    goff_s[threadIdx.x] = k.goff;
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(bar_go + 8 * i, 4);
    mbar_init(bar_acc, 1); mbar_init(bar_acc + 8, 1);
    for (int i = 0; i < nst; ++i) { mbar_init(bar_wfull + 8 * i, 1); mbar_init(bar_wempty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_modfree + 8 * i, 8); mbar_init(bar_r7 + 8 * i, 1); mbar_init(bar_staged + 8 * i, 2); }
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
    // Rounds run slot 0 then slot 1 (both slots consume each weight stage, loaded once per tile pair); a block waits only
    // for the go barriers its operands depend on (chunk j of the slot's previous epilogue for a 128-wide TMEM operand).
    const uint32_t idesc = umma_idesc_bf16(128);
    const uint64_t dt = umma_desc(0, CHUNK, 128);                 // LBO = 2048 (K-chunks of A, tiles of B), SBO = 128
    const uint32_t dhi = (uint32_t)(dt >> 32), dlo = (uint32_t)dt;
    auto mk = [&](uint32_t lo) { return ((uint64_t)dhi << 32) | lo; };
    uint32_t st0 = 0, wph0 = 0, ph_go0 = 0, ph_go1 = 0, ph_staged = 0;
    Tracer tr(0);
    if (lane != 0) tr.base = nullptr;
    const uint32_t ring_lo = dlo + (ring >> 4);
    for (long long p = blockIdx.x; p < npairs; p += gridDim.x) {
      const int nslots = (2 * p + 1 < ntiles) ? 2 : 1;
#pragma unroll 1
      for (int r = 0; r < NROUND; ++r) {
        const int b0 = plan.round_begin[r], b1 = plan.round_begin[r + 1];
        const uint32_t a_reg = plan.a_reg[r];
#pragma unroll 1
        for (int s = 0; s < nslots; ++s) {
          const uint32_t go = bar_go + 32 * s, ph = s ? ph_go1 : ph_go0;
          if (r == 0) mbar_wait(bar_staged + 8 * s, ph_staged);   // the stager warps have placed this tile's operands
          tr.ev(1, r, s);
          const uint32_t a_ss = dlo + ((sbase + s * MODB) >> 4);
          const uint32_t t_slot = tmem_base + s * 256;
          uint32_t st = st0, wph = wph0;
          long long w_go = 0, w_full = 0;                         // traced CTA only: cycles spent waiting
#pragma unroll 1
          for (int b = b0; b < b1; ++b) {
            const uint32_t units = plan.blk[b].units, a0 = plan.blk[b].a0, flags = plan.blk[b].flags, wm = plan.blk[b].wait;
            const long long tw0 = tr.base ? clock64() : 0;
            if (wm) {
              // (every go barrier is waited for in every round: a parity wait cannot tell phases two apart)
              if (wm & 1) mbar_wait(go, ph);
              if (wm & 2) mbar_wait(go + 8, ph);
              if (wm & 4) mbar_wait(go + 16, ph);
              if (wm & 8) mbar_wait(go + 24, ph);
              tc_fence_after();
            }
            const long long tw1 = tr.base ? clock64() : 0;
            if (s == 0) mbar_wait(bar_wfull + 8 * st, wph);       // slot 1 reuses the stage slot 0 waited for
            if (tr.base) { const long long tw2 = clock64(); w_go += tw1 - tw0; w_full += tw2 - tw1; }
            const uint32_t b_lo = ring_lo + st * (STAGE >> 4);
            const uint32_t d = t_slot + ((flags & BF_DQ) ? 128 : 0);
            const uint32_t first = (flags & BF_FRESH) ? 0u : 1u;
            if (flags & BF_SS) {
              const uint32_t ah = a_ss + a0 * (CHUNK >> 4);
              umma_ss_w(d, mk(ah), mk(b_lo), idesc, first);
              umma_ss_w(d, mk(ah + (LOD >> 4)), mk(b_lo), idesc, 1u);
              umma_ss_w(d, mk(ah), mk(b_lo + 256), idesc, 1u);
              if (units > 1) {
                umma_ss_w(d, mk(ah + 256), mk(b_lo + 512), idesc, 1u);
                umma_ss_w(d, mk(ah + 256 + (LOD >> 4)), mk(b_lo + 512), idesc, 1u);
                umma_ss_w(d, mk(ah + 256), mk(b_lo + 768), idesc, 1u);
              }
            } else {
              const uint32_t ah = t_slot + 128 * a_reg + 16 * a0;   // unit: [hi 8 columns | lo 8 columns]
              umma_ts_w(d, ah, mk(b_lo), idesc, first);
              umma_ts_w(d, ah + 8, mk(b_lo), idesc, 1u);
              umma_ts_w(d, ah, mk(b_lo + 256), idesc, 1u);
              if (units > 1) {
                umma_ts_w(d, ah + 16, mk(b_lo + 512), idesc, 1u);
                umma_ts_w(d, ah + 24, mk(b_lo + 512), idesc, 1u);
                umma_ts_w(d, ah + 16, mk(b_lo + 768), idesc, 1u);
              }
            }
            tr.ev(5, r, s);
            if (s == nslots - 1) umma_commit_w(bar_wempty + 8 * st);   // free once the last slot's MMAs have read it
            if (++st == (uint32_t)nst) { st = 0; wph ^= 1; }
          }
          umma_commit_w(bar_acc + 8 * s);
          if (r == 7) umma_commit_w(bar_r7 + 8 * s);              // fb has read the image features: chunks 0..15 are free
          if (s) ph_go1 ^= 1; else ph_go0 ^= 1;
          if (s == nslots - 1) { st0 = st; wph0 = wph; }
          tr.ev(2, r, s, (w_go << 20) | (w_full & 0xFFFFF));
        }
        if (r == 0) ph_staged ^= 1;
      }
    }
    __syncwarp();
  } else if (warp == 2 || warp == 3) {
    // ================= stager warps =================================================================
    // Everything a tile reads from global memory except the image features is brought in HERE, off the epilogue warps'
    // serial chain (a global load issued under this kernel's shared-memory / TMEM traffic takes 2-3 K cycles, a PE
    // rebuild ~4 K): per tile k of slot s and row,
    //   (a) any time:                  x -> PE(x) as bf16 hi/lo words -> global stash[parity] (L2); 1 - confidence -> ubuf
    //   (b) after tile k-1's last FiLM epilogue (modfree):  [film inputs | 1] chunks (K2's operand tile) -> chunks 16.. /
    //       24.. of the slot's 64 KB by cp.async: no registers, all copies of a slot in flight at once
    //   (c) after tile k-1's fb MMAs (r7):                  stash -> chunks 0..7 / 8..15 (the shared-memory operand of L0)
    // then `staged`.  Each lane serves two rows of each slot (warp 2: rows 0..63, warp 3: 64..127).
    const uint32_t Nu = (uint32_t)N;
    uint32_t ph_mod = 0, ph_r7w = 0;                             // both slots advance in lock step
    for (long long k = 0;; ++k) {
      const long long t0 = 2ll * (blockIdx.x + k * gridDim.x);
      if (t0 >= ntiles) break;
      const int nsl = (t0 + 1 < ntiles) ? 2 : 1;
      const int par = (int)(k & 1);
#pragma unroll 1
      for (int sl = 0; sl < nsl; ++sl) {                          // (a)
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          const int row = (warp - 2) * 64 + 32 * h + lane;
          const uint32_t q = (uint32_t)(t0 + sl) * 128u + row;
          const bool valid = q < Nu;
          float x[3] = {0.f, 0.f, 0.f};
          if (valid) { x[0] = __ldg(ndc_pts + 3 * q); x[1] = __ldg(ndc_pts + 3 * q + 1); x[2] = __ldg(ndc_pts + 3 * q + 2); }
          const float *conf = reinterpret_cast<const float *>(oper + (t0 + sl) * rec_bytes + (size_t)(2 * Kfd + 2 * Kbd) * 2048);
          ubuf_s[(sl * 2 + par) * 128 + row] = 1.0f - __ldg(conf + row);                                     // models.py:149
          float pe[64];
          pos_enc<10>(x, pe);
          pe[63] = 0.0f;
          uint4 *st4 = pe_stash + (((size_t)blockIdx.x * 2 + sl) * 2 + par) * 16 * 128 + row;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) split2(pe[8 * c + 2 * i], pe[8 * c + 2 * i + 1], hi[i], lo[i]);
            st4[c * 128] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            st4[(8 + c) * 128] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          }
        }
      }
#pragma unroll 1
      for (int sl = 0; sl < nsl; ++sl) {
        const uint32_t mod_s = sbase + sl * MODB;
        if (k > 0) mbar_wait_backoff(bar_modfree + 8 * sl, ph_mod);          // (b)
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {                              // K2 wrote the operand chunks: copy them verbatim
          const int row = (warp - 2) * 64 + 32 * h + lane;
          const unsigned char *rec = oper + (t0 + sl) * rec_bytes + row * 16;
#pragma unroll 1
          for (int c = 0; c < Kf8; ++c) {
            if (c < Kfd) {
              cp_async16(mod_s + (16 + c) * CHUNK + row * 16, rec + (size_t)c * 2048);
              cp_async16(mod_s + (24 + c) * CHUNK + row * 16, rec + (size_t)(Kfd + c) * 2048);
            } else {                                               // ones column alone (V even) / zero padding
              sts128(mod_s + (16 + c) * CHUNK + row * 16, (c == Kfd && !(V & 1)) ? 0x3F80u : 0u, 0u, 0u, 0u);
              sts128(mod_s + (24 + c) * CHUNK + row * 16, 0u, 0u, 0u, 0u);
            }
          }
        }
        cp_async_commit();
        if (k > 0) mbar_wait_backoff(bar_r7 + 8 * sl, ph_r7w);               // (c)
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          const int row = (warp - 2) * 64 + 32 * h + lane;
          const uint4 *st4 = pe_stash + (((size_t)blockIdx.x * 2 + sl) * 2 + par) * 16 * 128 + row;
          uint4 w[16];
#pragma unroll
          for (int c = 0; c < 16; ++c) w[c] = st4[c * 128];                  // this thread's own stores (a)
#pragma unroll
          for (int c = 0; c < 16; ++c) sts128(mod_s + row * 16 + c * CHUNK, w[c].x, w[c].y, w[c].z, w[c].w);
        }
        cp_async_wait_all();
        __threadfence();                                           // stash and ubuf are read by the epilogue warps
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_staged + 8 * sl);
      }
      if (k > 0) { ph_mod ^= 1; ph_r7w ^= 1; }
    }
  } else if (warp >= 4) {
    // ================= epilogue warpgroups =========================================================
    const int g = (warp - 4) >> 2, s = g >> 1, hf = g & 1;
    const int row = ((warp & 3) << 5) | lane;                    // sample row == TMEM lane
    const uint32_t mod_row = sbase + s * MODB + row * 16;        // column group cg of this row: mod_row + cg*CHUNK
    const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) << 5) << 16) + s * 256;
    const uint32_t t_P = t_lane, t_Q = t_lane + 128;
    const int cA = 32 * hf, cB = 32 * hf + 64;                   // this warpgroup's two 32-column chunks
    const uint32_t my_acc = pin(bar_acc + 8 * s), goA = pin(bar_go + 32 * s + 8 * hf), goB = pin(goA + 16);
    const int bar_id = 1 + s, bar_id2 = 3 + s;
    const float *bias_s = f32s + ucn::X3Layout::b_pts;
    uint32_t ph_acc = 0;
    Tracer tr(((warp & 3) == 0 && lane == 0) ? 1 + g : 6);
    if (!((warp & 3) == 0 && lane == 0)) tr.base = nullptr;
    float head[4], base4[4] = {0.f, 0.f, 0.f, 0.f}, u_cur = 0.f;
    const int dbg = g_dbg_p;             // timing experiment (results wrong): 1 = skip the 128-wide epilogues of rounds 0-4, 7, 8

    // hand a chunk (TMEM operand) / everything (shared-memory operands) to the MMA issuer
    auto arrive = [&](uint32_t bar) {
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar);
    };
    auto arrive_both = [&](bool smem_written) {
      tmem_wait_st();
      tc_fence_before();
      if (smem_written) fence_async_smem();
      __syncwarp();
      if (lane == 0) { mbar_arrive(goA); mbar_arrive(goB); }
    };
    // per-tile state of this thread's row
    struct TileCtx { uint32_t q, ray_first; int nrays, ri; bool valid; const unsigned char *rec; };
    const uint32_t Nu = (uint32_t)N, Su = (uint32_t)S;
    auto open_tile = [&](long long tile, TileCtx &c) {
      const uint32_t q0 = (uint32_t)tile * 128u;
      c.q = q0 + row;
      c.valid = c.q < Nu;
      c.ray_first = q0 / Su;
      const uint32_t q_last = q0 + 127u < Nu ? q0 + 127u : Nu - 1u;
      c.nrays = (int)(q_last / Su - c.ray_first) + 1;
      c.ri = c.valid ? (int)(c.q / Su - c.ray_first) : 0;
      c.rec = oper + tile * rec_bytes + row * 16;                 // this row's 16 bytes of chunk 0 of the tile's record
    };
    // [img feats | 1]: K2's hi / lo chunks, 4 chunks (32 channels) per batch, parked in the idle half of P (columns 64.. hi
    // words, 96.. lo words) until `mod` is dead
    auto park_imgf = [&](const TileCtx &c, int b) {
      uint32_t wh[16], wl[16];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint4 h4 = make_uint4(0, 0, 0, 0), l4 = h4;
        if (4 * b + k < Kbd) {
          h4 = __ldg(reinterpret_cast<const uint4 *>(c.rec + (size_t)(2 * Kfd + 4 * b + k) * 2048));
          l4 = __ldg(reinterpret_cast<const uint4 *>(c.rec + (size_t)(2 * Kfd + Kbd + 4 * b + k) * 2048));
        } else if (4 * b + k == Kbd) h4.x = 0x3F80u;               // the ones column (bias of pts_bias_confidence): bf16 1.0
        wh[4 * k] = h4.x; wh[4 * k + 1] = h4.y; wh[4 * k + 2] = h4.z; wh[4 * k + 3] = h4.w;
        wl[4 * k] = l4.x; wl[4 * k + 1] = l4.y; wl[4 * k + 2] = l4.z; wl[4 * k + 3] = l4.w;
      }
      tmem_st16(t_P + 64 + 16 * b, wh);
      tmem_st16(t_P + 96 + 16 * b, wl);
    };

    TileCtx cur;
    long long tile = 2ll * blockIdx.x + s;
    if (tile < ntiles) arrive_both(false);                       // round 0 of the first tile: P and Q are free
    for (int kt = 0; tile < ntiles; tile += 2ll * gridDim.x, ++kt) {
      const bool has_next = tile + 2ll * gridDim.x < ntiles;
      const int par = kt & 1;                                    // parity of the stager's double buffers (stash, ubuf)
      open_tile(tile, cur);
#pragma unroll 1
      for (int r = 0; r < NROUND; ++r) {
        mbar_wait_backoff(my_acc, ph_acc);
        ph_acc ^= 1;
        tc_fence_after();
        tr.ev(3, r, g);
        if (r == 0) {
          // Q holds mod = film(...) (+ bias, folded): move this thread's 64 columns to shared memory (the A operands
          // staged there are consumed), then -- once BOTH warpgroups have read their part of Q, which R1 overwrites --
          // the FiLM epilogue of L0 on P
#pragma unroll 1
          for (int h2 = 0; h2 < 2; ++h2) {
            const int c0 = h2 ? cB : cA;
            uint32_t m[32];
            tmem_ld32(t_Q + c0, m);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 8; ++j) sts128(mod_row + ((c0 >> 2) + j) * CHUNK, m[4 * j], m[4 * j + 1], m[4 * j + 2], m[4 * j + 3]);
          }
          tc_fence_before();
          named_bar_sync(bar_id2, 256);
        }
        if (r < 5) {
          const uint32_t t_d = (r & 1) ? t_Q : t_P;
          EPI_LOOP(0, t_d, 0, bias_s + 128 * r, true);
        } else if (r == 5) {
          if (hf == 0) {                                         // skip connection: PE words back as TMEM operand units 0..3 of P
            uint4 sw[16];
#pragma unroll
            u_cur = ubuf_s[(s * 2 + par) * 128 + row];           // (the stager reuses this buffer once round 7 is done)
            const uint4 *stash = pe_stash + (((size_t)blockIdx.x * 2 + s) * 2 + par) * 16 * 128 + row;
#pragma unroll
            for (int j = 0; j < 16; ++j) sw[j] = __ldcg(stash + j * 128);
#pragma unroll
            for (int u4 = 0; u4 < 4; ++u4) {
              const uint32_t w[16] = {sw[2 * u4].x, sw[2 * u4].y, sw[2 * u4].z, sw[2 * u4].w,
                                      sw[2 * u4 + 1].x, sw[2 * u4 + 1].y, sw[2 * u4 + 1].z, sw[2 * u4 + 1].w,
                                      sw[8 + 2 * u4].x, sw[8 + 2 * u4].y, sw[8 + 2 * u4].z, sw[8 + 2 * u4].w,
                                      sw[8 + 2 * u4 + 1].x, sw[8 + 2 * u4 + 1].y, sw[8 + 2 * u4 + 1].z, sw[8 + 2 * u4 + 1].w};
              tmem_st16(t_P + 16 * u4, w);
            }
            arrive_both(false);
          } else {
            // warpgroup 1 has nothing round 6 depends on: it releases its go barriers at once, then (off the critical path)
            // parks the tile's image features in P[64..128) -- columns round 6 does not touch -- and starts the asynchronous
            // copy of the tile's per-ray view-direction terms (ray_term_kernel) into shared memory for round 9
            arrive_both(false);
#pragma unroll 1
            for (int i = row >> 5; i < cur.nrays; i += 4)
              cp_async16(smem_u32(rayb_s + (s * rpt + i) * 128 + (row & 31) * 4),
                         rayterm + (size_t)(cur.ray_first + i) * 128 + (row & 31) * 4);
            cp_async_commit();
            park_imgf(cur, 0);
            if (Kb8 > 4) park_imgf(cur, 1);
            tmem_wait_st();                                      // read back by this thread in round 6
          }
        } else if (r == 6) {
          head[0] = head[1] = head[2] = head[3] = 0.f;
          EPI_PAIR(0, true, t_Q, 0, bias_s + 128 * 5, nullptr, hw.base, false);
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_modfree + 8 * s);       // this warp's last `mod` read of the tile: chunks 16.. may be restaged
          // [img feats | 1] from their parking columns (written by this very thread in round 5) to the shared-memory A
          // operand of fb, hi words -> chunks 0.., lo words -> chunks 8..: those column groups are `mod` of both
          // warpgroups, so warpgroup 1 moves them once both have finished their last FiLM epilogue of the tile
          named_bar_sync(bar_id2, 256);
          if (hf == 1) {
#pragma unroll 1
            for (int b2 = 0; b2 < (L.Kfb + 31) / 32; ++b2) {
              uint32_t wh[16], wl[16];
              tmem_ld16(t_P + 64 + 16 * b2, wh);
              tmem_ld16(t_P + 96 + 16 * b2, wl);
              tmem_wait_ld();
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (4 * b2 + k < L.Kfb / 8) {
                  sts128(mod_row + (4 * b2 + k) * CHUNK, wh[4 * k], wh[4 * k + 1], wh[4 * k + 2], wh[4 * k + 3]);
                  sts128(mod_row + LOD + (4 * b2 + k) * CHUNK, wl[4 * k], wl[4 * k + 1], wl[4 * k + 2], wl[4 * k + 3]);
                }
            }
          }
          arrive_both(true);
          if (hf == 1) {
            exch_s[s * 128 + row] = make_float4(head[0], head[1], head[2], head[3]);
            named_bar_arrive(bar_id, 256);
          } else {
            named_bar_sync(bar_id, 256);
            const float4 o = exch_s[s * 128 + row];
            const float *bb = f32s + ucn::X3Layout::b_base;
            base4[0] = head[0] + o.x + bb[0]; base4[1] = head[1] + o.y + bb[1];
            base4[2] = head[2] + o.z + bb[2]; base4[3] = head[3] + o.w + bb[3];
          }
        } else if (r == 7) {
          EPI_LOOP(3, t_P, t_Q, nullptr, false);
          arrive_both(false);
        } else if (r == 8) {
          if (hf == 1) cp_async_wait_all();                      // per-ray terms landed (ordered before round 9 by the go barrier)
          const float *bf = f32s + ucn::X3Layout::b_feat;
          EPI_LOOP(1, t_Q, 0, bf, true);
        } else {
          // round 9's MMAs are complete: Q is free -> the next tile's film MMAs may start (go 0,1, once the stager warps
          // have placed its operands); its L0 MMAs overwrite P and wait for this epilogue's reads (go 2,3)
          if (has_next) { __syncwarp(); if (lane == 0) mbar_arrive(goA); }
          head[0] = head[1] = head[2] = head[3] = 0.f;
          const float *rb = rayb_s + (s * rpt + cur.ri) * 128;
          {
            const float *bvv = f32s + ucn::X3Layout::b_vv;
            if (hf == 0) {
              epi32<2, true, 0, 7>(t_P, 0, bvv, rb, mod_row, hw.adapt, head);
              epi32<2, true, 64, 8>(t_P, 0, bvv, rb, mod_row, hw.adapt, head);
            } else {
              epi32<2, true, 32, 7>(t_P, 0, bvv, rb, mod_row, hw.adapt, head);
              epi32<2, true, 96, 8>(t_P, 0, bvv, rb, mod_row, hw.adapt, head);
            }
          }
          if (has_next) { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(goB); }
          if (hf == 1) {
            exch_s[s * 128 + row] = make_float4(head[0], head[1], head[2], head[3]);
            named_bar_arrive(bar_id, 256);
          } else {
            named_bar_sync(bar_id, 256);
            const float4 o = exch_s[s * 128 + row];
            const float *bb = f32s + ucn::X3Layout::b_adapt;
            const float ad[4] = {head[0] + o.x + bb[0], head[1] + o.y + bb[1], head[2] + o.z + bb[2], head[3] + o.w + bb[3]};
            const float u = u_cur, cf = 1.0f - u;                // models.py:177-180
            float4 out;
            out.x = 1.0f / (1.0f + __expf(-(base4[0] * cf + ad[0] * u)));
            out.y = 1.0f / (1.0f + __expf(-(base4[1] * cf + ad[1] * u)));
            out.z = 1.0f / (1.0f + __expf(-(base4[2] * cf + ad[2] * u)));
            out.w = fmaxf(ad[3] * cf + base4[3] * u, 0.0f);
            if (cur.valid) reinterpret_cast<float4 *>(raw)[cur.q] = out;
          }
        }
        tr.ev(4, r, g);
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
static int make_plan_p(int V, int S, Plan &P) {
  const ucn::X3Layout L(V);
  memset(&P, 0, sizeof(P));
  int nb = 0, r = 0;
  // units_total K16 units of one matrix: blocks of two; a0 = first unit (TS) or first K-chunk (SS: two chunks per unit)
  // wait0: go barriers the FIRST block waits for; chunked: block j waits for chunk j instead (K = 128 operands)
  auto add = [&](size_t mat_off, int units_total, int a0, int flags, int wait0, bool chunked) {
    for (int u0 = 0; u0 < units_total; u0 += 2) {
      Blk &b = P.blk[nb++];
      b.goff = (uint32_t)(mat_off + (size_t)u0 * ucn::X3Layout::kUnit);
      b.units = (uint8_t)(units_total - u0 >= 2 ? 2 : 1);
      b.a0 = (uint8_t)((flags & BF_SS) ? a0 + 2 * u0 : a0 + u0);
      b.flags = (uint8_t)flags;
      if (u0 != 0) b.flags &= ~BF_FRESH;
      b.wait = (uint8_t)(chunked ? (1 << (u0 / 2)) : (u0 == 0 ? wait0 : 0));
    }
  };
  auto end_round = [&](int a_reg) {
    P.a_reg[r] = (uint8_t)a_reg;
    P.round_begin[++r] = (uint8_t)nb;
  };
  P.round_begin[0] = 0;
  // R0: film -> Q needs the staged operands (go 0,1: arrived when both warpgroups' staging is visible and Q is free);
  //     L0 -> P additionally needs P, which the previous tile's last epilogue releases with go 2,3
  add(L.film, L.Kfilm / 16, 16, BF_FRESH | BF_DQ | BF_SS, 0x3, false);
  add(L.pts0, 4, 0, BF_FRESH | BF_SS, 0xC, false);
  end_round(0);
  for (int i = 0; i < 4; ++i) {                                       // R1..R4: L1..L4, D = Q,P,Q,P, A = the other region
    const int dq = (i & 1) == 0;
    add(L.pts14[i], 8, 0, BF_FRESH | (dq ? BF_DQ : 0), 0, true);
    end_round(dq ? 0 : 1);
  }
  add(L.pts5h, 8, 0, BF_FRESH | BF_DQ, 0, true);  end_round(0);       // R5: L5 hidden part, A = P (h4) -> Q
  add(L.pts5pe, 4, 0, BF_DQ, 0xF, false);         end_round(0);       // R6: L5 PE part, A = P[0..64) -> Q (warpgroup 1 arrives at once)
  add(L.fb, L.Kfb / 16, 0, BF_FRESH | BF_SS, 0xF, false);  end_round(0);   // R7: fb, A = image features (shared memory) -> P
  add(L.feat, 8, 0, BF_FRESH | BF_DQ, 0xF, false);  end_round(0);     // R8: feature_linear, A = P -> Q (whole: R7's epilogue reads Q)
  add(L.vvh, 8, 0, BF_FRESH, 0, true);              end_round(1);     // R9: views|view_confi, A = Q -> P
  if (r != NROUND || nb > MAXBLK) return -1;
  P.nblk = (uint8_t)nb;
  const int rpt = 127 / S + 2;                                       // rays a 128-sample tile can touch
  if (rpt > 255) return -1;
  P.rpt = (uint8_t)rpt;
  for (int nst = 5; nst >= 3; --nst) {
    uint32_t o = 2 * MODB + nst * STAGE;
    P.off_f32 = o;  o += (ucn::X3Layout::kSmemFloats * 4 + 127) & ~127u;
    P.off_exch = o; o += 2 * 128 * 16;
    P.off_rayb = o; o += 2u * rpt * 128 * 4;
    P.off_ubuf = o; o += 2u * 2 * 128 * 4;
    P.off_bars = o; o += (8 + 2 + 2 * nst + 6) * 8 + 16;
    P.smem_total = o;
    P.nst = (uint8_t)nst;
    if (o <= 232448u - 1024u) return 0;              // 1 KB of static shared memory (block tables)
  }
  return -2;
}

// Per-ray view-direction term of views|view_confi (network/models.py:166-173): the 27 PE(dir) inputs are constant along a
// ray, so their contribution to the 128 hidden columns is one 128-vector per ray, vd^T PE(dir), applied in round 9's
// epilogue.  32 rays per block: phase 1 builds the 27 encodings, phase 2 thread = column.
__global__ void __launch_bounds__(128) ray_term_kernel(const float *__restrict__ vd, const float *__restrict__ angle, int R,
                                                        float *__restrict__ out) {
  __shared__ float pe_s[32][28];
  const int r0 = blockIdx.x * 32, t = threadIdx.x;
  if (t < 32 && r0 + t < R) {
    const float a[3] = {angle[3 * (r0 + t)], angle[3 * (r0 + t) + 1], angle[3 * (r0 + t) + 2]};
    float pe[27];
    pos_enc<4>(a, pe);
#pragma unroll
    for (int k = 0; k < 27; ++k) pe_s[t][k] = pe[k];
  }
  float w[27];
#pragma unroll
  for (int k = 0; k < 27; ++k) w[k] = vd[k * 128 + t];
  __syncthreads();
  const int n = min(32, R - r0);
  for (int i = 0; i < n; ++i) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 27; ++k) acc = fmaf(w[k], pe_s[i][k], acc);
    out[(size_t)(r0 + i) * 128 + t] = acc;
  }
}

// grow-only device scratch per (device, stream, tag): launches on one stream are ordered, so consecutive calls may reuse it
static void *scratch_for(void *stream, int tag, size_t bytes) {
  struct Ent { int dev; void *stream; int tag; size_t bytes; void *buf; };
  static std::mutex mu;
  static std::vector<Ent> tab;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(mu);
  for (Ent &e : tab)
    if (e.dev == dev && e.stream == stream && e.tag == tag) {
      if (e.bytes >= bytes) return e.buf;
      cudaStreamSynchronize((cudaStream_t)stream);           // the old buffer may still be in use by the previous launch
      cudaFree(e.buf);
      e.buf = nullptr; e.bytes = 0;
      if (cudaMalloc(&e.buf, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
      e.bytes = bytes;
      return e.buf;
    }
  if (tab.size() >= 128) return nullptr;
  void *p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  tab.push_back({dev, stream, tag, bytes, p});
  return p;
}

// Scratch for the PE words of a tile (stager warps -> round 0's shared-memory operand and round 5's TMEM operand): 32 KB
// per tile, double buffered per slot, 128 KB per CTA; it lives in L2.  One buffer per (device, stream): launches on one stream are ordered.
static uint4 *pe_stash_for(void *stream, unsigned blocks) {
  return static_cast<uint4 *>(scratch_for(stream, 0, (size_t)blocks * 2 * 2 * 16 * 128 * sizeof(uint4)));   // [cta][slot][parity]
}

extern "C" int ucn_debug_set_flags_p(int flags) {
  cudaError_t e = cudaMemcpyToSymbol(g_dbg_p, &flags, sizeof(flags));
  return e == cudaSuccess ? UCN_OK : ucn_cuda_status(e, "ucn_debug_set_flags_p");
}
extern "C" int ucn_debug_set_trace_p(long long *buf) {
  cudaError_t e = cudaMemcpyToSymbol(g_trace_p, &buf, sizeof(buf));
  return e == cudaSuccess ? UCN_OK : ucn_cuda_status(e, "ucn_debug_set_trace_p");
}

int ucn_mlp_forward_x3p(const ucn_weights *w, const float *ndc_pts, const void *operand_tiles, const float *angle, int R,
                        int S, float *raw, void *stream) {
  UCN_CHECK_ARG(w->bf16x3, "ucn_mlp_forward: split-bf16 weight pack missing");
  UCN_CHECK_ARG(w->V <= 7, "ucn_mlp_forward: tensor-core modes support at most 7 source views");
  Plan P;
  const int rc = make_plan_p(w->V, S, P);
  if (rc == -2) return ucn_set_error(UCN_E_UNSUPPORTED, "ucn_mlp_forward: BF16X3_TC needs more samples per ray (shared-memory budget)");
  if (rc != 0) return ucn_set_error(UCN_E_ARG, "ucn_mlp_forward: cannot build the x3p schedule");
  const long long N = (long long)R * S;
  UCN_CHECK_ARG(N < (1ll << 31), "ucn_mlp_forward: more than 2^31 samples in one call; split the ray range");
  static int attr[kUcnMaxDevices];
  {
    cudaError_t e = ucn_ensure_dyn_smem(mlp_x3p_kernel, (int)P.smem_total, attr);
    if (e != cudaSuccess) return ucn_cuda_status(e, "mlp_x3p smem attribute");
  }
  const int sms = ucn_sm_count();
  const long long npairs = ((N + 127) / 128 + 1) / 2;
  const unsigned blocks = (unsigned)(npairs < sms ? npairs : sms);
  uint4 *stash = pe_stash_for(stream, (unsigned)sms);
  float *rayterm = static_cast<float *>(scratch_for(stream, 1, (size_t)R * 128 * sizeof(float)));
  if (!stash || !rayterm) return ucn_cuda_status(cudaErrorMemoryAllocation, "ucn_mlp_forward: scratch (PE stash, per-ray terms)");
  const ucn::X3Layout L(w->V);
  const float *f32g = reinterpret_cast<const float *>(reinterpret_cast<const unsigned char *>(w->bf16x3) + L.f32);
  ray_term_kernel<<<(R + 31) / 32, 128, 0, (cudaStream_t)stream>>>(f32g + ucn::X3Layout::vd, angle, R, rayterm);
  UCN_LAUNCH_CHECK("ray_term_kernel");
  // head weights travel as a kernel parameter (constant bank); ucn_pack_weights keeps the host copy of every pack
  HeadW hw;
  if (!ucn_headw_lookup(w->bf16x3, hw.base)) {                 // a pack not written by ucn_pack_weights: fetch once
    cudaError_t e = cudaMemcpyAsync(hw.base, f32g + ucn::X3Layout::w_base, sizeof(float) * 1024, cudaMemcpyDeviceToHost, (cudaStream_t)stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
    if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_mlp_forward: head weights");
    ucn_headw_store(w->bf16x3, hw.base);
  }
  mlp_x3p_kernel<<<blocks, kThreads, P.smem_total, (cudaStream_t)stream>>>(
      P, hw, reinterpret_cast<const unsigned char *>(w->bf16x3), w->V, ndc_pts,
      static_cast<const unsigned char *>(operand_tiles), (long long)ucn_operand_tile_bytes(w->V), rayterm, N, S, raw, stash);
  UCN_LAUNCH_CHECK("mlp_x3p_kernel");
  return UCN_OK;
}
