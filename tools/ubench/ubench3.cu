#include <string.h>
#include <math.h>
// Scratch microbenchmarks that size the K3 design (not product code): TMEM read rate, L2->smem bulk-copy
// rate when every SM streams the same weight pack, tcgen05.mma issue rate with/without concurrent smem stores.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(b) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory"); } while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                 "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_bf16(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }


__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}"
               ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc), "r"(0u) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
// correctness: D[128][128] = A[128][32] * B[128][32]^T with A in TMEM (packed bf16 pairs), B in smem (no-swizzle K-major)
__global__ void __launch_bounds__(192, 1) k_ts_check(const uint16_t *A, const uint16_t *B, float *D) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint32_t tptr;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  for (int i = threadIdx.x; i < 128 * 32; i += blockDim.x) {
    const int n = i / 32, k = i % 32;
    reinterpret_cast<uint16_t *>(smem)[((k / 8) * 128 + n) * 8 + (k % 8)] = B[n * 32 + k];
  }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tptr)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tptr;
  if (warp < 4) {
    const int row = warp * 32 + lane;
    uint32_t r[16];
    for (int j = 0; j < 16; ++j) r[j] = (uint32_t)A[row * 32 + 2 * j] | ((uint32_t)A[row * 32 + 2 * j + 1] << 16);
    tmem_st16(tb + ((uint32_t)(warp * 32) << 16) + 0, r);
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp == 4 && lane == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 3) << 17) | ((128u >> 4) << 24);
    for (int ks = 0; ks < 2; ++ks)
      umma_ts(tb + 256, tb + 8 * ks, umma_desc(sbase + ks * 2 * 2048, 2048, 128), idesc, ks ? 1u : 0u);
    umma_commit(smem_u32(&bar));
  }
  __syncwarp();
  if (warp < 4) {
    mbar_wait(smem_u32(&bar), 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < 128; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(tb + ((uint32_t)(warp * 32) << 16) + 256 + c0, r);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int j = 0; j < 32; ++j) D[row * 128 + c0 + j] = __uint_as_float(r[j]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512u) : "memory");
}

// rate: TS-mode MMA stream (A in TMEM) alone / with nw warps doing tmem ld+st / with nw warps doing smem 16B loads+stores
__global__ void __launch_bounds__(1024, 1) k_ts_rate(int mode, int nw, int iters, int mma_iters, int burst, long long *cycles, float *sink) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint32_t tptr;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tptr)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tptr;
  long long t0 = clock64();
  float acc = 0.f;
  if (warp == 0) {
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 3) << 17) | ((128u >> 4) << 24);
      uint32_t ph = 0;
      for (int i = 0; i < mma_iters; ++i) {
        const uint32_t b = sbase + (i & 7) * 4096;
        // D at cols 0..127 / 256..383, A at 128.. / 256+128..  (columns 384.. are the st scratch in modes 3/4: keep D off them)
        umma_ts(tb + (i & 1) * 128, tb + 256 + (i & 7) * 8, umma_desc(b, 2048, 128), idesc, 1u);
        if (burst && (i % burst) == burst - 1) { umma_commit(smem_u32(&bar)); mbar_wait(smem_u32(&bar), ph); ph ^= 1; }
      }
      umma_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), ph);
    }
    __syncwarp();
  } else if (warp >= 4 && warp < 4 + nw) {
    const uint32_t tl = tb + ((uint32_t)((warp & 3) << 5) << 16);
    if (mode == 1) {          // epilogue-like TMEM traffic: ld 32 cols of D, st 16 cols of A
      for (int i = 0; i < iters / 2; ++i) {
        uint32_t r[32];
        tmem_ld32(tl + ((i & 3) * 32) + ((warp >> 2) & 1) * 256, r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        uint32_t w[16];
        for (int j = 0; j < 16; ++j) w[j] = r[2 * j] ^ r[2 * j + 1];
        tmem_st16(tl + 448 + ((i & 3) * 16), w);   // scratch columns 448..511 (not used by the MMA stream)
        acc += __uint_as_float(r[0]);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    } else if (mode == 3) {   // st only
      for (int i = 0; i < iters / 2; ++i) {
        uint32_t w[16];
        for (int j = 0; j < 16; ++j) w[j] = i + j;
        tmem_st16(tl + 384 + ((i & 7) * 16), w);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    } else if (mode == 4) {   // ld32 + wait + 2 x st16, wait::st every 4 iterations (epilogue-like batches)
      for (int i = 0; i < iters / 2; ++i) {
        uint32_t r[32];
        tmem_ld32(tl + ((i & 3) * 32) + ((warp >> 2) & 1) * 256, r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        uint32_t w[16], v[16];
        for (int j = 0; j < 16; ++j) { w[j] = r[2 * j] ^ r[2 * j + 1]; v[j] = r[2 * j] + r[2 * j + 1]; }
        tmem_st16(tl + 384 + ((i & 3) * 32), w);
        tmem_st16(tl + 384 + ((i & 3) * 32) + 16, v);
        if ((i & 3) == 3) asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        acc += __uint_as_float(r[0]);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    } else if (mode == 2) {   // smem traffic: LDS.128 + STS.128 per iteration
      const uint32_t base = sbase + 64 * 1024 + ((warp - 4) & 15) * 4096 + lane * 16;
      for (int i = 0; i < iters; ++i) {
        uint32_t a, b, c, d;
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(base + (i & 7) * 512) : "memory");
        acc += __uint_as_float(a ^ b ^ c ^ d);
      }
    }
  }
  long long t1 = clock64();
  if (lane == 0) cycles[blockIdx.x * 32 + warp] = t1 - t0;
  if (acc == 123.f) sink[0] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512u) : "memory");
}

static uint16_t f2bf(float f) { uint32_t u; memcpy(&u, &f, 4); u += 0x7fff + ((u >> 16) & 1); return (uint16_t)(u >> 16); }
static float bf2f(uint16_t h) { uint32_t u = (uint32_t)h << 16; float f; memcpy(&f, &u, 4); return f; }

int main() {
  static uint16_t hA[128 * 32], hB[128 * 32];
  static float hD[128 * 128];
  srand(1);
  for (int i = 0; i < 128 * 32; ++i) { hA[i] = f2bf((rand() % 2001 - 1000) / 500.0f); hB[i] = f2bf((rand() % 2001 - 1000) / 500.0f); }
  uint16_t *dA, *dB; float *dD;
  CK(cudaMalloc(&dA, sizeof(hA))); CK(cudaMalloc(&dB, sizeof(hB))); CK(cudaMalloc(&dD, sizeof(hD)));
  CK(cudaMemcpy(dA, hA, sizeof(hA), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB, sizeof(hB), cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(k_ts_check, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  k_ts_check<<<1, 192, 64 * 1024>>>(dA, dB, dD);
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(hD, dD, sizeof(hD), cudaMemcpyDeviceToHost));
  double maxerr = 0;
  for (int m = 0; m < 128; ++m) for (int n = 0; n < 128; ++n) {
    double ref = 0;
    for (int k = 0; k < 32; ++k) ref += (double)bf2f(hA[m * 32 + k]) * bf2f(hB[n * 32 + k]);
    double e = fabs(ref - hD[m * 128 + n]);
    if (e > maxerr) maxerr = e;
  }
  printf("TS-mode check (A in TMEM as packed bf16 pairs, col j = k 2j,2j+1): max abs err %.3e  D[0][0]=%f D[5][7]=%f\n", maxerr, hD[0], hD[5 * 128 + 7]);

  long long *cyc; float *sink;
  CK(cudaMalloc(&cyc, 148 * 32 * 8)); CK(cudaMalloc(&sink, 4));
  CK(cudaFuncSetAttribute(k_ts_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  long long h[148 * 32];
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  struct Cfg { int mode, nw, mma_iters, burst; const char *what; };
  const Cfg cfgs[] = {{3, 8, 0, 0, "st16 only, no MMA"}, {3, 16, 0, 0, "st16 only, no MMA"}, {4, 8, 0, 0, "ld32+2xst16, no MMA"}, {4, 16, 0, 0, "ld32+2xst16, no MMA"},
                      {3, 8, 8000, 0, "st16 + MMA stream (deep queue)"}, {3, 8, 8000, 24, "st16 + MMA bursts of 24"}, {3, 8, 8000, 6, "st16 + MMA bursts of 6"},
                      {4, 8, 8000, 0, "ld32+2xst16 + MMA stream"}, {4, 8, 8000, 24, "ld32+2xst16 + MMA bursts of 24"}, {4, 16, 8000, 24, "ld32+2xst16 + MMA bursts of 24"},
                      {1, 8, 8000, 24, "ld32+st16 (scratch 448) + MMA bursts of 24"}};
  for (const Cfg &c : cfgs) {
    const int iters = 4000;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      k_ts_rate<<<148, 32 * (4 + c.nw), 200 * 1024>>>(c.mode, c.nw, iters, c.mma_iters, c.burst, cyc, sink);
      cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    printf("%-44s nw %2d: %.3f ms  mma %.1f cyc/mma  epilogue warp %.1f cyc/iter\n", c.what, c.nw, ms, c.mma_iters ? (double)h[0] / c.mma_iters : 0.0, (double)h[4] / (iters / 2));
  }
  return 0;
}
