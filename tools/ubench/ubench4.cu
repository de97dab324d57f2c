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

// mode 0: TMEM read rate with nw warps.  mode 1: MMA rate (N given) alone.  mode 2: MMA + nw warps of 16B smem stores.
// mode 3: MMA + nw warps of tmem loads.
__global__ void __launch_bounds__(1024, 1) k_tc(int mode_, int nw, int iters, int N, long long *cycles, float *sink, const unsigned char *src, int pace, int dsame, int rnd) {
  const int mode = mode_ & 7; const bool fill = mode_ & 8;
  __shared__ __align__(8) uint64_t full[4];
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint32_t tptr;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) {
    uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    // two bf16 values in [-2,2) with random mantissas when rnd, else the constant 1.0 pattern
    const uint32_t lo = 0x3f00u | ((h & 1) << 15) | ((h >> 1) & 0xff), hi = 0x3f00u | (((h >> 9) & 1) << 15) | ((h >> 10) & 0xff);
    reinterpret_cast<uint32_t *>(smem)[i] = rnd ? (lo | (hi << 16)) : 0x3c003c00u;
  }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&full[i]), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
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
  if (warp == 0 && (mode >= 1)) {
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
      for (int i = 0; i < iters; ++i) {
        const uint32_t a = sbase + (i & 7) * 4096, b = sbase + 64 * 1024 + (i & 7) * 4096;
        umma_bf16(tb + (dsame ? 0 : (i & 1) * 256), umma_desc(a, 2048, 128), umma_desc(b, N * 16, 128), idesc, 1u);
      }
      umma_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), 0);
    }
    __syncwarp();
  } else if (warp == 1 && fill) {
    if (lane == 0) {   // stream 16 KB stages into smem [144K, 208K) for roughly the duration of the MMA stream
      const int total = iters / 6;
      for (int c = 0; c < total; ++c) {
        const int st = c & 3;
        if (c >= 4) mbar_wait(smem_u32(&full[st]), ((c >> 2) - 1) & 1);
        mbar_expect_tx(smem_u32(&full[st]), 16384);
        bulk_g2s(sbase + 144 * 1024 + st * 16384, src + (size_t)(c % 32) * 16384, 16384, smem_u32(&full[st]));
        if (pace) { long long t = clock64(); while (clock64() - t < pace) {} }
      }
      for (int c = total; c < total + 4 && c >= 4; ++c) mbar_wait(smem_u32(&full[c & 3]), ((c >> 2) - 1) & 1);
    }
    __syncwarp();
  } else if (warp >= 4 && warp < 4 + nw) {
    const uint32_t tl = tb + ((uint32_t)((warp & 3) << 5) << 16);
    if (mode == 0 || mode == 3) {
      for (int i = 0; i < iters; ++i) {
        uint32_t r[32];
        tmem_ld32(tl + ((i * 32 + (warp >> 2) * 64) & 511 & ~31), r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc += __uint_as_float(r[0]) + __uint_as_float(r[31]);
      }
    } else if (mode == 2) {
      const uint32_t base = sbase + 128 * 1024 + ((warp - 4) & 7) * 4096 + lane * 16;
      for (int i = 0; i < iters; ++i) {
        asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + (i & 7) * 512), "r"(i) : "memory");
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

// every CTA streams the same `bytes` buffer `reps` times through a ring of nst stages of `stage` bytes
__global__ void __launch_bounds__(128, 1) k_bulk(const unsigned char *src, int bytes, int stage, int nst, int reps, long long *cycles) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full[16];
  const uint32_t sbase = smem_u32(smem);
  if (threadIdx.x == 0) { for (int i = 0; i < nst; ++i) mbar_init(smem_u32(&full[i]), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  long long t0 = clock64();
  if (threadIdx.x == 0) {
    const int nblk = bytes / stage, total = nblk * reps;
    int issued = 0;
    for (; issued < nst && issued < total; ++issued) {
      mbar_expect_tx(smem_u32(&full[issued]), stage);
      bulk_g2s(sbase + issued * stage, src + (size_t)(issued % nblk) * stage, stage, smem_u32(&full[issued]));
    }
    for (int c = 0; c < total; ++c) {
      const int st = c % nst;
      mbar_wait(smem_u32(&full[st]), (c / nst) & 1);
      if (issued < total) {
        mbar_expect_tx(smem_u32(&full[st]), stage);
        bulk_g2s(sbase + st * stage, src + (size_t)(issued % nblk) * stage, stage, smem_u32(&full[st]));
        ++issued;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
  long long *cyc; float *sink; unsigned char *src;
  CK(cudaMalloc(&cyc, 148 * 32 * 8)); CK(cudaMalloc(&sink, 4)); CK(cudaMalloc(&src, 1 << 20)); CK(cudaMemset(src, 0, 1 << 20));
  CK(cudaFuncSetAttribute(k_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024));
  CK(cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  long long h[148 * 32];
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int pace = 0, dsame = 0, rnd = 0;
  auto run_tc = [&](int mode, int nw, int iters, int N, const char *what) {
    for (int rep = 0; rep < 2; ++rep) {
      CK(cudaMemset(cyc, 0, 148 * 32 * 8));
      cudaEventRecord(e0);
      k_tc<<<148, 32 * (4 + (nw > 0 ? nw : 1)), 208 * 1024>>>(mode, nw, iters, N, cyc, sink, src, pace, dsame, rnd);
      cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    long long c0 = h[0], c4 = h[4];
    printf("%-46s mode %d nw %2d N %3d iters %6d: %.3f ms  warp0 %lld cyc (%.1f/iter)  warp4 %lld cyc (%.1f/iter)\n", what, mode, nw, N, iters, ms, c0, (double)c0 / iters, c4, (double)c4 / iters);
  };
  run_tc(1, 0, 8000, 128, "mma alone");
  rnd = 1; dsame = 1; run_tc(1, 0, 8000, 128, "mma alone, RANDOM data, same acc"); run_tc(1, 0, 80000, 128, "mma alone, RANDOM data, same acc, 10x longer"); run_tc(1, 0, 400000, 128, "mma alone, RANDOM data, 50x longer"); rnd = 0; dsame = 0; run_tc(1, 0, 400000, 128, "mma alone, const data, 50x longer");
  dsame = 1; run_tc(1, 0, 8000, 128, "mma alone, same accumulator"); run_tc(1, 0, 8000, 256, "mma alone N256, same accumulator"); run_tc(1, 0, 8000, 64, "mma alone N64, same accumulator"); dsame = 0;
  run_tc(1 | 8, 0, 8000, 128, "mma + bulk fills 16KB back-to-back");
  pace = 300; run_tc(1 | 8, 0, 8000, 128, "mma + bulk fills 16KB every ~400 cyc");
  pace = 700; run_tc(1 | 8, 0, 8000, 128, "mma + bulk fills 16KB every ~800 cyc");
  pace = 300; run_tc(2 | 8, 8, 8000, 128, "mma + fills/400 + 8 warps STS");
  pace = 300; run_tc(3 | 8, 8, 8000, 128, "mma + fills/400 + 8 warps tmem ld");
  return 0;
}
