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


// MMA stream variants (one issuing thread, nothing else running):
//  v&1: commit to a dummy mbarrier after every 6 MMAs   v&2: real-kernel operand pattern (A_hi/A_lo 32 KB apart, B_hi/B_lo 4 KB apart,
//  three MMAs per K step)   v&4: mbarrier try_wait (already-complete barrier) before every 6 MMAs   v&8: alternate D between two accumulators
__global__ void __launch_bounds__(128, 1) k_stream(int v, int iters, long long *cycles) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint32_t tptr;
  __shared__ __align__(8) uint64_t bar, dummy[8], done;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_init(smem_u32(&done), 1); for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&dummy[i]), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
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
  if (warp == 1 && lane == 0) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&done)) : "memory");   // phase 0 of `done` completes: try_wait(parity 0) on it always succeeds
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(128 >> 3) << 17) | ((128u >> 4) << 24);
    for (int blk = 0; blk < iters / 6; ++blk) {
      if (v & 4) mbar_wait(smem_u32(&done), 0);
      const uint32_t ring = sbase + 128 * 1024 + (blk % 4) * 16384, slot = sbase + (blk & 1) * 65536;
      const uint32_t d = tb + ((v & 8) ? (blk & 1) * 256 : 0);
      for (int u = 0; u < 2; ++u) {
        if (v & 2) {
          const uint32_t ah = slot + ((blk % 4) * 4 + 2 * u) * 2048, al = ah + 32768, bh = ring + u * 8192, bl = bh + 4096;
          umma_bf16(d, umma_desc(ah, 2048, 128), umma_desc(bh, 2048, 128), idesc, 1u);
          umma_bf16(d, umma_desc(al, 2048, 128), umma_desc(bh, 2048, 128), idesc, 1u);
          umma_bf16(d, umma_desc(ah, 2048, 128), umma_desc(bl, 2048, 128), idesc, 1u);
        } else {
          for (int j = 0; j < 3; ++j) {
            const int i = blk * 6 + u * 3 + j;
            umma_bf16(d, umma_desc(sbase + (i & 7) * 4096, 2048, 128), umma_desc(sbase + 64 * 1024 + (i & 7) * 4096, 2048, 128), idesc, 1u);
          }
        }
      }
      if (v & 1) umma_commit(smem_u32(&dummy[blk & 7]));
    }
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
  }
  __syncwarp();
  long long t1 = clock64();
  if (lane == 0) cycles[blockIdx.x * 4 + warp] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512u) : "memory");
}
int main() {
  long long *cyc; CK(cudaMalloc(&cyc, 148 * 4 * 8));
  CK(cudaFuncSetAttribute(k_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  long long h[148 * 4];
  for (int v = 0; v < 16; ++v) {
    const int iters = 12000;
    for (int rep = 0; rep < 2; ++rep) { k_stream<<<148, 128, 200 * 1024>>>(v, iters, cyc); CK(cudaDeviceSynchronize()); }
    CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    printf("variant %2d (commit/6 %d, real operand pattern %d, try_wait/6 %d, alternate D %d): %.1f cycles per MMA\n", v, v & 1, (v >> 1) & 1, (v >> 2) & 1, (v >> 3) & 1, (double)h[1] / iters);
  }
  return 0;
}
