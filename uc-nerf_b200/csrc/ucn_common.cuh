// ucn_common.cuh -- shared device helpers for libucnerf.so (sm_100a only).
//
// Arithmetic-order contract: everything that feeds a sample POSITION or a gather INDEX is written
// with explicit round-to-nearest intrinsics (__fmul_rn/__fadd_rn/__fdiv_rn never contract into FMA)
// and fmaf() exactly where the reference's K=3 sgemm fuses, so the kernels are bit-identical to the
// oracle (oracle/ucnerf_oracle.c), which is pinned bit-exact to the reference's own output.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ucnerf.h"

#define UCN_CHECK_ARG(cond, msg)                 \
  do {                                           \
    if (!(cond)) return ucn_set_error(UCN_E_ARG, msg); \
  } while (0)

int ucn_set_error(int code, const char *msg);
int ucn_cuda_status(cudaError_t e, const char *where);
void ucn_count_launch(int n = 1);
void ucn_headw_store(const void *x3_pack, const float *w1024);   // ucn_pack.cu: host copy of K3's head weights per pack
bool ucn_headw_lookup(const void *x3_pack, float *w1024);
int ucn_sm_count();            // multiprocessors of the CURRENT device (cached per device)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: `done` is the caller's static per-device
// table of the byte counts already applied (one process may drive several GPUs).
constexpr int kUcnMaxDevices = 64;
template <typename Kernel>
inline cudaError_t ucn_ensure_dyn_smem(Kernel k, int bytes, int (&done)[kUcnMaxDevices]) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kUcnMaxDevices) return cudaErrorInvalidDevice;
  if (done[dev] >= bytes) return cudaSuccess;
  e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) done[dev] = bytes;
  return e;
}

#define UCN_LAUNCH_CHECK(where)                                    \
  do {                                                             \
    ucn_count_launch();                                            \
    cudaError_t e__ = cudaGetLastError();                          \
    if (e__ != cudaSuccess) return ucn_cuda_status(e__, where);    \
  } while (0)

namespace ucn {

// y_j = sum_k x_k * M[j*ld + k], accumulated as the k-ordered FMA chain sgemm uses for K=3.
__device__ __forceinline__ void mm3t(const float *__restrict__ M, int ld, const float x[3], float y[3]) {
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    float acc = __fmul_rn(x[0], M[j * ld + 0]);
    acc = fmaf(x[1], M[j * ld + 1], acc);
    acc = fmaf(x[2], M[j * ld + 2], acc);
    y[j] = acc;
  }
}

// world -> camera -> pixel -> [0,1]-normalised xy + projective depth
// (reference: utils/utils.py:333-347).
__device__ __forceinline__ void project(const float *__restrict__ w2c, const float *__restrict__ K, float iw,
                                        float ih, const float p[3], float &nx, float &ny, float &pz) {
  float pc[3], pp[3];
  mm3t(w2c, 4, p, pc);
  pc[0] = __fadd_rn(pc[0], w2c[3]);
  pc[1] = __fadd_rn(pc[1], w2c[7]);
  pc[2] = __fadd_rn(pc[2], w2c[11]);
  if (fabsf(pc[2]) < 1e-4f) pc[2] = 1e-4f;
  mm3t(K, 3, pc, pp);
  nx = __fdiv_rn(__fadd_rn(__fdiv_rn(pp[0], pp[2]), 0.0f), iw);
  ny = __fdiv_rn(__fadd_rn(__fdiv_rn(pp[1], pp[2]), 0.0f), ih);
  pz = pp[2];
}

// ATen grid_sampler coordinate maps (ATen/native/GridSampler.h:27-36,58-60), border padding.
// (x / 2 is written x * 0.5f: the two are the same correctly-rounded operation for every float, subnormals included,
//  and the multiply is one instruction where an IEEE divide is ~12 -- 17 of them per sample in K2.)
__device__ __forceinline__ float unnorm(float g, int size, bool align_corners) {
  if (align_corners) return __fmul_rn(__fmul_rn(__fadd_rn(g, 1.0f), 0.5f), (float)(size - 1));
  return __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(g, 1.0f), (float)size), 1.0f), 0.5f);
}
__device__ __forceinline__ float clipc(float x, int size) { return fminf((float)(size - 1), fmaxf(x, 0.0f)); }

// grid value the reference feeds grid_sample: c*2 - 1 (utils/utils.py:775,852-857)
__device__ __forceinline__ float to_grid(float c) { return __fsub_rn(__fmul_rn(c, 2.0f), 1.0f); }

struct Lerp1 {   // floor index + the two weights along one axis
  int i0;
  float w0, w1;  // weight of i0 and of i0+1
  bool has1;     // i0+1 inside the map
};
__device__ __forceinline__ Lerp1 axis(float g, int size, bool align) {
  const float x = clipc(unnorm(g, size, align), size);
  const float f = floorf(x);
  Lerp1 a;
  a.i0 = (int)f;
  a.w1 = __fsub_rn(x, f);
  a.w0 = __fsub_rn(__fadd_rn(f, 1.0f), x);
  a.has1 = a.i0 + 1 < size;
  return a;
}

__device__ __forceinline__ float4 ldg4(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }

// Ray origin/direction for pixel (ys, xs) (reference: utils/utils.py:259-266).
__device__ __forceinline__ void ray_dir(const ucn_frame &f, float ys, float xs, float d[3]) {
  float dirs[3] = {__fdiv_rn(__fsub_rn(xs, f.K[2]), f.K[0]), __fdiv_rn(__fsub_rn(ys, f.K[5]), f.K[4]), 1.0f};
  mm3t(f.c2w, 4, dirs, d);
}

__device__ __forceinline__ void stage_near_far(const ucn_frame &f, float ys, float xs, float nr[3], float fr[3]) {
  const int py = (int)ys, px = (int)xs;          // .long(): trunc toward zero
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int div = k == 0 ? 4 : (k == 1 ? 2 : 1);
    const int qy = py / div, qx = px / div;      // torch.div(..., rounding_mode='trunc')
    const float2 nf = __ldg(reinterpret_cast<const float2 *>(f.nearfar[k]) + (size_t)qy * f.w[k] + qx);
    nr[k] = nf.x;
    fr[k] = nf.y;
  }
}

__device__ __forceinline__ void pixel_of_ray(const ucn_frame &f, const float *__restrict__ pix_yx, int ray0, int R,
                                             int r, float &ys, float &xs) {
  if (pix_yx) {
    ys = pix_yx[r];
    xs = pix_yx[R + r];
  } else {
    const int g = ray0 + r;
    ys = (float)(g / f.W);
    xs = (float)(g % f.W);
  }
}

}  // namespace ucn
