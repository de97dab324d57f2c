// ucn_composite.cu -- K4: alpha compositing as a warp-per-ray multiplicative (transmittance) scan.
//
// replaces (reference): network/renderer.py:25-36 raw2alpha, :109-140 raw2outputs.
//   alpha = 1 - exp(-sigma)   (dists are ignored by the reference, quirk Q3)
//   T_i   = prod_{j<i} (1 - alpha_j + 1e-10)      weights = alpha * T
//   rgb_map = sum w * rgb     depth_map = sum w * z     acc = sum w
//
// Lanes take samples i = 32*g + lane (coalesced float4 loads of raw); each group of 32 does an
// inclusive product scan with 5 shuffles and carries the running transmittance to the next group.
// Roofline: HBM-bound, 20*S + 16 algorithmic bytes per ray.
#include "ucn_common.cuh"

namespace {

constexpr int kWarps = 8;

__global__ void __launch_bounds__(kWarps * 32)
composite_kernel(const float4 *__restrict__ raw, const float *__restrict__ z, int R, int S, int white_bkgd,
                 float *__restrict__ rgb, float *__restrict__ depth, float *__restrict__ acc_out,
                 float *__restrict__ weights) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (r >= R) return;
  float T = 1.0f, cr = 0.f, cg = 0.f, cb = 0.f, cd = 0.f, ca = 0.f;
  for (int base = 0; base < S; base += 32) {
    const int i = base + lane;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    float zi = 0.f;
    if (i < S) {
      s = __ldg(raw + (size_t)r * S + i);
      zi = __ldg(z + (size_t)r * S + i);
    }
    const float alpha = i < S ? 1.0f - expf(-s.w) : 0.0f;
    float p = i < S ? (1.0f - alpha) + 1e-10f : 1.0f;         // inclusive product of (1-alpha+eps)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float up = __shfl_up_sync(0xffffffffu, p, o);
      if (lane >= o) p *= up;
    }
    float excl = __shfl_up_sync(0xffffffffu, p, 1);
    if (lane == 0) excl = 1.0f;
    const float w = alpha * (T * excl);
    if (weights && i < S) weights[(size_t)r * S + i] = w;
    cr += w * s.x; cg += w * s.y; cb += w * s.z; cd += w * zi; ca += w;
    T *= __shfl_sync(0xffffffffu, p, 31);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cr += __shfl_xor_sync(0xffffffffu, cr, o); cg += __shfl_xor_sync(0xffffffffu, cg, o);
    cb += __shfl_xor_sync(0xffffffffu, cb, o); cd += __shfl_xor_sync(0xffffffffu, cd, o);
    ca += __shfl_xor_sync(0xffffffffu, ca, o);
  }
  if (lane == 0) {
    if (white_bkgd) { const float bg = 1.0f - ca; cr += bg; cg += bg; cb += bg; }   // renderer.py:138-139
    rgb[3 * r] = cr; rgb[3 * r + 1] = cg; rgb[3 * r + 2] = cb;
    depth[r] = cd;
    if (acc_out) acc_out[r] = ca;
  }
}

}  // namespace

extern "C" int ucn_composite(const float *raw, const float *z, int R, int S, int white_bkgd, float *rgb, float *depth,
                             float *acc, float *weights, void *stream) {
  UCN_CHECK_ARG(raw && z && rgb && depth, "ucn_composite: null argument");
  UCN_CHECK_ARG(S > 0, "ucn_composite: S <= 0");
  if (R <= 0) return UCN_OK;
  composite_kernel<<<(R + kWarps - 1) / kWarps, kWarps * 32, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4 *>(raw), z, R, S, white_bkgd, rgb, depth, acc, weights);
  UCN_LAUNCH_CHECK("composite_kernel");
  return UCN_OK;
}
