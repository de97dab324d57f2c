// ucn_rays.cu -- K1: fused ray generation + depth-bounded sampling + target-view NDC, and the
// per-ray view-direction feature.  One warp per ray; the S depth candidates live in shared memory.
//
// replaces (reference): utils/utils.py:217-271 get_rays_mvs, :393-397 sample_points_uniform,
//   :600-739 build_rays_test, :323-373 get_ndc_coordinate; network/renderer.py:163-174,232-239.
//
// Roofline: HBM-bound.  Algorithmic bytes per ray = 4*S (z) + 36 (rays_d + six near/far reads) in
// whole-frame mode; the optional drop-in outputs (pts, 4x ndc) add 60*S.
#include "ucn_common.cuh"

namespace {

constexpr int kWarpsPerBlock = 8;

// Philox4x32-10 (Salmon et al. 2011), one call -> 4 uniforms in [0,1).  Used only when the caller
// passes no t_rand: a whole-frame kernel cannot replay torch's per-chunk generator stream
// (SURVEY.md 7.4-2), so parity tests always pass t_rand explicitly.
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += 0x9E3779B9u;
    key.y += 0xBB67AE85u;
  }
  return ctr;
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
raygen_sample_kernel(const __grid_constant__ ucn_frame f, const float *__restrict__ pix_yx, int ray0, int R,
                     const float *__restrict__ t_vals, const float *__restrict__ t_rand, uint64_t seed, int P,
                     float *__restrict__ z_out, float *__restrict__ rays_d_out, float *__restrict__ pts_out,
                     float *__restrict__ ndc_out) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * kWarpsPerBlock + warp;
  if (r >= R) return;
  float *zs = smem + (size_t)warp * 2 * P;                       // [P] sorted candidates | [P] the three sweeps
  const int S = f.S, n = S / 3;

  float ys, xs;
  ucn::pixel_of_ray(f, pix_yx, ray0, R, r, ys, xs);
  float d[3], nr[3], fr[3];
  ucn::ray_dir(f, ys, xs, d);
  ucn::stage_near_far(f, ys, xs, nr, fr);

  // three uniform sweeps, near*(1-t) + far*t with separate roundings (utils/utils.py:396), computed ONCE into shared
  // memory (sw[k*n + j]): the rank-merge below probes them ~12 times per element, and a probe is then one LDS instead of
  // a t_vals load + 5 roundings + selects
  float *sw = zs + P;
  auto sweep = [&](int k, int j) {           // (selects, not indexing: nr/fr stay in registers)
    const float t = __ldg(t_vals + j);
    const float nk = k == 0 ? nr[0] : (k == 1 ? nr[1] : nr[2]), fk = k == 0 ? fr[0] : (k == 1 ? fr[1] : fr[2]);
    return __fadd_rn(__fmul_rn(nk, __fsub_rn(1.0f, t)), __fmul_rn(fk, t));
  };
#pragma unroll
  for (int k = 0; k < 3; ++k)
    for (int j = lane; j < n; j += 32) sw[k * n + j] = sweep(k, j);
  __syncwarp();
  // torch.sort of their concatenation (utils/utils.py:698-700), values only.  Each sweep is non-decreasing (checked
  // below), so the sorted position of element j of sweep k is j + the number of smaller elements in the other two
  // sweeps: two binary searches per element instead of a 28-stage bitonic network.  Ties between sweeps are
  // broken by sweep index (<= against earlier sweeps, < against later ones), which makes the ranks a permutation;
  // equal values are interchangeable in a value-only sort.
  bool monotone = true;
#pragma unroll
  for (int k = 0; k < 3; ++k)
    for (int j = lane; j + 1 < n; j += 32) monotone = monotone && (sw[k * n + j] <= sw[k * n + j + 1]);
  if (__all_sync(0xffffffffu, monotone)) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
    for (int j = lane; j < n; j += 32) {
      const float v = sw[k * n + j];
      int rank = j;
#pragma unroll
      for (int m = 0; m < 3; ++m) {
        if (m == k) continue;
        // number of elements of sweep m that sort before v: first index with sweep(m, idx) > v (m < k) or >= v (m > k)
        const float *sm = sw + m * n;
        int lo = 0, hi = n;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          const float w = sm[mid];
          const bool before = m < k ? (w <= v) : (w < v);
          if (before) lo = mid + 1; else hi = mid;
        }
        rank += lo;
      }
      zs[rank] = v;
    }
    __syncwarp();
  } else {
  // degenerate sweeps (never seen with near < far; kept for exactness): bitonic sort of P (power of two) values
  for (int i = lane; i < P; i += 32) {
    float v = __int_as_float(0x7f800000);   // +inf padding sorts to the end
    if (i < S) {
      const int k = i / n, j = i - k * n;
      v = sweep(k, j);
    }
    zs[i] = v;
  }
  __syncwarp();
  for (int k2 = 2; k2 <= P; k2 <<= 1) {
    for (int j = k2 >> 1; j > 0; j >>= 1) {
      for (int i = lane; i < P; i += 32) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const float a = zs[i], b = zs[ixj];
          const bool up = (i & k2) == 0;
          if ((a > b) == up) { zs[i] = b; zs[ixj] = a; }
        }
      }
      __syncwarp();
    }
  }
  }

  if (rays_d_out && lane < 3) rays_d_out[(size_t)r * 3 + lane] = d[lane];
  const float o[3] = {f.c2w[3], f.c2w[7], f.c2w[11]};
  const float iw = (float)(f.W - 1), ih = (float)(f.H - 1);
  const size_t plane = (size_t)R * S * 3;
  const int g = pix_yx ? r : ray0 + r;                         // RNG counter: global ray id
  // a lane owns 4 consecutive samples: ONE Philox4x32-10 call (counter = sample / 4) yields their 4 uniforms -- the same
  // uniform per (ray, sample) as a call per sample, a quarter of the integer multiplies
  for (int i4 = 4 * lane; i4 < S; i4 += 128) {
    uint32_t w4[4] = {0u, 0u, 0u, 0u};
    if (!t_rand) {
      const uint4 rnd = philox4x32(make_uint4((uint32_t)g, (uint32_t)(i4 >> 2), 0u, 0u),
                                   make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
      w4[0] = rnd.x; w4[1] = rnd.y; w4[2] = rnd.z; w4[3] = rnd.w;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
    const int i = i4 + e;
    if (i >= S) break;
    const float zc = zs[i];
    const float lower = i == 0 ? zc : __fmul_rn(0.5f, __fadd_rn(zc, zs[i - 1]));
    const float upper = i == S - 1 ? zc : __fmul_rn(0.5f, __fadd_rn(zs[i + 1], zc));
    const float u = t_rand ? __ldg(t_rand + (size_t)r * S + i) : u01(w4[e]);
    const float zj = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), u));      // :706
    const size_t q = (size_t)r * S + i;
    z_out[q] = zj;
    if (pts_out || ndc_out) {
      const float p[3] = {__fadd_rn(o[0], __fmul_rn(zj, d[0])), __fadd_rn(o[1], __fmul_rn(zj, d[1])),
                          __fadd_rn(o[2], __fmul_rn(zj, d[2]))};                   // :716
      if (pts_out) { pts_out[3 * q] = p[0]; pts_out[3 * q + 1] = p[1]; pts_out[3 * q + 2] = p[2]; }
      if (ndc_out) {
        float nx, ny, pz;
        ucn::project(f.w2c, f.K, iw, ih, p, nx, ny, pz);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float nn = k < 3 ? nr[k] : f.near, ff = k < 3 ? fr[k] : f.far;
          float *dst = ndc_out + k * plane + 3 * q;
          dst[0] = nx; dst[1] = ny;
          dst[2] = __fdiv_rn(__fsub_rn(pz, nn), __fsub_rn(ff, nn));
        }
      }
    }
    }
  }
}

__global__ void dir_feature_kernel(const __grid_constant__ ucn_frame f, const float *__restrict__ rays_d, int R,
                                   int ray_offset, float *__restrict__ angle) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  const float d0 = rays_d[3 * r], d1 = rays_d[3 * r + 1], d2 = rays_d[3 * r + 2];
  const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1)), __fmul_rn(d2, d2)));
  const float u[3] = {__fdiv_rn(d0, nrm), __fdiv_rn(d1, nrm), __fdiv_rn(d2, nrm)};
  float a[3];
  ucn::mm3t((r + ray_offset) < f.dir_switch ? f.dir_rot0 : f.dir_rot1, 3, u, a);
  angle[3 * r] = a[0]; angle[3 * r + 1] = a[1]; angle[3 * r + 2] = a[2];
}

// data/ray_utils.py:152-197 ray_marcher: uniform (or disparity-uniform) samples between per-ray near/far, optional
// stratified jitter, world points.  One thread per (ray, sample); every product and sum rounds on its own, as the
// reference's separate elementwise kernels do.
__device__ __forceinline__ float marcher_z(float near, float far, float t, int lindisp) {
  const float omt = __fsub_rn(1.0f, t);
  if (!lindisp) return __fadd_rn(__fmul_rn(near, omt), __fmul_rn(far, t));
  return __fdiv_rn(1.0f, __fadd_rn(__fmul_rn(__fdiv_rn(1.0f, near), omt), __fmul_rn(__fdiv_rn(1.0f, far), t)));
}
__global__ void ray_marcher_kernel(const float *__restrict__ rays, int R, int S, const float *__restrict__ t_vals,
                                   const float *__restrict__ t_rand, float perturb, int lindisp,
                                   float *__restrict__ xyz, float *__restrict__ z_out) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (int64_t)R * S) return;
  const int r = (int)(q / S), j = (int)(q - (int64_t)r * S);
  const float *ray = rays + (size_t)r * 8;
  const float near = ray[6], far = ray[7];
  float z = marcher_z(near, far, t_vals[j], lindisp);
  if (perturb > 0.0f) {
    const float lower = j == 0 ? z : __fmul_rn(0.5f, __fadd_rn(marcher_z(near, far, t_vals[j - 1], lindisp), z));
    const float upper = j == S - 1 ? z : __fmul_rn(0.5f, __fadd_rn(z, marcher_z(near, far, t_vals[j + 1], lindisp)));
    z = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), __fmul_rn(perturb, t_rand[q])));
  }
  z_out[q] = z;
#pragma unroll
  for (int c = 0; c < 3; ++c) xyz[3 * q + c] = __fadd_rn(ray[c], __fmul_rn(ray[3 + c], z));
}

}  // namespace

extern "C" int ucn_ray_marcher(const float *rays, int R, int S, const float *t_vals, const float *t_rand, float perturb,
                               int lindisp, float *xyz, float *z, void *stream) {
  UCN_CHECK_ARG(rays && t_vals && xyz && z, "ucn_ray_marcher: null argument");
  UCN_CHECK_ARG(S > 0 && R >= 0, "ucn_ray_marcher: bad sizes");
  UCN_CHECK_ARG(!(perturb > 0.0f) || t_rand, "ucn_ray_marcher: perturb > 0 needs t_rand");
  if (R == 0) return UCN_OK;
  const int64_t n = (int64_t)R * S;
  ray_marcher_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rays, R, S, t_vals, t_rand, perturb,
                                                                                    lindisp, xyz, z);
  UCN_LAUNCH_CHECK("ray_marcher_kernel");
  return UCN_OK;
}

extern "C" int ucn_raygen_sample(const ucn_frame *f, const float *pix_yx, int ray0, int R, const float *t_vals,
                                 const float *t_rand, uint64_t seed, float *z, float *rays_d, float *pts,
                                 float *ndc, void *stream) {
  UCN_CHECK_ARG(f && t_vals && z, "ucn_raygen_sample: null frame/t_vals/z");
  UCN_CHECK_ARG(f->S > 0 && f->S % 3 == 0 && f->S <= UCN_MAX_S, "ucn_raygen_sample: S must be a multiple of 3, <= 1024");
  UCN_CHECK_ARG(R >= 0 && ray0 >= 0, "ucn_raygen_sample: negative ray range");
  UCN_CHECK_ARG(pix_yx || ray0 + R <= f->H * f->W, "ucn_raygen_sample: ray range exceeds the frame");
  if (R == 0) return UCN_OK;
  int P = 32;
  while (P < f->S) P <<= 1;
  const size_t smem = (size_t)kWarpsPerBlock * 2 * P * sizeof(float);
  static int attr[kUcnMaxDevices];
  if (smem > 48 * 1024) {
    cudaError_t e = ucn_ensure_dyn_smem(raygen_sample_kernel, (int)smem, attr);
    if (e != cudaSuccess) return ucn_cuda_status(e, "raygen smem attribute");
  }
  const int blocks = (R + kWarpsPerBlock - 1) / kWarpsPerBlock;
  raygen_sample_kernel<<<blocks, kWarpsPerBlock * 32, smem, (cudaStream_t)stream>>>(
      *f, pix_yx, ray0, R, t_vals, t_rand, seed, P, z, rays_d, pts, ndc);
  UCN_LAUNCH_CHECK("raygen_sample_kernel");
  return UCN_OK;
}

extern "C" int ucn_dir_feature(const ucn_frame *f, const float *rays_d, int R, int ray_offset, float *angle,
                               void *stream) {
  UCN_CHECK_ARG(f && rays_d && angle, "ucn_dir_feature: null argument");
  if (R <= 0) return UCN_OK;
  dir_feature_kernel<<<(R + 255) / 256, 256, 0, (cudaStream_t)stream>>>(*f, rays_d, R, ray_offset, angle);
  UCN_LAUNCH_CHECK("dir_feature_kernel");
  return UCN_OK;
}
