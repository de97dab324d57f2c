// ucn_metrics.cu -- device-resident evaluation metrics of a rendered frame (the caller of the
// render path).  The reference moves every frame to the host and evaluates it in numpy:
//   utils/utils.py:18-22       get_psnr          -10 log10(mean((img - tar)^2)) per image
//   utils/evaluation.py:8-27   compute_errors    abs_rel, sq_rel, rmse, rmse_log, a1, a2, a3 over the valid pixels
//   utils/evaluation.py:30-74  depth_evaluation  valid = min < gt < max (and mask); the prediction is scaled by
//                                                median(valid gt) / median(valid pred) over ALL frames, clamped, then
//                                                compute_errors per frame
// Here the frame stays in HBM: one streaming reduction per quantity (fp64 accumulators, so the sums do not depend on
// the reduction order beyond 1e-15) and an exact median by radix select on the order-preserving integer image of the
// float (four 8-bit passes, both middle ranks of both arrays at once) -- no sort, no host copy of the frame.  The
// per-pixel arithmetic (ratio scaling, clamps, gt/pred quotients against 1.25^k) is the reference's fp32 arithmetic
// operation for operation, so the threshold counts a1..a3 are bit-exact.  All kernels are HBM-bound streaming reads:
// 8 B/pixel (+1 B mask) per pass.
#include "ucn_common.cuh"

namespace {

constexpr int kThreads = 256;

struct SelectState {            // lives in the caller's workspace
  unsigned int prefix[4];       // selections: 0 gt lower-middle, 1 gt upper-middle, 2 pred lower-middle, 3 pred upper-middle
  unsigned long long rank[4];   // rank still to be skipped inside the current prefix
  unsigned long long count;     // valid pixels
  unsigned int hist[4][256];
};

__device__ __forceinline__ unsigned int fkey(float f) {        // monotone float -> uint
  const unsigned int u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float fkey_inv(unsigned int k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}
__device__ __forceinline__ bool valid_px(float gt, const unsigned char *mask, long long i, float lo, float hi) {
  return gt > lo && gt < hi && (!mask || mask[i] != 0);          // evaluation.py:40,46-47
}

__global__ void __launch_bounds__(kThreads)
select_hist_kernel(const float *__restrict__ gt, const float *__restrict__ pred, const unsigned char *__restrict__ mask,
                   long long n, float lo, float hi, SelectState *st, int pass) {
  __shared__ unsigned int h[4][256];
  for (int i = threadIdx.x; i < 1024; i += kThreads) (&h[0][0])[i] = 0u;
  __syncthreads();
  const int shift = 24 - 8 * pass;
  const unsigned int himask = pass == 0 ? 0u : ~((1u << (shift + 8)) - 1u);
  const unsigned int p0 = st->prefix[0], p1 = st->prefix[1], p2 = st->prefix[2], p3 = st->prefix[3];
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += (long long)gridDim.x * kThreads) {
    const float g = gt[i];
    if (!valid_px(g, mask, i, lo, hi)) continue;
    const unsigned int kg = fkey(g), kp = fkey(pred[i]);
    const unsigned int dg = (kg >> shift) & 255u, dp = (kp >> shift) & 255u;
    if ((kg & himask) == p0) atomicAdd(&h[0][dg], 1u);
    if ((kg & himask) == p1) atomicAdd(&h[1][dg], 1u);
    if ((kp & himask) == p2) atomicAdd(&h[2][dp], 1u);
    if ((kp & himask) == p3) atomicAdd(&h[3][dp], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 1024; i += kThreads) {
    const unsigned int v = (&h[0][0])[i];
    if (v) atomicAdd(&st->hist[0][0] + i, v);
  }
}

// one thread per selection walks its 256-bin histogram
__global__ void select_pick_kernel(SelectState *st, int pass, float *out4, long long *out_count) {
  const int r = threadIdx.x;
  if (r < 4) {
    if (pass == 0) {
      unsigned long long c = 0;
      for (int d = 0; d < 256; ++d) c += st->hist[r][d];
      if (r == 0) { st->count = c; if (out_count) *out_count = (long long)c; }
      st->rank[r] = c == 0 ? 0 : ((r & 1) ? c / 2 : (c - 1) / 2);      // numpy median: mean of ranks (n-1)/2 and n/2
    }
    const int shift = 24 - 8 * pass;
    unsigned long long k = st->rank[r], below = 0;
    int d = 0;
    for (; d < 255; ++d) {
      const unsigned long long c = st->hist[r][d];
      if (k < below + c) break;
      below += c;
    }
    st->prefix[r] |= (unsigned int)d << shift;
    st->rank[r] = k - below;
    if (pass == 3) out4[r] = fkey_inv(st->prefix[r]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) (&st->hist[0][0])[i] = 0u;
}

__device__ __forceinline__ void block_sum_to(double *vals, int nvals, double *out) {
  __shared__ double part[kThreads / 32][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int k = 0; k < nvals; ++k) {
    double v = vals[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) part[w][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < nvals) {
    double v = 0.0;
    for (int i = 0; i < kThreads / 32; ++i) v += part[i][threadIdx.x];
    atomicAdd(out + threadIdx.x, v);
  }
}

__global__ void __launch_bounds__(kThreads)
sq_err_kernel(const float *__restrict__ a, const float *__restrict__ b, long long n, double *out) {
  double s[1] = {0.0};
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += (long long)gridDim.x * kThreads) {
    const float d = a[i] - b[i];                                   // fp32 difference and square, as numpy does
    s[0] += (double)(d * d);
  }
  block_sum_to(s, 1, out);
}

// out8: sums of |gt-p|/gt, (gt-p)^2/gt, (gt-p)^2, (log gt - log p)^2, [t<1.25], [t<1.25^2], [t<1.25^3], and the count
__global__ void __launch_bounds__(kThreads)
depth_err_kernel(const float *__restrict__ gt, const float *__restrict__ pred, const unsigned char *__restrict__ mask,
                 long long n, float lo, float hi, float ratio, double *out8) {
  double s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += (long long)gridDim.x * kThreads) {
    const float g = gt[i];
    if (!valid_px(g, mask, i, lo, hi)) continue;
    float p = __fmul_rn(pred[i], ratio);                           // evaluation.py:63-65
    if (p < lo) p = lo;
    if (p > hi) p = hi;
    const float t = fmaxf(__fdiv_rn(g, p), __fdiv_rn(p, g));       // :11
    const float d = g - p, d2 = d * d;
    const float dl = logf(g) - logf(p);
    s[0] += (double)__fdiv_rn(fabsf(d), g);
    s[1] += (double)__fdiv_rn(d2, g);
    s[2] += (double)d2;
    s[3] += (double)(dl * dl);
    s[4] += t < 1.25f ? 1.0 : 0.0;
    s[5] += t < 1.5625f ? 1.0 : 0.0;
    s[6] += t < 1.953125f ? 1.0 : 0.0;
    s[7] += 1.0;
  }
  block_sum_to(s, 8, out8);
}

// SSIM map of one channel plane, summed over the interior (utils/evaluation.py:93 calls
// skimage.metrics.structural_similarity(gt, pred, data_range=1, channel_axis=2) with its defaults: 7 x 7 uniform window,
// sample covariance (x 49/48), K1 = 0.01, K2 = 0.03, the mean taken over the map cropped by 3 pixels per side, then over
// the channels).  Block = 32 x 8 output pixels of one plane; the (32+6) x (8+6) input patch of both images goes through
// shared memory, the 7 x 7 sums are taken in fp64.
constexpr int SS_W = 32, SS_H = 8, SS_R = 3;
__global__ void __launch_bounds__(SS_W * SS_H)
ssim_kernel(const float *__restrict__ a, const float *__restrict__ b, int H, int W, double c1, double c2,
            double *__restrict__ out_sum) {
  __shared__ float sa[SS_H + 2 * SS_R][SS_W + 2 * SS_R], sb[SS_H + 2 * SS_R][SS_W + 2 * SS_R];
  __shared__ double red[SS_W * SS_H / 32];
  const float *pa = a + (size_t)blockIdx.z * H * W, *pb = b + (size_t)blockIdx.z * H * W;
  const int x0 = SS_R + blockIdx.x * SS_W, y0 = SS_R + blockIdx.y * SS_H;      // first interior pixel of the tile
  for (int i = threadIdx.x; i < (SS_H + 2 * SS_R) * (SS_W + 2 * SS_R); i += SS_W * SS_H) {
    const int ly = i / (SS_W + 2 * SS_R), lx = i % (SS_W + 2 * SS_R);
    const int y = y0 - SS_R + ly, x = x0 - SS_R + lx;
    const bool ok = y < H && x < W;
    sa[ly][lx] = ok ? pa[(size_t)y * W + x] : 0.f;
    sb[ly][lx] = ok ? pb[(size_t)y * W + x] : 0.f;
  }
  __syncthreads();
  const int tx = threadIdx.x % SS_W, ty = threadIdx.x / SS_W;
  const int x = x0 + tx, y = y0 + ty;
  double s = 0.0;
  if (x < W - SS_R && y < H - SS_R) {
    double ua = 0, ub = 0, uaa = 0, ubb = 0, uab = 0;
#pragma unroll
    for (int dy = 0; dy < 2 * SS_R + 1; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2 * SS_R + 1; ++dx) {
        const double p = sa[ty + dy][tx + dx], q = sb[ty + dy][tx + dx];
        ua += p; ub += q; uaa += p * p; ubb += q * q; uab += p * q;
      }
    const double np_ = 49.0, cov = np_ / (np_ - 1.0);
    ua /= np_; ub /= np_; uaa /= np_; ubb /= np_; uab /= np_;
    const double va = cov * (uaa - ua * ua), vb = cov * (ubb - ub * ub), vab = cov * (uab - ua * ub);
    s = ((2.0 * ua * ub + c1) * (2.0 * vab + c2)) / ((ua * ua + ub * ub + c1) * (va + vb + c2));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < SS_W * SS_H / 32; ++i) t += red[i];
    atomicAdd(out_sum, t);
  }
}

inline unsigned grid_for(long long n) {
  const long long b = (n + kThreads - 1) / kThreads;
  const long long cap = 8ll * ucn_sm_count();                      // persistent grid-stride: 8 CTAs per SM
  return (unsigned)(b < cap ? (b < 1 ? 1 : b) : cap);
}

}  // namespace

extern "C" size_t ucn_metrics_workspace_bytes(void) { return sizeof(SelectState); }

// *out_sum += sum_i (a[i] - b[i])^2 (fp64 accumulation; the caller zeroes out_sum).  PSNR = -10 log10(sum / n).
extern "C" int ucn_sq_err_sum(const float *a, const float *b, int64_t n, double *out_sum, void *stream) {
  UCN_CHECK_ARG(a && b && out_sum, "ucn_sq_err_sum: null argument");
  if (n <= 0) return UCN_OK;
  sq_err_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(a, b, n, out_sum);
  UCN_LAUNCH_CHECK("sq_err_kernel");
  return UCN_OK;
}

// *out_sum += sum over the C channel planes and the interior pixels (3 .. H-4, 3 .. W-4) of the SSIM map of a vs b
// ([C][H][W] fp32, data_range = 1 scaled by `data_range`); mean SSIM = sum / (C (H-6) (W-6)).  The caller zeroes out_sum.
extern "C" int ucn_ssim_sum(const float *a, const float *b, int C, int H, int W, float data_range, double *out_sum,
                            void *stream) {
  UCN_CHECK_ARG(a && b && out_sum, "ucn_ssim_sum: null argument");
  UCN_CHECK_ARG(C >= 1 && C <= 65535 && H >= 7 && W >= 7, "ucn_ssim_sum: the 7 x 7 window needs images of at least 7 x 7 pixels");
  const double c1 = (0.01 * data_range) * (0.01 * data_range), c2 = (0.03 * data_range) * (0.03 * data_range);
  dim3 grid((unsigned)((W - 2 * SS_R + SS_W - 1) / SS_W), (unsigned)((H - 2 * SS_R + SS_H - 1) / SS_H), (unsigned)C);
  UCN_CHECK_ARG(grid.y < 65536u, "ucn_ssim_sum: image too tall");
  ssim_kernel<<<grid, SS_W * SS_H, 0, (cudaStream_t)stream>>>(a, b, H, W, c1, c2, out_sum);
  UCN_LAUNCH_CHECK("ssim_kernel");
  return UCN_OK;
}

// Exact medians of the valid ground-truth and predicted depths (valid: min_depth < gt < max_depth and mask != 0; mask
// may be NULL).  out_mid4 = {gt lower-middle, gt upper-middle, pred lower-middle, pred upper-middle} (ranks (c-1)/2 and
// c/2 of the c valid pixels; numpy's median is the mean of the two), *out_count = c.
extern "C" int ucn_depth_median(const float *gt, const float *pred, const unsigned char *mask, int64_t n, float min_depth,
                                float max_depth, void *workspace, size_t workspace_bytes, float *out_mid4,
                                int64_t *out_count, void *stream) {
  UCN_CHECK_ARG(gt && pred && workspace && out_mid4 && out_count, "ucn_depth_median: null argument");
  UCN_CHECK_ARG(workspace_bytes >= sizeof(SelectState), "ucn_depth_median: workspace too small");
  UCN_CHECK_ARG(n > 0, "ucn_depth_median: n <= 0");
  cudaStream_t st = (cudaStream_t)stream;
  SelectState *state = reinterpret_cast<SelectState *>(workspace);
  cudaError_t e = cudaMemsetAsync(state, 0, sizeof(SelectState), st);
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_depth_median: memset");
  for (int pass = 0; pass < 4; ++pass) {
    select_hist_kernel<<<grid_for(n), kThreads, 0, st>>>(gt, pred, mask, n, min_depth, max_depth, state, pass);
    UCN_LAUNCH_CHECK("select_hist_kernel");
    select_pick_kernel<<<1, 256, 0, st>>>(state, pass, out_mid4, reinterpret_cast<long long *>(out_count));
    UCN_LAUNCH_CHECK("select_pick_kernel");
  }
  return UCN_OK;
}

// out8 += {sum |gt-p|/gt, sum (gt-p)^2/gt, sum (gt-p)^2, sum (log gt - log p)^2, #[t<1.25], #[t<1.25^2], #[t<1.25^3],
// #valid} with p = clamp(pred * ratio, min_depth, max_depth), t = max(gt/p, p/gt) (the caller zeroes out8)
extern "C" int ucn_depth_errors(const float *gt, const float *pred, const unsigned char *mask, int64_t n, float min_depth,
                                float max_depth, float ratio, double *out8, void *stream) {
  UCN_CHECK_ARG(gt && pred && out8, "ucn_depth_errors: null argument");
  if (n <= 0) return UCN_OK;
  depth_err_kernel<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(gt, pred, mask, n, min_depth, max_depth, ratio, out8);
  UCN_LAUNCH_CHECK("depth_err_kernel");
  return UCN_OK;
}
