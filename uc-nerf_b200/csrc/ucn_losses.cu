// ucn_losses.cu -- the renderer-side loss of the training step, value AND gradient in one launch
// ("losses are ~40 tiny kernels on 2 048 rays").
//
// replaces (reference): train.py:164-188 (img2mse, the weighted sparse-depth term, the weighted sum),
//   network/mvs_models.py:491-509 EdgePreservingSmoothnessLoss, utils/loss.py:54-112 compute_scale_and_shift +
//   gradient_loss + GradientLoss -- and the autograd graph torch builds for them (about 150 launches for 2 048 rays).
// One CTA: the work is a few thousand elements, so the cost is one launch instead of a launch-latency chain.  The
// gradients are the closed forms restated in oracle/losses.py (the least-squares shift cancels in the gradient
// differences; the scale carries d scale / d pred), evaluated in fp64 per patch.
#include "ucn_common.cuh"

namespace {

constexpr int kT = 256;
constexpr int kMaxPs = 8;

__device__ double block_sum(double v, double *sh) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double t = 0.0;
  for (int i = 0; i < kT / 32; ++i) t += sh[i];
  return t;
}
__device__ __forceinline__ double sgn(double v) { return v > 0.0 ? 1.0 : (v < 0.0 ? -1.0 : 0.0); }

__global__ void __launch_bounds__(kT)
render_losses_kernel(const float *__restrict__ rgb, const float *__restrict__ depth, const float *__restrict__ target,
                     const long long *__restrict__ prow, const long long *__restrict__ pcol, int R, int N_rays,
                     int patch_num, int ps, const float *__restrict__ sparse_d, const float *__restrict__ sparse_w,
                     const float *__restrict__ dpt, int W, float *__restrict__ terms, float *__restrict__ d_rgb,
                     float *__restrict__ d_depth) {
  __shared__ double sh[kT / 32];
  const int tid = threadIdx.x;
  const int half = patch_num / 2, pp = ps * ps, n_si = patch_num - half;
  // ---- 5 * img2mse (train.py:164,188) --------------------------------------------------------------------------------
  double s = 0.0;
  const double k_img = 5.0 * 2.0 / (3.0 * R);
  for (int i = tid; i < 3 * R; i += kT) {
    const double d = (double)rgb[i] - (double)target[i];
    s += d * d;
    d_rgb[i] = (float)(k_img * d);
  }
  const double img = block_sum(s, sh) / (3.0 * R);
  // ---- 0.05 * mean((depth - sparse)^2 * weight) over the sparse-depth rays (train.py:166-170,186) -------------------
  const int Ns = R - N_rays;
  s = 0.0;
  for (int j = tid; j < R; j += kT) {
    float g = 0.0f;
    if (j >= N_rays) {
      const long long t = prow[j] * W + pcol[j];
      const double d = (double)depth[j] - (double)sparse_d[t], w = (double)sparse_w[t];
      s += d * d * w;
      g = (float)(0.05 * 2.0 * d * w / Ns);
    }
    d_depth[j] = g;                         // patch rays are overwritten below; the other image rays have no depth term
  }
  const double nd = block_sum(s, sh) / Ns;
  __syncthreads();
  // ---- one thread per patch ------------------------------------------------------------------------------------------
  double sm = 0.0, si = 0.0;
  for (int p = tid; p < patch_num; p += kT) {
    double x[kMaxPs * kMaxPs], t[kMaxPs * kMaxPs], g[kMaxPs * kMaxPs];
    for (int k = 0; k < pp; ++k) {
      const int ray = p * pp + k;
      x[k] = (double)depth[ray];
      t[k] = (double)dpt[prow[ray] * W + pcol[ray]];
      g[k] = 0.0;
    }
    if (p < half) {
      // edge-preserving smoothness (mvs_models.py:491-509): four neighbour directions, bilateral weights from the guide
      const int dy[4] = {0, 1, 1, -1}, dx[4] = {1, 0, 1, 1};
      for (int dir = 0; dir < 4; ++dir) {
        const int ny = dy[dir] == 0 ? ps : ps - 1, nx = dx[dir] == 0 ? ps : ps - 1;      // pairs per patch in this direction
        const double inv = 1.0 / ((double)half * ny * nx) / 4.0;
        for (int y = 0; y < ps; ++y)
          for (int xx = 0; xx + dx[dir] < ps; ++xx) {
            const int y1 = y + dy[dir];
            if (y1 < 0 || y1 >= ps) continue;
            const int a = y * ps + xx, b = y1 * ps + xx + dx[dir];
            const double w = exp(-fabs(t[a] - t[b]) / 0.1), v = w * (x[a] - x[b]);
            sm += fabs(v) * inv;
            const double q = w * sgn(v) * inv;
            g[a] += q;
            g[b] -= q;
          }
      }
      for (int k = 0; k < pp; ++k) d_depth[p * pp + k] = (float)(0.05 * g[k]);
    } else {
      // scale-invariant gradient loss (utils/loss.py:54-112): least-squares scale per patch, L1 of the aligned
      // residual's horizontal + vertical differences, / number of patches
      double a00 = 0, a01 = 0, b0 = 0, b1 = 0;
      for (int k = 0; k < pp; ++k) { a00 += x[k] * x[k]; a01 += x[k]; b0 += x[k] * t[k]; b1 += t[k]; }
      const double n = pp, det = a00 * n - a01 * a01;
      const double sc = det != 0.0 ? (n * b0 - a01 * b1) / det : 0.0;
      double G = 0.0;
      for (int y = 0; y < ps; ++y)
        for (int xx = 0; xx < ps; ++xx) {
          const int a = y * ps + xx;
          if (xx + 1 < ps) {
            const int b = a + 1;
            const double r = sc * (x[b] - x[a]) - (t[b] - t[a]), sg = sgn(r);
            si += fabs(r) / n_si; G += sg * (x[b] - x[a]); g[b] += sg; g[a] -= sg;
          }
          if (y + 1 < ps) {
            const int b = a + ps;
            const double r = sc * (x[b] - x[a]) - (t[b] - t[a]), sg = sgn(r);
            si += fabs(r) / n_si; G += sg * (x[b] - x[a]); g[b] += sg; g[a] -= sg;
          }
        }
      for (int k = 0; k < pp; ++k) {
        const double ds = det != 0.0 ? ((n * t[k] - b1) - sc * (2.0 * n * x[k] - 2.0 * a01)) / det : 0.0;
        d_depth[p * pp + k] = (float)(0.008 * (G * ds + sc * g[k]) / n_si);
      }
    }
  }
  sm = block_sum(sm, sh);
  si = block_sum(si, sh);
  if (tid == 0) {
    terms[0] = (float)(0.05 * nd + 0.05 * sm + 0.008 * si + 5.0 * img);     // train.py:185-188
    terms[1] = (float)img; terms[2] = (float)nd; terms[3] = (float)sm; terms[4] = (float)si;
  }
}

}  // namespace

// terms5 = {loss, img, nerf_depth, smooth, scale_invariant}; d_rgb [R][3] and d_depth [R] = d loss / d (rgb, depth).
// Rays [0, patch_num*patch_size^2) are the patches (first half: smoothness, second half: scale-invariant term), rays
// [N_rays, R) the sparse-depth rays; pix_rows / pix_cols [R] int64 index the [H][W] maps.
extern "C" int ucn_render_losses(const float *rgb, const float *depth, const float *target_rgb, const int64_t *pix_rows,
                                 const int64_t *pix_cols, int R, int N_rays, int patch_num, int patch_size,
                                 const float *sparse_depths, const float *sparse_weights, const float *dpt, int H, int W,
                                 float *terms5, float *d_rgb, float *d_depth, void *stream) {
  UCN_CHECK_ARG(rgb && depth && target_rgb && pix_rows && pix_cols && sparse_depths && sparse_weights && dpt && terms5 &&
                d_rgb && d_depth, "ucn_render_losses: null argument");
  UCN_CHECK_ARG(patch_size >= 2 && patch_size <= kMaxPs, "ucn_render_losses: 2 <= patch_size <= 8");
  UCN_CHECK_ARG(patch_num >= 2 && patch_num * patch_size * patch_size <= N_rays && N_rays < R && H > 0 && W > 0,
                "ucn_render_losses: need patch_num >= 2, patches inside the first N_rays rays, and sparse-depth rays after them");
  render_losses_kernel<<<1, kT, 0, (cudaStream_t)stream>>>(rgb, depth, target_rgb, reinterpret_cast<const long long *>(pix_rows),
                                                           reinterpret_cast<const long long *>(pix_cols), R, N_rays, patch_num,
                                                           patch_size, sparse_depths, sparse_weights, dpt, W, terms5, d_rgb, d_depth);
  UCN_LAUNCH_CHECK("render_losses_kernel");
  return UCN_OK;
}
