// ucn_select.cu -- ray SELECTION of the training sampler in one launch.
//
// replaces (reference): utils/utils.py:450-479 (the two get_rays_with_random_patches calls + get_rays_mvs(isRandom) +
//   get_rays_mvs_coord of build_rays), :169-215 get_rays_with_random_patches(random=False: torch.multinomial on the
//   confidence map, patch offsets clamped to [0, H/ps - 2], a uniform shift in [0, ps) per patch, a Python loop over the
//   patches), :217-247 get_rays_mvs (uniform pixels, optional centre crop), :274-320 get_rays_mvs_coord (the sparse-depth
//   coordinates appended).  The PyTorch form of this (rays.select_training_pixels: ~45 small launches, 0.55 ms of the
//   6.8 ms training step) stays for seeded reproducibility with a torch.Generator.
//
// Weighted sampling without replacement of `half` pixels from w = confidence and `half` from w = 1 - confidence is the
// exponential race torch.multinomial itself runs (q ~ Exp(1), take the top-k of w / q): key_i = -log(u_i) / w_i with a
// Philox4x32-10 uniform per (seed, pixel, weight set), the `half` SMALLEST keys win.  Every block reduces its slice to
// its `half` best per set; the last block to finish (threadfence + counter) merges the candidates, places the patches,
// draws the uniform pixels and appends the coordinates.  Same distribution as the reference, not the same stream.
#include "ucn_common.cuh"

#include <mutex>

namespace {

constexpr int kT = 256;
constexpr int kPerBlock = 2048;        // pixels per block
constexpr int kMaxHalf = 64;           // patches per weight set

__device__ __forceinline__ uint4 philox(uint4 ctr, uint2 key) {
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
__device__ __forceinline__ float u01_open(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }   // (0, 1)

struct Cand { float key; int idx; };

// the `k` smallest keys of keys[0..n) (shared memory) -> out[0..k), by k rounds of a block-wide arg-min; ties by index
__device__ void block_select(float *keys, const int *ids, int n, int k, Cand *out, float *red_k, int *red_i) {
  for (int round = 0; round < k; ++round) {
    float best = __int_as_float(0x7f800000);
    int bi = 0x7fffffff, bslot = -1;
    for (int i = threadIdx.x; i < n; i += kT) {
      const float v = keys[i];
      const int id = ids ? ids[i] : i;
      if (v < best || (v == best && id < bi)) { best = v; bi = id; bslot = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o), os = __shfl_xor_sync(0xffffffffu, bslot, o);
      if (ov < best || (ov == best && oi < bi)) { best = ov; bi = oi; bslot = os; }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { red_k[threadIdx.x >> 5] = best; red_i[2 * (threadIdx.x >> 5)] = bi; red_i[2 * (threadIdx.x >> 5) + 1] = bslot; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float b = red_k[0];
      int id = red_i[0], sl = red_i[1];
      for (int w = 1; w < kT / 32; ++w)
        if (red_k[w] < b || (red_k[w] == b && red_i[2 * w] < id)) { b = red_k[w]; id = red_i[2 * w]; sl = red_i[2 * w + 1]; }
      out[round].key = b;
      out[round].idx = id;
      if (sl >= 0) keys[sl] = __int_as_float(0x7f800000);      // taken
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kT)
select_pixels_kernel(const float *__restrict__ conf, int H, int W, int n_rand, int half, int ps, int crop,
                     const long long *__restrict__ coords, int n_depth, uint64_t seed, Cand *__restrict__ cand,
                     unsigned int *__restrict__ counter, float *__restrict__ pix) {
  __shared__ float keys[2][kPerBlock];
  __shared__ float red_k[kT / 32];
  __shared__ int red_i[2 * (kT / 32)];
  __shared__ Cand win[2][kMaxHalf];
  __shared__ int is_last;
  const int n = H * W, base = blockIdx.x * kPerBlock;
  const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  for (int i = threadIdx.x; i < kPerBlock; i += kT) {
    float k0 = __int_as_float(0x7f800000), k1 = k0;
    const int g = base + i;
    if (g < n) {
      const float c = conf[g];
      const uint4 r = philox(make_uint4((uint32_t)g, 0u, 1u, 0u), key);
      if (c > 0.0f) k0 = -__logf(u01_open(r.x)) / c;                       // w = confidence
      if (1.0f - c > 0.0f) k1 = -__logf(u01_open(r.y)) / (1.0f - c);       // w = 1 - confidence (utils/utils.py:460-468)
    }
    keys[0][i] = k0; keys[1][i] = k1;
  }
  __syncthreads();
  const int R = 2 * half * ps * ps + n_rand + n_depth;
  for (int set = 0; set < 2; ++set) {
    block_select(keys[set], nullptr, kPerBlock, half, win[set], red_k, red_i);
    for (int j = threadIdx.x; j < half; j += kT) {
      Cand c = win[set][j];
      c.idx += base;
      cand[((size_t)blockIdx.x * 2 + set) * half + j] = c;
    }
  }
  // ---- last block: merge, place patches, uniform pixels, coordinates ------------------------------------------------
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int nc = gridDim.x * half;                      // candidates per set (launcher: <= kPerBlock)
  int *ids = reinterpret_cast<int *>(keys[1]);
  for (int set = 0; set < 2; ++set) {
    for (int i = threadIdx.x; i < nc; i += kT) {
      const Cand c = cand[((size_t)(i / half) * 2 + set) * half + i % half];
      keys[0][i] = c.key; ids[i] = c.idx;
    }
    __syncthreads();
    block_select(keys[0], ids, nc, half, win[set], red_k, red_i);
    __syncthreads();
  }
  const int pp = ps * ps;
  for (int e = threadIdx.x; e < 2 * half * pp; e += kT) {        // rows = pi*ps + si + a, cols = pj*ps + sj + b  (:187-205)
    const int p = e / pp, a = (e % pp) / ps, b = e % ps;
    const int set = p / half, j = p % half;
    const int sel = win[set][j].idx;
    int pi = (sel / W) / ps, pj = (sel % W) / ps;
    pi = min(max(pi, 0), H / ps - 2);
    pj = min(max(pj, 0), W / ps - 2);
    const uint4 r = philox(make_uint4((uint32_t)p, 0u, 2u, 0u), key);
    const int si = (int)(r.x % (uint32_t)ps), sj = (int)(r.y % (uint32_t)ps);
    pix[e] = (float)(pi * ps + si + a);
    pix[R + e] = (float)(pj * ps + sj + b);
  }
  const int o1 = 2 * half * pp;
  const int x0 = crop ? W / 6 : 0, xw = crop ? W - 2 * (W / 6) : W, y0 = crop ? H / 6 : 0, yh = crop ? H - 2 * (H / 6) : H;
  for (int e = threadIdx.x; e < n_rand; e += kT) {               // uniform pixels (:241-250)
    const uint4 r = philox(make_uint4((uint32_t)e, 0u, 3u, 0u), key);
    pix[o1 + e] = (float)(y0 + (int)(r.y % (uint32_t)yh));
    pix[R + o1 + e] = (float)(x0 + (int)(r.x % (uint32_t)xw));
  }
  const int o2 = o1 + n_rand;
  for (int e = threadIdx.x; e < n_depth; e += kT) {              // sparse-depth coordinates [N][2] = (row, col)  (:304-306)
    pix[o2 + e] = (float)coords[2 * e];
    pix[R + o2 + e] = (float)coords[2 * e + 1];
  }
  if (threadIdx.x == 0) *counter = 0;                            // ready for the next launch on this stream
}

}  // namespace

extern "C" size_t ucn_select_workspace_bytes(int H, int W, int patch_num) {
  const int half = patch_num / 2, blocks = (H * W + kPerBlock - 1) / kPerBlock;
  return 256 + (size_t)blocks * 2 * (half > 0 ? half : 1) * sizeof(Cand);
}

// pix_yx [2][R] fp32 (row, col), R = 2*(patch_num/2)*ps^2 + n_rand + n_depth: patch pixels drawn by confidence, then by
// 1 - confidence, then n_rand uniform pixels (centre crop when crop != 0), then the n_depth coordinates.
// workspace: ucn_select_workspace_bytes() bytes, ZERO-FILLED once by the caller before the first use.
extern "C" int ucn_select_training_pixels(const float *confidence, int H, int W, int n_rand, int patch_num, int patch_size,
                                          int crop, const long long *coords_rc, int n_depth, uint64_t seed, void *workspace,
                                          float *pix_yx, void *stream) {
  UCN_CHECK_ARG(confidence && workspace && pix_yx, "ucn_select_training_pixels: null argument");
  const int half = patch_num / 2;
  UCN_CHECK_ARG(H > 0 && W > 0 && patch_size > 0 && n_rand >= 0 && n_depth >= 0 && (n_depth == 0 || coords_rc),
                "ucn_select_training_pixels: bad argument");
  UCN_CHECK_ARG(half >= 0 && half <= kMaxHalf, "ucn_select_training_pixels: at most 128 patches");
  UCN_CHECK_ARG(H / patch_size >= 2 && W / patch_size >= 2, "ucn_select_training_pixels: image smaller than two patches");
  const int blocks = (H * W + kPerBlock - 1) / kPerBlock;
  UCN_CHECK_ARG((long long)blocks * half <= kPerBlock, "ucn_select_training_pixels: image too large for one merge block");
  unsigned int *counter = static_cast<unsigned int *>(workspace);
  Cand *cand = reinterpret_cast<Cand *>(static_cast<unsigned char *>(workspace) + 256);
  select_pixels_kernel<<<blocks, kT, 0, (cudaStream_t)stream>>>(confidence, H, W, n_rand, half, patch_size, crop, coords_rc, n_depth,
                                                                seed, cand, counter, pix_yx);
  UCN_LAUNCH_CHECK("select_pixels_kernel");
  return UCN_OK;
}

// ---- the same selection for callers that may not allocate (no counterpart in the reference) ---------------------------
// Added: code under CUDA graph capture, tight serving loops and harnesses that
// forbid cudaMalloc cannot provide (or let us grow) a workspace.  The workspace of select_pixels_kernel lives in a
// file-scope device array instead: 256 B (the block counter) + kPerBlock candidates per weight set = 33 024 B, which is
// the largest workspace ucn_select_training_pixels itself accepts (blocks * half <= kPerBlock), held by every GPU that
// loads the library until exit.  In host code `g_select_scratch` is only a stand-in for the symbol; the device address is
// taken once per device with cudaGetSymbolAddress (that device current) and passed to the kernel.
// Every call shares the array: the counter word is cleared on the caller's stream before each launch (the array keeps the
// previous call's candidates, which the kernel overwrites before reading), and two calls that OVERLAP on one GPU
// (different streams, or different host threads' streams) race on it -- serialise them or use the workspace form.
namespace {
constexpr size_t kSelectScratchBytes = 256 + (size_t)kPerBlock * 2 * sizeof(Cand);
static __device__ __align__(256) unsigned char g_select_scratch[kSelectScratchBytes];

cudaError_t select_scratch_address(unsigned char **out) {
  static unsigned char *addr[kUcnMaxDevices];        // per device: the symbol has one instance per GPU
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kUcnMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(mu);
  if (!addr[dev]) {
    void *p = nullptr;
    e = cudaGetSymbolAddress(&p, g_select_scratch);
    if (e != cudaSuccess) return e;
    addr[dev] = static_cast<unsigned char *>(p);
  }
  *out = addr[dev];
  return cudaSuccess;
}
}  // namespace

extern "C" size_t ucn_select_static_scratch_bytes(void) { return kSelectScratchBytes; }

extern "C" int ucn_select_training_pixels_static(const float *confidence, int H, int W, int n_rand, int patch_num,
                                                 int patch_size, int crop, const long long *coords_rc, int n_depth,
                                                 uint64_t seed, float *pix_yx, void *stream) {
  UCN_CHECK_ARG(confidence && pix_yx, "ucn_select_training_pixels_static: null argument");
  const int half = patch_num / 2;
  UCN_CHECK_ARG(H > 0 && W > 0 && patch_size > 0 && n_rand >= 0 && n_depth >= 0 && (n_depth == 0 || coords_rc),
                "ucn_select_training_pixels_static: bad argument");
  UCN_CHECK_ARG(half >= 0 && half <= kMaxHalf, "ucn_select_training_pixels_static: at most 128 patches");
  UCN_CHECK_ARG(H / patch_size >= 2 && W / patch_size >= 2, "ucn_select_training_pixels_static: image smaller than two patches");
  const long long blocks = ((long long)H * W + kPerBlock - 1) / kPerBlock;
  if (256 + (size_t)blocks * 2 * (half > 0 ? half : 1) * sizeof(Cand) > kSelectScratchBytes || blocks * half > kPerBlock)
    return ucn_set_error(UCN_E_ARG, "ucn_select_training_pixels_static: ceil(H*W/2048) * (patch_num/2) exceeds 2048 -- the "
                                    "33 024-byte static scratch holds no more candidates; use ucn_select_training_pixels");
  unsigned char *scratch = nullptr;
  cudaError_t e = select_scratch_address(&scratch);
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_select_training_pixels_static: cudaGetSymbolAddress");
  e = cudaMemsetAsync(scratch, 0, 256, (cudaStream_t)stream);          // the block counter must start at zero
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_select_training_pixels_static: clear counter");
  select_pixels_kernel<<<(unsigned)blocks, kT, 0, (cudaStream_t)stream>>>(
      confidence, H, W, n_rand, half, patch_size, crop, coords_rc, n_depth, seed, reinterpret_cast<Cand *>(scratch + 256),
      reinterpret_cast<unsigned int *>(scratch), pix_yx);
  UCN_LAUNCH_CHECK("select_pixels_kernel (static scratch)");
  return UCN_OK;
}
