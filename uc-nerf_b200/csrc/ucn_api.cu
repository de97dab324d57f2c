// ucn_api.cu -- library plumbing (errors, launch accounting, device check), MLP mode dispatch and the
// whole-frame chain K1 -> K2 -> K3 -> K4 that replaces the reference's per-chunk Python loop
// (train.py:251-275).
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <stdlib.h>

#include "ucn_common.cuh"

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

int ucn_set_error(int code, const char *msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
int ucn_cuda_status(cudaError_t e, const char *where) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s)", where, cudaGetErrorString(e), cudaGetErrorName(e));
  return (int)e;
}
void ucn_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int ucn_sm_count() {
  static std::atomic<int> cache[kUcnMaxDevices];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kUcnMaxDevices) return 148;
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

int ucn_mlp_forward_fp32(const ucn_weights *w, const float *ndc_pts, const float *feat, const float *angle, int R,
                         int S, float *raw, void *stream);
int ucn_mlp_forward_tc(const ucn_weights *w, const float *ndc_pts, const float *feat, const float *angle, int R,
                       int S, float *raw, void *stream);
int ucn_mlp_forward_x3(const ucn_weights *w, const float *ndc_pts, const float *feat, long long f_tile_stride,
                       long long f_k_stride, long long f_row_stride, const float *angle, int R, int S, float *raw,
                       void *stream);
int ucn_mlp_forward_x3p(const ucn_weights *w, const float *ndc_pts, const void *operand_tiles, const float *angle, int R,
                        int S, float *raw, void *stream);
// Which split-bf16 kernel the fused frame path (ucn_render_rays) uses for UCN_MLP_BF16X3_TC:
//   2 = ucn_mlp_x3p.cu (default): TMEM-resident activations, pipelined; reads K2's operand tiles (ucn_gather_operands_from_z)
//   1 = ucn_mlp_x3.cu (round 1): activations in shared memory; reads tile-transposed or row-major fp32 feature rows
// UCN_X3_KERNEL=1 in the environment or ucn_mlp_x3_variant(1) selects the older pair for A/B timing; both meet the gates.
static int g_x3_kernel = 0;
static int x3_kernel() {
  if (g_x3_kernel == 0) {
    const char *e = getenv("UCN_X3_KERNEL");
    g_x3_kernel = (e && e[0] == '1') ? 1 : 2;
  }
  return g_x3_kernel;
}
extern "C" int ucn_mlp_x3_variant(int set) {
  if (set == 1 || set == 2) g_x3_kernel = set;
  return x3_kernel();
}

extern "C" int ucn_version(void) { return 100; }
extern "C" const char *ucn_last_error(void) { return g_err; }
extern "C" int64_t ucn_launch_count(int reset) {
  return reset ? g_launches.exchange(0) : g_launches.load();
}

extern "C" int ucn_check_device(void) {
  int dev = 0, major = 0, minor = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return ucn_cuda_status(e, "cudaGetDevice");
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  if (major != 10) {
    snprintf(g_err, sizeof(g_err), "libucnerf is built for sm_100a only; device %d is sm_%d%d", dev, major, minor);
    return UCN_E_UNSUPPORTED;
  }
  return UCN_OK;
}

// ---- peer-mapped buffers (one process per GPU; CUDA IPC over NVLink) ---------------------------------------------
// The owner allocates with cudaMalloc (its own allocation, so the IPC handle maps exactly this buffer) and exports
// a handle; every other process opens it WITH ITS OWN DEVICE CURRENT, which is what makes the mapping addressable
// by that device's kernels (cudaIpcMemLazyEnablePeerAccess) -- K4 then stores its tile straight into the owner's HBM.
static_assert(sizeof(cudaIpcMemHandle_t) == UCN_PEER_HANDLE_BYTES, "IPC handle size");
extern "C" int ucn_peer_alloc(size_t bytes, void **dptr, unsigned char *handle) {
  UCN_CHECK_ARG(dptr && handle && bytes > 0, "ucn_peer_alloc: null argument / zero size");
  cudaError_t e = cudaMalloc(dptr, bytes);
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_peer_alloc: cudaMalloc");
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, *dptr);
  if (e != cudaSuccess) { cudaFree(*dptr); *dptr = nullptr; return ucn_cuda_status(e, "ucn_peer_alloc: cudaIpcGetMemHandle"); }
  memcpy(handle, &h, sizeof(h));
  return UCN_OK;
}
extern "C" int ucn_peer_open(const unsigned char *handle, void **dptr) {
  UCN_CHECK_ARG(dptr && handle, "ucn_peer_open: null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return ucn_cuda_status(e, "ucn_peer_open: cudaIpcOpenMemHandle");
  return UCN_OK;
}
extern "C" int ucn_peer_close(void *dptr) {
  if (!dptr) return UCN_OK;
  cudaError_t e = cudaIpcCloseMemHandle(dptr);
  return e == cudaSuccess ? UCN_OK : ucn_cuda_status(e, "ucn_peer_close");
}
extern "C" int ucn_peer_free(void *dptr) {
  if (!dptr) return UCN_OK;
  cudaError_t e = cudaFree(dptr);
  return e == cudaSuccess ? UCN_OK : ucn_cuda_status(e, "ucn_peer_free");
}

extern "C" int ucn_mlp_forward(const ucn_weights *w, const float *ndc_pts, const float *feat, const float *angle,
                               int R, int S, float *raw, int mode, void *stream) {
  UCN_CHECK_ARG(w && ndc_pts && feat && angle && raw, "ucn_mlp_forward: null argument");
  UCN_CHECK_ARG(w->V >= 1 && w->V <= UCN_MAX_V, "ucn_mlp_forward: 1 <= V <= 8");
  UCN_CHECK_ARG(S > 0, "ucn_mlp_forward: S <= 0");
  if (R <= 0) return UCN_OK;
  if (mode == UCN_MLP_FP32) return ucn_mlp_forward_fp32(w, ndc_pts, feat, angle, R, S, raw, stream);
  if (mode == UCN_MLP_BF16_TC) return ucn_mlp_forward_tc(w, ndc_pts, feat, angle, R, S, raw, stream);
  if (mode == UCN_MLP_BF16X3_TC) {
    const long long F = 24 + 12 * w->V + 1;                       // row-major [N][F] rows (the reference layout)
    return ucn_mlp_forward_x3(w, ndc_pts, feat, 128 * F, 1, F, angle, R, S, raw, stream);
  }
  return ucn_set_error(UCN_E_ARG, "ucn_mlp_forward: unknown mode");
}

extern "C" int ucn_mlp_forward_tiles(const ucn_weights *w, const float *ndc_pts, const float *feat_tiles,
                                     const float *angle, int R, int S, float *raw, int mode, void *stream) {
  UCN_CHECK_ARG(w && ndc_pts && feat_tiles && angle && raw, "ucn_mlp_forward_tiles: null argument");
  UCN_CHECK_ARG(w->V >= 1 && w->V <= UCN_MAX_V, "ucn_mlp_forward_tiles: 1 <= V <= 8");
  UCN_CHECK_ARG(S > 0, "ucn_mlp_forward_tiles: S <= 0");
  if (mode != UCN_MLP_BF16X3_TC) return ucn_set_error(UCN_E_UNSUPPORTED, "ucn_mlp_forward_tiles: only UCN_MLP_BF16X3_TC reads tiled rows");
  if (R <= 0) return UCN_OK;
  const long long F = 24 + 12 * w->V + 1;
  return ucn_mlp_forward_x3(w, ndc_pts, feat_tiles, 128 * F, 128, 1, angle, R, S, raw, stream);
}

extern "C" int ucn_mlp_forward_operands(const ucn_weights *w, const float *ndc_pts, const void *operand_tiles,
                                        const float *angle, int R, int S, float *raw, int mode, void *stream) {
  UCN_CHECK_ARG(w && ndc_pts && operand_tiles && angle && raw, "ucn_mlp_forward_operands: null argument");
  UCN_CHECK_ARG(w->V >= 1 && w->V <= 7, "ucn_mlp_forward_operands: 1 <= V <= 7");
  UCN_CHECK_ARG(S > 0, "ucn_mlp_forward_operands: S <= 0");
  if (mode != UCN_MLP_BF16X3_TC) return ucn_set_error(UCN_E_UNSUPPORTED, "ucn_mlp_forward_operands: only UCN_MLP_BF16X3_TC reads operand tiles");
  if (R <= 0) return UCN_OK;
  return ucn_mlp_forward_x3p(w, ndc_pts, operand_tiles, angle, R, S, raw, stream);
}

// workspace carve-up for ucn_render_rays (all 256-byte aligned)
namespace {
struct Workspace {
  size_t z, rays_d, angle, ndc, feat, raw, total;
  Workspace(const ucn_frame *f, int R) {
    const size_t N = (size_t)R * f->S, F = 24 + 12 * (size_t)f->V + 1;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o += (bytes + 255) & ~size_t(255); return r; };
    z = take(N * 4); rays_d = take((size_t)R * 12); angle = take((size_t)R * 12);
    const size_t tiles = (N + 127) / 128, rows_b = 128 * F * 4, oper_b = ucn_operand_tile_bytes(f->V);
    ndc = take(N * 12); feat = take(tiles * (rows_b > oper_b ? rows_b : oper_b)); raw = take(N * 16);   // whole tiles, either layout
    total = o;
  }
};
}  // namespace

extern "C" size_t ucn_frame_workspace_bytes(const ucn_frame *f, int R, int mode) {
  (void)mode;
  if (!f || R <= 0) return 0;
  return Workspace(f, R).total;
}

extern "C" int ucn_render_rays(const ucn_frame *f, const ucn_weights *w, int ray0, int R, const float *t_vals,
                               const float *t_rand, uint64_t seed, int mode, void *workspace, size_t workspace_bytes,
                               float *out_rgb, float *out_depth, void *stream) {
  UCN_CHECK_ARG(f && w && t_vals && workspace && out_rgb && out_depth, "ucn_render_rays: null argument");
  UCN_CHECK_ARG(w->V == f->V, "ucn_render_rays: weights were packed for a different number of source views");
  if (R <= 0) return UCN_OK;
  const Workspace ws(f, R);
  if (workspace_bytes < ws.total) return ucn_set_error(UCN_E_WORKSPACE, "ucn_render_rays: workspace too small");
  char *base = reinterpret_cast<char *>(workspace);
  float *z = reinterpret_cast<float *>(base + ws.z), *rays_d = reinterpret_cast<float *>(base + ws.rays_d);
  float *angle = reinterpret_cast<float *>(base + ws.angle), *ndc = reinterpret_cast<float *>(base + ws.ndc);
  float *feat = reinterpret_cast<float *>(base + ws.feat), *raw = reinterpret_cast<float *>(base + ws.raw);
  int rc;
  if ((rc = ucn_raygen_sample(f, nullptr, ray0, R, t_vals, t_rand, seed, z, rays_d, nullptr, nullptr, stream))) return rc;
  if ((rc = ucn_dir_feature(f, rays_d, R, ray0, angle, stream))) return rc;
  if (mode == UCN_MLP_BF16X3_TC && x3_kernel() == 2 && f->V <= 7) {   // tensor-core path: K2 writes the MLP's operand tiles
    if ((rc = ucn_gather_operands_from_z(f, z, nullptr, ray0, R, feat, ndc, stream))) return rc;
    if ((rc = ucn_mlp_forward_operands(w, ndc, feat, angle, R, f->S, raw, mode, stream))) return rc;
  } else if (mode == UCN_MLP_BF16X3_TC) {   // round-1 kernel: tile-transposed fp32 interchange rows
    if ((rc = ucn_gather_tiles_from_z(f, z, nullptr, ray0, R, feat, ndc, stream))) return rc;
    if ((rc = ucn_mlp_forward_tiles(w, ndc, feat, angle, R, f->S, raw, mode, stream))) return rc;
  } else {
    if ((rc = ucn_gather_from_z(f, z, nullptr, ray0, R, feat, ndc, stream))) return rc;
    if ((rc = ucn_mlp_forward(w, ndc, feat, angle, R, f->S, raw, mode, stream))) return rc;
  }
  if ((rc = ucn_composite(raw, z, R, f->S, 0, out_rgb, out_depth, nullptr, nullptr, stream))) return rc;
  return UCN_OK;
}
