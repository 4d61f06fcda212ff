// psfm_abi.cu -- extern "C" boundary of libpsfm_b200.so (declared in include/psfm.h).
// Validates arguments, converts to the internal parameter block and enqueues kernels on the
// caller's stream.  No allocation, no synchronisation, no global mutable state besides the
// thread-local error string.
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include "../../include/psfm.h"
#include "profile.h"
#include <atomic>
#include <mutex>
#include <vector>

namespace psfm {
size_t loss_workspace_bytes(const psfm_loss_cfg* cfg, const psfm_loss_io* io);
cudaError_t loss_forward_dispatch(const psfm_loss_cfg*, const psfm_loss_io*, void*, void*, cudaStream_t);
cudaError_t loss_backward_dispatch(const psfm_loss_cfg*, const psfm_loss_io*, void*, const void*, cudaStream_t);
cudaError_t loss_fused_dispatch(const psfm_loss_cfg*, const psfm_loss_io*, void*, void*, cudaStream_t);
cudaError_t warp_forward_dispatch(int, int, int, int, int, int, const void*, const void*, const void*, const void*,
                                  const void*, void*, cudaStream_t);
size_t warp_workspace_bytes(int, int, int, int);
cudaError_t warp_backward_dispatch(int, int, int, int, int, int, const void*, const void*, const void*, const void*,
                                   const void*, const void*, void*, void*, void*, void*, cudaStream_t);
cudaError_t resize_dispatch(int, int, int, int, int, int, int, const void*, void*, cudaStream_t);
cudaError_t resize_backward_dispatch(int, int, int, int, int, int, int, const void*, void*, cudaStream_t);
cudaError_t scale_dispatch(void*, const void*, size_t, int, cudaStream_t);
}  // namespace psfm

static thread_local char g_err[512] = "";

// ---- launch accounting / optional event timing ---------------------------------------------
namespace psfm {
static std::atomic<unsigned long long> g_launches{0};
static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;
struct ProfRec { cudaEvent_t a, b; int tag; };
static std::vector<ProfRec> g_prof;

void count_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

ProfScope::ProfScope(cudaStream_t st, int tag, int launches) : st_(st), slot_(-1) {
  count_launches(launches);
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (g_prof.size() >= 65536) return;
  ProfRec r;
  r.tag = tag;
  if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return;
  cudaEventRecord(r.a, st);
  slot_ = (int)g_prof.size();
  g_prof.push_back(r);
}
ProfScope::~ProfScope() {
  if (slot_ < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaEventRecord(g_prof[slot_].b, st_);
}
}  // namespace psfm

int psfm_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

static int cuda_result(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return PSFM_OK;
  return psfm_fail(PSFM_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

static int check_cfg(const psfm_loss_cfg* c) {
  if (!c) return psfm_fail(PSFM_ERR_INVALID, "cfg is NULL");
  if (c->B < 1) return psfm_fail(PSFM_ERR_INVALID, "B must be >= 1 (got %d)", c->B);
  if (c->C < 1 || c->C > PSFM_MAX_CTX) return psfm_fail(PSFM_ERR_INVALID, "C must be in 1..%d (got %d)", PSFM_MAX_CTX, c->C);
  if (c->n < 1 || c->n > PSFM_MAX_SCALES)
    return psfm_fail(PSFM_ERR_INVALID, "n must be in 1..%d (got %d)", PSFM_MAX_SCALES, c->n);
  if (c->H < 2 || c->W < 2) return psfm_fail(PSFM_ERR_INVALID, "image must be at least 2x2 (got %dx%d)", c->H, c->W);
  for (int s = 0; s < c->n; ++s)
    if (c->Hs[s] < 2 || c->Ws[s] < 2)
      return psfm_fail(PSFM_ERR_INVALID, "scale %d must be at least 2x2 (got %dx%d)", s, c->Hs[s], c->Ws[s]);
  if (c->real_dtype != PSFM_F32 && c->real_dtype != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "real_dtype must be f32|f64");
  if (c->real_dtype == PSFM_F64 && c->img_dtype != PSFM_F64)
    return psfm_fail(PSFM_ERR_INVALID, "f64 arithmetic requires f64 images");
  if (c->real_dtype == PSFM_F32 && c->img_dtype != PSFM_F32 && c->img_dtype != PSFM_BF16 && c->img_dtype != PSFM_U8)
    return psfm_fail(PSFM_ERR_INVALID, "f32 arithmetic requires f32, bf16 or u8 images");
  if (c->img_dtype == PSFM_U8) {
    bool ok = c->reduce_op == PSFM_REDUCE_MIN && c->ssim_w > 0.0 && !(c->clip_loss > 0.0) && c->C <= 8 && c->H >= 4 && c->W >= 4;
    for (int s = 0; s < c->n; ++s) ok = ok && c->Hs[s] == c->H && c->Ws[s] == c->W;
    if (!ok)
      return psfm_fail(PSFM_ERR_INVALID, "u8 images are read by the sweep kernels only, with every scale at image resolution "
                       "(photometric_reduce_op='min', ssim_loss_weight > 0, clip_loss = 0; inverse depths full-size or "
                       "upsampled in-kernel); convert to float before the loss otherwise");
  }
  if (c->padding != PSFM_PAD_ZEROS && c->padding != PSFM_PAD_BORDER)
    return psfm_fail(PSFM_ERR_INVALID, "padding must be zeros|border");
  if (c->reduce_op != PSFM_REDUCE_MIN && c->reduce_op != PSFM_REDUCE_MEAN)
    return psfm_fail(PSFM_ERR_INVALID, "Unknown photometric_reduce_op: %d", c->reduce_op);
  if (c->automask && c->reduce_op != PSFM_REDUCE_MIN)
    return psfm_fail(PSFM_ERR_INVALID, "For automasking only the min photometric_reduce_op is supported.");
  for (int s = 0; s < c->n; ++s) {
    const bool ups = (c->Hd[s] > 0 || c->Wd[s] > 0) && (c->Hd[s] != c->Hs[s] || c->Wd[s] != c->Ws[s]);
    if (!ups) continue;
    if (c->Hd[s] < 1 || c->Wd[s] < 1 || c->Hd[s] > c->Hs[s] || c->Wd[s] > c->Ws[s])
      return psfm_fail(PSFM_ERR_INVALID, "scale %d: in-kernel upsampling needs 1 <= Hd <= Hs and 1 <= Wd <= Ws (got %dx%d -> %dx%d)", s,
                       c->Hd[s], c->Wd[s], c->Hs[s], c->Ws[s]);
    if (c->reduce_op != PSFM_REDUCE_MIN || !(c->ssim_w > 0.0) || c->clip_loss > 0.0 || c->C > 8 || c->Hs[s] < 4 || c->Ws[s] < 4)
      return psfm_fail(PSFM_ERR_INVALID, "in-kernel upsampling of the inverse depths is implemented by the sweep kernels only "
                       "(photometric_reduce_op='min', ssim_loss_weight > 0, clip_loss = 0); upsample before the loss otherwise");
  }
  if (c->camera_model != PSFM_CAM_PINHOLE && c->camera_model != PSFM_CAM_FISHEYE)
    return psfm_fail(PSFM_ERR_INVALID, "camera_model must be pinhole|fisheye (got %d)", c->camera_model);
  if (c->camera_model == PSFM_CAM_FISHEYE) {
    bool ok = c->reduce_op == PSFM_REDUCE_MIN && c->ssim_w > 0.0 && !(c->clip_loss > 0.0) && c->C <= 8;
    for (int s = 0; s < c->n; ++s) ok = ok && c->Hs[s] >= 4 && c->Ws[s] >= 4;
    if (!ok)
      return psfm_fail(PSFM_ERR_INVALID, "the fisheye camera model is implemented by the sweep kernels only "
                       "(photometric_reduce_op='min', ssim_loss_weight > 0, clip_loss = 0, scales of at least 4x4)");
  }
  const int groups = (c->automask ? 2 : 1) * c->C * (c->ssim_w > 0.0 ? 1 : 3);
  if (groups > 254) return psfm_fail(PSFM_ERR_INVALID, "too many candidates (%d)", groups);
  return PSFM_OK;
}

static int check_io(const psfm_loss_cfg* c, const psfm_loss_io* io, bool need_grads) {
  if (!io) return psfm_fail(PSFM_ERR_INVALID, "io is NULL");
  if (!io->K || !io->ref_K) return psfm_fail(PSFM_ERR_INVALID, "K / ref_K is NULL");
  for (int j = 0; j < c->C; ++j) {
    if (!io->pose[j]) return psfm_fail(PSFM_ERR_INVALID, "pose[%d] is NULL", j);
    if (need_grads && !io->d_pose[j]) return psfm_fail(PSFM_ERR_INVALID, "d_pose[%d] is NULL", j);
  }
  for (int s = 0; s < c->n; ++s) {
    if (!io->image[s] || !io->inv_depth[s]) return psfm_fail(PSFM_ERR_INVALID, "image/inv_depth[%d] is NULL", s);
    for (int j = 0; j < c->C; ++j)
      if (!io->ctx[s][j]) return psfm_fail(PSFM_ERR_INVALID, "ctx[%d][%d] is NULL", s, j);
    if (need_grads && !io->d_inv_depth[s]) return psfm_fail(PSFM_ERR_INVALID, "d_inv_depth[%d] is NULL", s);
  }
  return PSFM_OK;
}

extern "C" {

int psfm_abi_version(void) { return PSFM_ABI_VERSION; }

unsigned long long psfm_launch_count(void) { return psfm::g_launches.load(); }

int psfm_profile_enable(int on) {
  psfm::g_prof_on.store(on ? 1 : 0);
  return PSFM_OK;
}

int psfm_profile_read(float* ms, int* tags, int max_records) {
  std::lock_guard<std::mutex> lk(psfm::g_prof_mu);
  int n = 0;
  for (auto& r : psfm::g_prof) {
    cudaEventSynchronize(r.b);
    float t = 0.f;
    cudaEventElapsedTime(&t, r.a, r.b);
    if (n < max_records && ms && tags) { ms[n] = t; tags[n] = r.tag; ++n; }
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  psfm::g_prof.clear();
  return n;
}

const char* psfm_last_error(void) { return g_err; }

size_t psfm_loss_workspace_bytes(const psfm_loss_cfg* cfg) {
  if (check_cfg(cfg) != PSFM_OK) return 0;
  return psfm::loss_workspace_bytes(cfg, nullptr);
}

size_t psfm_loss_workspace_bytes_io(const psfm_loss_cfg* cfg, const psfm_loss_io* io) {
  if (check_cfg(cfg) != PSFM_OK || !io) return 0;
  return psfm::loss_workspace_bytes(cfg, io);
}

int psfm_loss_forward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* workspace, size_t workspace_bytes,
                      void* loss_out, void* stream) {
  int rc = check_cfg(cfg);
  if (rc) return rc;
  if ((rc = check_io(cfg, io, false))) return rc;
  for (int s = 0; s < cfg->n; ++s) {
    if (!io->sel[s]) return psfm_fail(PSFM_ERR_INVALID, "sel[%d] is NULL", s);
    if (cfg->Hd[s] > 0 && (cfg->Hd[s] != cfg->Hs[s] || cfg->Wd[s] != cfg->Ws[s]))
      return psfm_fail(PSFM_ERR_INVALID, "in-kernel upsampling: use psfm_loss_forward_backward");
  }
  if (cfg->camera_model == PSFM_CAM_FISHEYE) return psfm_fail(PSFM_ERR_INVALID, "fisheye cameras: use psfm_loss_forward_backward");
  if (cfg->img_dtype == PSFM_U8) return psfm_fail(PSFM_ERR_INVALID, "u8 images: use psfm_loss_forward_backward");
  if (!loss_out || !workspace) return psfm_fail(PSFM_ERR_INVALID, "loss_out / workspace is NULL");
  if (workspace_bytes < psfm::loss_workspace_bytes(cfg, io))
    return psfm_fail(PSFM_ERR_WORKSPACE, "workspace too small: %zu < %zu", workspace_bytes, psfm::loss_workspace_bytes(cfg, io));
  return cuda_result(psfm::loss_forward_dispatch(cfg, io, workspace, loss_out, (cudaStream_t)stream), "psfm_loss_forward");
}

int psfm_loss_backward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* workspace, size_t workspace_bytes,
                       const void* grad_loss, void* stream) {
  int rc = check_cfg(cfg);
  if (rc) return rc;
  if ((rc = check_io(cfg, io, true))) return rc;
  for (int s = 0; s < cfg->n; ++s) {
    if (!io->sel[s]) return psfm_fail(PSFM_ERR_INVALID, "sel[%d] is NULL", s);
    if (cfg->Hd[s] > 0 && (cfg->Hd[s] != cfg->Hs[s] || cfg->Wd[s] != cfg->Ws[s]))
      return psfm_fail(PSFM_ERR_INVALID, "in-kernel upsampling: use psfm_loss_forward_backward");
  }
  if (cfg->camera_model == PSFM_CAM_FISHEYE) return psfm_fail(PSFM_ERR_INVALID, "fisheye cameras: use psfm_loss_forward_backward");
  if (cfg->img_dtype == PSFM_U8) return psfm_fail(PSFM_ERR_INVALID, "u8 images: use psfm_loss_forward_backward");
  if (!grad_loss || !workspace) return psfm_fail(PSFM_ERR_INVALID, "grad_loss / workspace is NULL");
  if (workspace_bytes < psfm::loss_workspace_bytes(cfg, io))
    return psfm_fail(PSFM_ERR_WORKSPACE, "workspace too small: %zu < %zu", workspace_bytes, psfm::loss_workspace_bytes(cfg, io));
  return cuda_result(psfm::loss_backward_dispatch(cfg, io, workspace, grad_loss, (cudaStream_t)stream),
                     "psfm_loss_backward");
}

int psfm_loss_forward_backward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* workspace,
                               size_t workspace_bytes, void* loss_out, void* stream) {
  int rc = check_cfg(cfg);
  if (rc) return rc;
  if ((rc = check_io(cfg, io, true))) return rc;
  if (!loss_out || !workspace) return psfm_fail(PSFM_ERR_INVALID, "loss_out / workspace is NULL");
  if (workspace_bytes < psfm::loss_workspace_bytes(cfg, io))
    return psfm_fail(PSFM_ERR_WORKSPACE, "workspace too small: %zu < %zu", workspace_bytes, psfm::loss_workspace_bytes(cfg, io));
  return cuda_result(psfm::loss_fused_dispatch(cfg, io, workspace, loss_out, (cudaStream_t)stream),
                     "psfm_loss_forward_backward");
}

int psfm_scale_inplace(void* x, const void* scale, size_t count, int real_dtype, void* stream) {
  if ((!x && count) || !scale) return psfm_fail(PSFM_ERR_INVALID, "x / scale is NULL");
  if (real_dtype != PSFM_F32 && real_dtype != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "real_dtype must be f32|f64");
  return cuda_result(psfm::scale_dispatch(x, scale, count, real_dtype, (cudaStream_t)stream), "psfm_scale_inplace");
}

static int check_warp(int B, int H, int W, int img_dtype, int real_dtype, int padding) {
  if (B < 1 || H < 2 || W < 2) return psfm_fail(PSFM_ERR_INVALID, "bad shape B=%d H=%d W=%d", B, H, W);
  if (real_dtype != PSFM_F32 && real_dtype != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "real_dtype must be f32|f64");
  if (real_dtype == PSFM_F64 && img_dtype != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "f64 arithmetic requires f64 images");
  if (real_dtype == PSFM_F32 && img_dtype != PSFM_F32 && img_dtype != PSFM_BF16)
    return psfm_fail(PSFM_ERR_INVALID, "f32 arithmetic requires f32 or bf16 images");
  if (padding != PSFM_PAD_ZEROS && padding != PSFM_PAD_BORDER) return psfm_fail(PSFM_ERR_INVALID, "padding must be zeros|border");
  return PSFM_OK;
}

int psfm_warp_forward(int B, int H, int W, int img_dtype, int real_dtype, int padding, const void* ref_image,
                      const void* depth, const void* K, const void* ref_K, const void* pose, void* out, void* stream) {
  int rc = check_warp(B, H, W, img_dtype, real_dtype, padding);
  if (rc) return rc;
  if (!ref_image || !depth || !K || !ref_K || !pose || !out) return psfm_fail(PSFM_ERR_INVALID, "NULL pointer argument");
  return cuda_result(psfm::warp_forward_dispatch(B, H, W, img_dtype, real_dtype, padding, ref_image, depth, K, ref_K,
                                                 pose, out, (cudaStream_t)stream), "psfm_warp_forward");
}

size_t psfm_warp_workspace_bytes(int B, int H, int W, int real_dtype) {
  if (B < 1 || H < 1 || W < 1) return 0;
  return psfm::warp_workspace_bytes(B, H, W, real_dtype);
}

int psfm_warp_backward(int B, int H, int W, int img_dtype, int real_dtype, int padding, const void* ref_image,
                       const void* depth, const void* K, const void* ref_K, const void* pose, const void* grad_out,
                       void* d_depth, void* d_pose, void* d_ref_image, void* workspace, size_t workspace_bytes,
                       void* stream) {
  int rc = check_warp(B, H, W, img_dtype, real_dtype, padding);
  if (rc) return rc;
  if (!ref_image || !depth || !K || !ref_K || !pose || !grad_out || !d_depth || !d_pose || !workspace)
    return psfm_fail(PSFM_ERR_INVALID, "NULL pointer argument");
  if (d_ref_image && img_dtype == PSFM_BF16)
    return psfm_fail(PSFM_ERR_INVALID, "gradient w.r.t. a bf16 reference image is not supported");
  if (workspace_bytes < psfm::warp_workspace_bytes(B, H, W, real_dtype))
    return psfm_fail(PSFM_ERR_WORKSPACE, "workspace too small");
  return cuda_result(psfm::warp_backward_dispatch(B, H, W, img_dtype, real_dtype, padding, ref_image, depth, K, ref_K,
                                                  pose, grad_out, d_depth, d_pose, d_ref_image, workspace,
                                                  (cudaStream_t)stream), "psfm_warp_backward");
}

int psfm_resize(int N, int Hi, int Wi, int Ho, int Wo, int dtype, int mode, const void* in, void* out, void* stream) {
  if (N < 1 || Hi < 1 || Wi < 1 || Ho < 1 || Wo < 1) return psfm_fail(PSFM_ERR_INVALID, "bad resize shape");
  if (dtype != PSFM_F32 && dtype != PSFM_BF16 && dtype != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "bad dtype");
  if (mode < 0 || mode > 2)
    return psfm_fail(PSFM_ERR_INVALID, "mode must be 0 (bilinear, align_corners=True), 1 (nearest) or 2 (bilinear, align_corners=False)");
  if (!in || !out) return psfm_fail(PSFM_ERR_INVALID, "NULL pointer argument");
  return cuda_result(psfm::resize_dispatch(N, Hi, Wi, Ho, Wo, dtype, mode, in, out, (cudaStream_t)stream), "psfm_resize");
}

int psfm_resize_backward(int N, int Hi, int Wi, int Ho, int Wo, int dtype, int mode, const void* grad_out, void* grad_in,
                         void* stream) {
  if (N < 1 || Hi < 1 || Wi < 1 || Ho < 1 || Wo < 1) return psfm_fail(PSFM_ERR_INVALID, "bad resize shape");
  if (dtype != PSFM_F32 && dtype != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "gradients are f32 or f64");
  if (mode < 0 || mode > 2) return psfm_fail(PSFM_ERR_INVALID, "mode must be 0, 1 or 2");
  if (!grad_out || !grad_in) return psfm_fail(PSFM_ERR_INVALID, "NULL pointer argument");
  return cuda_result(psfm::resize_backward_dispatch(N, Hi, Wi, Ho, Wo, dtype, mode, grad_out, grad_in, (cudaStream_t)stream),
                     "psfm_resize_backward");
}

}  // extern "C"
