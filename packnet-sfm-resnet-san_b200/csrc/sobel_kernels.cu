// sobel_kernels.cu -- the multi-scale Sobel gradient term of SSISilogLoss (reference losses/ssi_silog_loss.py:12-50
// Gradient2D, :116-174 compute_gradient_loss; enabled by `gradient_weight: 0.2` in
// configs/train_resnet_san_ncdb_dual_head_640x384.yaml), forward and backward.
//
// Reference, per scale s < gradient_scales: pred and gt DEPTH maps (1 / max(inv, 1e-6), invalid pixels of both filled with
// 1e-6 by the sparse caller, i.e. 1e6 m) are bilinearly resized by 0.5^s (align_corners=False: the mean of the 2x2 centre
// pixels of every 2^s block), the mask by nearest (its pixel (y 2^s, x 2^s)); Sobel x / y of both maps (3x3, no padding);
// mean |Sobel(pred) - Sobel(gt)| over the masked interior, x plus y; scales with a masked pixel are averaged.
// Resize and Sobel are linear, so the kernels work on ONE map D = depth(pred) - depth(gt), exactly 0 at invalid pixels
// (the reference subtracts two sums that both carry the 1e6 fills: same value, minus its cancellation noise):
//   pyramid   D_s for all scales straight from the inputs (one launch)
//   sobel     per scale: Sobel of D_s at the masked interior pixels, |gx| and |gy| sums + count as per-CTA partials, and
//             a one-byte code (sign gx, sign gy) per pixel kept for the backward
//   finish    fixed-order sums -> loss, per-scale counts (state)
//   backward  one thread per full-size pixel GATHERS its gradient: the <= 1 pixel of each D_s it feeds, the <= 8 Sobel
//             centres around that pixel, their sign codes; times d depth / d inv = -depth^2.  No atomics: deterministic.
// Algorithmic bytes: forward reads pred + gt (8 B/px), backward reads them again and writes d pred (12 B/px).
#include <cuda_runtime.h>
#include <stdint.h>
#include "psfm_device.cuh"
#include "profile.h"

namespace psfm {

constexpr int kST = 256;
constexpr int kMaxGS = 8;
constexpr int kSobelState = 4 + 2 * kMaxGS;   // {loss, scales used, 0, 0, count_s x 8, loss_s x 8}

struct SobelCfg {
  int B, H, W, S;
  int Hs[kMaxGS], Ws[kMaxGS];
  unsigned int off[kMaxGS + 1];   // first pixel of level s in the pyramids (levels hold B * Hs * Ws pixels)
  unsigned int nblk[kMaxGS];      // CTAs of the sobel launch that carry pixels of level s
};

template <typename real>
__device__ __forceinline__ bool sg_valid(const real* gt, const uint8_t* mask, size_t i) {
  return __ldg(gt + i) > real(0) && (!mask || __ldg(mask + i) != 0);
}

// depth(pred) - depth(gt) at a full-size pixel, 0 where invalid (inv2depth: utils/depth.py:103-120)
template <typename real>
__device__ __forceinline__ real sg_diff(const real* pred, const real* gt, const uint8_t* mask, size_t i) {
  if (!sg_valid<real>(gt, mask, i)) return real(0);
  return real(1) / rmax(__ldg(pred + i), real(1e-6)) - real(1) / rmax(__ldg(gt + i), real(1e-6));
}

// one thread per pyramid pixel
template <typename real>
__global__ void __launch_bounds__(kST) sobel_pyramid_kernel(SobelCfg c, const real* pred, const real* gt, const uint8_t* mask,
                                                            real* D) {
  const unsigned int t = blockIdx.x * kST + threadIdx.x;
  if (t >= c.off[c.S]) return;
  int s = 0;
  while (t >= c.off[s + 1]) ++s;
  const unsigned int r = t - c.off[s], plane = (unsigned int)c.Hs[s] * c.Ws[s];
  const unsigned int b = r / plane, q = r - b * plane;
  const int yy = (int)(q / c.Ws[s]), xx = (int)(q - (unsigned int)yy * c.Ws[s]);
  const size_t base = (size_t)b * c.H * c.W;
  if (s == 0) {
    D[t] = sg_diff<real>(pred, gt, mask, base + q);
    return;
  }
  // src = (dst + 0.5) 2^s - 0.5: halfway between pixels 2^s dst + 2^(s-1) - 1 and + 2^(s-1)
  const int y0 = (yy << s) + (1 << (s - 1)) - 1, x0 = (xx << s) + (1 << (s - 1)) - 1;
  const real a = sg_diff<real>(pred, gt, mask, base + (size_t)y0 * c.W + x0);
  const real bq = sg_diff<real>(pred, gt, mask, base + (size_t)y0 * c.W + x0 + 1);
  const real cq = sg_diff<real>(pred, gt, mask, base + (size_t)(y0 + 1) * c.W + x0);
  const real d = sg_diff<real>(pred, gt, mask, base + (size_t)(y0 + 1) * c.W + x0 + 1);
  D[t] = real(0.5) * (real(0.5) * a + real(0.5) * bq) + real(0.5) * (real(0.5) * cq + real(0.5) * d);
}

__device__ __forceinline__ int sgn_code(double v) { return v > 0 ? 2 : (v < 0 ? 0 : 1); }

// grid (max CTAs of a level, S): CTA (k, s) owns pixels [k kST, (k + 1) kST) of level s over the whole batch
template <typename real>
__global__ void __launch_bounds__(kST) sobel_kernel(SobelCfg c, const real* gt, const uint8_t* mask, const real* D,
                                                    uint8_t* codes, double* part) {
  __shared__ real scratch[3 * (kST / 32)];
  const int s = blockIdx.y;
  if (blockIdx.x >= c.nblk[s]) return;
  const int Hs = c.Hs[s], Ws = c.Ws[s];
  const unsigned int plane = (unsigned int)Hs * Ws, total = plane * c.B;
  const unsigned int r = blockIdx.x * kST + threadIdx.x;
  real v[3] = {real(0), real(0), real(0)};
  if (r < total) {
    const unsigned int b = r / plane, q = r - b * plane;
    const int yy = (int)(q / Ws), xx = (int)(q - (unsigned int)yy * Ws);
    uint8_t code = 0xff;
    // mask resized by nearest: pixel (yy 2^s, xx 2^s); Sobel output (H-2, W-2) <-> centres 1..H-2 (:158-165)
    if (Hs >= 3 && Ws >= 3 && yy >= 1 && yy <= Hs - 2 && xx >= 1 && xx <= Ws - 2 &&
        sg_valid<real>(gt, mask, (size_t)b * c.H * c.W + (size_t)(yy << s) * c.W + (xx << s))) {
      const real* p = D + c.off[s] + (size_t)b * plane + (size_t)yy * Ws + xx;
      const real nw = p[-Ws - 1], n = p[-Ws], ne = p[-Ws + 1], w = p[-1], e = p[1], sw = p[Ws - 1], so = p[Ws], se = p[Ws + 1];
      const real gx = (ne + real(2) * e + se) - (nw + real(2) * w + sw);
      const real gy = (sw + real(2) * so + se) - (nw + real(2) * n + ne);
      v[0] = rabs(gx); v[1] = rabs(gy); v[2] = real(1);
      code = (uint8_t)(sgn_code((double)gx) | (sgn_code((double)gy) << 2));
    }
    codes[c.off[s] + r] = code;
  }
  block_reduce<real, 3>(v, scratch);
  if (threadIdx.x < 3) part[((size_t)s * gridDim.x + blockIdx.x) * 3 + threadIdx.x] = (double)v[threadIdx.x];
}

// one CTA of 32 S threads: warp s sums level s (lane-strided, fixed shuffle tree); thread 0 combines (:166-174)
template <typename real>
__global__ void sobel_finish_kernel(SobelCfg c, unsigned int stride, const double* part, double* state, real* out) {
  __shared__ double sums[kMaxGS][3];
  const int s = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double a[3] = {0.0, 0.0, 0.0};
  for (unsigned int i = lane; i < c.nblk[s]; i += 32)
    for (int k = 0; k < 3; ++k) a[k] += part[((size_t)s * stride + i) * 3 + k];
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a[k] += __shfl_down_sync(0xffffffffu, a[k], o);
    if (lane == 0) sums[s][k] = a[k];
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  double total = 0.0, used = 0.0;
  for (int k = 0; k < kSobelState; ++k) state[k] = 0.0;
  for (int l = 0; l < c.S; ++l) {
    const double n = sums[l][2];
    state[4 + l] = n;
    if (n > 0) {
      const double ls = sums[l][0] / n + sums[l][1] / n;
      state[4 + kMaxGS + l] = ls;
      total += ls;
      used += 1.0;
    }
  }
  const double loss = used > 0 ? total / used : 0.0;
  state[0] = loss;
  state[1] = used;
  out[0] = (real)loss;
}

// grid (CTAs of a plane, B): d loss / d pred_inv, gathered
template <typename real>
__global__ void __launch_bounds__(kST) sobel_grad_kernel(SobelCfg c, const real* pred, const real* gt, const uint8_t* mask,
                                                         const uint8_t* codes, const double* state, const real* grad_loss,
                                                         real* d_pred) {
  const unsigned int q = blockIdx.x * kST + threadIdx.x, plane = (unsigned int)c.H * c.W;
  if (q >= plane) return;
  const int b = blockIdx.y;
  const size_t i = (size_t)b * plane + q;
  const real pi = __ldg(pred + i);
  const double used = state[1];
  if (!sg_valid<real>(gt, mask, i) || !(pi >= real(1e-6)) || !(used > 0)) { d_pred[i] = real(0); return; }
  const int y = (int)(q / c.W), x = (int)(q - (unsigned int)y * c.W);
  double g = 0.0;
  for (int s = 0; s < c.S; ++s) {
    const double n = state[4 + s];
    if (!(n > 0)) continue;
    int yy = y, xx = x;
    double wgt = 1.0;
    if (s > 0) {
      const int half = 1 << (s - 1), m = (1 << s) - 1, ry = y & m, rx = x & m;
      if ((ry != half - 1 && ry != half) || (rx != half - 1 && rx != half)) continue;
      yy = y >> s; xx = x >> s;
      wgt = 0.25;
    }
    const int Hs = c.Hs[s], Ws = c.Ws[s];
    if (yy >= Hs || xx >= Ws) continue;
    const uint8_t* cm = codes + c.off[s] + (size_t)b * Hs * Ws;
    int acc = 0;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        if (dy == 0 && dx == 0) continue;
        const int cy = yy - dy, cx = xx - dx;          // Sobel centre that reads this pixel at offset (dy, dx)
        if (cy < 0 || cy >= Hs || cx < 0 || cx >= Ws) continue;
        const int code = cm[(size_t)cy * Ws + cx];
        if (code == 0xff) continue;
        const int sx = (code & 3) - 1, sy = (code >> 2) - 1;
        acc += sx * dx * (2 - (dy < 0 ? -dy : dy)) + sy * dy * (2 - (dx < 0 ? -dx : dx));
      }
    g += wgt * (double)acc / (n * used);
  }
  const real depth = real(1) / pi;
  d_pred[i] = (real)(-g) * depth * depth * grad_loss[0];
}

static inline size_t al256s(size_t x) { return (x + 255) / 256 * 256; }

static bool sobel_cfg(int B, int H, int W, int S, SobelCfg& c) {
  if (B < 1 || H < 1 || W < 1 || S < 1 || S > kMaxGS) return false;
  c.B = B; c.H = H; c.W = W; c.S = S;
  size_t off = 0;
  for (int s = 0; s < kMaxGS; ++s) {
    c.Hs[s] = s < S ? (H >> s) : 0;       // floor(H * 0.5^s): F.interpolate's output size
    c.Ws[s] = s < S ? (W >> s) : 0;
    c.off[s] = (unsigned int)off;
    const size_t n = (size_t)B * c.Hs[s] * c.Ws[s];
    c.nblk[s] = (unsigned int)((n + kST - 1) / kST);
    off += n;
    if (off >= ((size_t)1 << 31)) return false;
  }
  c.off[kMaxGS] = (unsigned int)off;
  for (int s = S; s < kMaxGS; ++s) c.off[s] = (unsigned int)off;
  return true;
}

}  // namespace psfm

int psfm_fail(int code, const char* fmt, ...);      // psfm_abi.cu
using namespace psfm;

extern "C" {

size_t psfm_sobel_gradient_code_bytes(int B, int H, int W, int scales) {
  SobelCfg c;
  if (!sobel_cfg(B, H, W, scales, c)) return 0;
  return al256s(c.off[kMaxGS]);
}

size_t psfm_sobel_gradient_workspace_bytes(int B, int H, int W, int scales) {
  SobelCfg c;
  if (!sobel_cfg(B, H, W, scales, c)) return 0;
  return al256s((size_t)c.off[kMaxGS] * sizeof(double)) + al256s((size_t)scales * c.nblk[0] * 3 * sizeof(double));
}

int psfm_sobel_gradient_loss(int B, int H, int W, int dt, int scales, const void* pred_inv, const void* gt_inv,
                             const void* mask, void* loss_out, void* state, void* codes, void* workspace,
                             size_t workspace_bytes, void* stream) {
  if (dt != PSFM_F32 && dt != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "psfm_sobel_gradient_loss: real_dtype must be f32|f64");
  SobelCfg c;
  if (!sobel_cfg(B, H, W, scales, c))
    return psfm_fail(PSFM_ERR_INVALID, "psfm_sobel_gradient_loss: need B, H, W >= 1, 1 <= scales <= %d and fewer than 2^31 pyramid pixels", kMaxGS);
  if (!pred_inv || !gt_inv || !loss_out || !state || !codes) return psfm_fail(PSFM_ERR_INVALID, "psfm_sobel_gradient_loss: bad argument");
  if (!workspace || workspace_bytes < psfm_sobel_gradient_workspace_bytes(B, H, W, scales))
    return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_sobel_gradient_loss: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned int total = c.off[kMaxGS];
  double* part = (double*)((char*)workspace + al256s((size_t)total * sizeof(double)));
  count_launches(3);
  if (dt == PSFM_F64) {
    typedef double real;
    sobel_pyramid_kernel<real><<<(total + kST - 1) / kST, kST, 0, st>>>(c, (const real*)pred_inv, (const real*)gt_inv, (const uint8_t*)mask, (real*)workspace);
    sobel_kernel<real><<<dim3(c.nblk[0], scales), kST, 0, st>>>(c, (const real*)gt_inv, (const uint8_t*)mask, (const real*)workspace, (uint8_t*)codes, part);
    sobel_finish_kernel<real><<<1, 32 * scales, 0, st>>>(c, c.nblk[0], part, (double*)state, (real*)loss_out);
  } else {
    typedef float real;
    sobel_pyramid_kernel<real><<<(total + kST - 1) / kST, kST, 0, st>>>(c, (const real*)pred_inv, (const real*)gt_inv, (const uint8_t*)mask, (real*)workspace);
    sobel_kernel<real><<<dim3(c.nblk[0], scales), kST, 0, st>>>(c, (const real*)gt_inv, (const uint8_t*)mask, (const real*)workspace, (uint8_t*)codes, part);
    sobel_finish_kernel<real><<<1, 32 * scales, 0, st>>>(c, c.nblk[0], part, (double*)state, (real*)loss_out);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return psfm_fail(PSFM_ERR_CUDA, "psfm_sobel_gradient_loss: %s", cudaGetErrorString(e));
  return PSFM_OK;
}

int psfm_sobel_gradient_loss_backward(int B, int H, int W, int dt, int scales, const void* pred_inv, const void* gt_inv,
                                      const void* mask, const void* state, const void* codes, const void* grad_loss,
                                      void* d_pred, void* stream) {
  if (dt != PSFM_F32 && dt != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "psfm_sobel_gradient_loss_backward: real_dtype must be f32|f64");
  SobelCfg c;
  if (!sobel_cfg(B, H, W, scales, c)) return psfm_fail(PSFM_ERR_INVALID, "psfm_sobel_gradient_loss_backward: bad shape");
  if (!pred_inv || !gt_inv || !state || !codes || !grad_loss || !d_pred)
    return psfm_fail(PSFM_ERR_INVALID, "psfm_sobel_gradient_loss_backward: bad argument");
  const dim3 grid((unsigned)(((size_t)H * W + kST - 1) / kST), B);
  count_launches(1);
  if (dt == PSFM_F64)
    sobel_grad_kernel<double><<<grid, kST, 0, (cudaStream_t)stream>>>(c, (const double*)pred_inv, (const double*)gt_inv, (const uint8_t*)mask,
                                                                      (const uint8_t*)codes, (const double*)state, (const double*)grad_loss, (double*)d_pred);
  else
    sobel_grad_kernel<float><<<grid, kST, 0, (cudaStream_t)stream>>>(c, (const float*)pred_inv, (const float*)gt_inv, (const uint8_t*)mask,
                                                                     (const uint8_t*)codes, (const double*)state, (const float*)grad_loss, (float*)d_pred);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return psfm_fail(PSFM_ERR_CUDA, "psfm_sobel_gradient_loss_backward: %s", cudaGetErrorString(e));
  return PSFM_OK;
}

}  // extern "C"
