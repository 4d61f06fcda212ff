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
//   finish    fixed-order sums -> loss, per-scale counts (state), the backward's per-scale coefficients (behind the codes)
//   backward  one thread per full-size pixel GATHERS its gradient: the <= 1 pixel of each D_s it feeds, the <= 8 Sobel
//             centres around that pixel, their sign codes; times d depth / d inv = -depth^2.  No atomics: deterministic.
// Algorithmic bytes: forward reads pred + gt (8 B/px), backward reads them again and writes d pred (12 B/px).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include "psfm_device.cuh"
#include "profile.h"

namespace psfm {

constexpr int kST = 256;
constexpr int kNoCode = 5;  // sign code of a pixel without a Sobel response: decodes to (0, 0)
constexpr int kSP = 8;      // pixels per thread of sobel_kernel
constexpr int kFT = 512;    // threads of the finish CTA
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
  const real g = __ldg(gt + i), p = __ldg(pred + i);     // both loads in flight before the validity test
  if (!(g > real(0)) || (mask && __ldg(mask + i) == 0)) return real(0);
  return real(1) / rmax(p, real(1e-6)) - real(1) / rmax(g, real(1e-6));
}

// one thread per pyramid pixel, from pyramid pixel `first` on (any shape; the levels >= 3 behind sobel_pyramid4_kernel)
template <typename real>
__global__ void __launch_bounds__(kST) sobel_pyramid_kernel(SobelCfg c, unsigned int first, const real* pred, const real* gt,
                                                            const uint8_t* mask, real* D) {
  const unsigned int t = first + blockIdx.x * kST + threadIdx.x;
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

// four consecutive elements through the widest loads (16-byte aligned address)
__device__ __forceinline__ void ld4(const float* p, float v[4]) {
  const float4 t = __ldg(reinterpret_cast<const float4*>(p));
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void ld4(const double* p, double v[4]) {
  const double2 a = __ldg(reinterpret_cast<const double2*>(p)), b = __ldg(reinterpret_cast<const double2*>(p) + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4(float* p, const float v[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4(double* p, const double v[4]) {
  reinterpret_cast<double2*>(p)[0] = make_double2(v[0], v[1]);
  reinterpret_cast<double2*>(p)[1] = make_double2(v[2], v[3]);
}
template <typename real>
__device__ __forceinline__ real mean4(real a, real b, real cq, real d) {      // the expression of sobel_pyramid_kernel
  return real(0.5) * (real(0.5) * a + real(0.5) * b) + real(0.5) * (real(0.5) * cq + real(0.5) * d);
}

// H and W multiples of 4: one thread per 4 x 4 block of full-size pixels -- 16-byte loads of its four rows, the 16
// differences once, and from them the block's pixels of level 0 (4 x 4), level 1 (2 x 2: the 2 x 2 block means) and level 2
// (1: the mean of the block's centre 2 x 2).  Levels >= 3 straddle blocks and come from sobel_pyramid_kernel.
template <typename real>
__global__ void __launch_bounds__(kST) sobel_pyramid4_kernel(SobelCfg c, const real* pred, const real* gt, const uint8_t* mask,
                                                             real* D) {
  const unsigned int nbx = (unsigned int)c.W >> 2, nby = (unsigned int)c.H >> 2, per = nbx * nby;
  const unsigned int t = blockIdx.x * kST + threadIdx.x;
  if (t >= per * c.B) return;
  const unsigned int b = t / per, q = t - b * per, Y = q / nbx, X = q - Y * nbx;
  const size_t base = (size_t)b * c.H * c.W + (size_t)(4 * Y) * c.W + 4 * X;
  real g[4][4], p[4][4];
  unsigned int m[4] = {0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u};
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    ld4(gt + base + (size_t)r * c.W, g[r]);
    ld4(pred + base + (size_t)r * c.W, p[r]);
    if (mask) m[r] = __ldg(reinterpret_cast<const unsigned int*>(mask + base + (size_t)r * c.W));
  }
  real d[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool ok = g[r][k] > real(0) && ((m[r] >> (8 * k)) & 0xffu) != 0;
      d[r][k] = ok ? real(1) / rmax(p[r][k], real(1e-6)) - real(1) / rmax(g[r][k], real(1e-6)) : real(0);
    }
    st4(D + base + (size_t)r * c.W, d[r]);
  }
  if (c.S > 1) {
    const int W1 = c.Ws[1];
    real* D1 = D + c.off[1] + (size_t)b * c.Hs[1] * W1 + (size_t)(2 * Y) * W1 + 2 * X;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      D1[(size_t)j * W1] = mean4<real>(d[2 * j][0], d[2 * j][1], d[2 * j + 1][0], d[2 * j + 1][1]);
      D1[(size_t)j * W1 + 1] = mean4<real>(d[2 * j][2], d[2 * j][3], d[2 * j + 1][2], d[2 * j + 1][3]);
    }
  }
  if (c.S > 2) D[c.off[2] + (size_t)b * c.Hs[2] * c.Ws[2] + (size_t)Y * c.Ws[2] + X] = mean4<real>(d[1][1], d[1][2], d[2][1], d[2][2]);
}

__device__ __forceinline__ int sgn_code(double v) { return v > 0 ? 2 : (v < 0 ? 0 : 1); }

// grid (max CTAs of a level, S): CTA (k, s) owns pixels [k kST kSP, (k + 1) kST kSP) of level s over the whole batch,
// kSP per thread at stride kST (coalesced), so that the finish kernel has few partials to add
// QUAD0 (W a multiple of 4, 16-byte aligned maps): at level 0 a thread takes quads of a row -- three rows of D as one
// 16-byte load + two scalars each (18 values for 4 responses instead of 32), gt and mask as vectors, the 4 codes as one word;
// same response expressions, same pixels per CTA (the partials keep their layout)
template <typename real, bool QUAD0>
__global__ void __launch_bounds__(kST) sobel_kernel(SobelCfg c, const real* gt, const uint8_t* mask, const real* D,
                                                    uint8_t* codes, double* part) {
  __shared__ real scratch[3 * (kST / 32)];
  const int s = blockIdx.y;
  if (blockIdx.x >= c.nblk[s]) return;
  const int Hs = c.Hs[s], Ws = c.Ws[s];
  const unsigned int plane = (unsigned int)Hs * Ws, total = plane * c.B;
  const bool has_out = Hs >= 3 && Ws >= 3;
  real v[3] = {real(0), real(0), real(0)};
  if (QUAD0 && s == 0) {
#pragma unroll
    for (int j = 0; j < kSP / 4; ++j) {
      const unsigned int r = ((blockIdx.x * (kSP / 4) + j) * kST + threadIdx.x) * 4;
      if (r >= total) break;
      const unsigned int b = r / plane, q = r - b * plane;
      const int yy = (int)(q / Ws), xx = (int)(q - (unsigned int)yy * Ws);
      unsigned int word = 0x01010101u * (unsigned int)kNoCode;
      if (has_out && yy >= 1 && yy <= Hs - 2) {
        real g4[4];
        ld4(gt + r, g4);
        const unsigned int mk = mask ? __ldg(reinterpret_cast<const unsigned int*>(mask + r)) : 0x01010101u;
        const real* p = D + r;
        real n[6], m[6], o[6];
        ld4(p - Ws, n + 1); ld4(p, m + 1); ld4(p + Ws, o + 1);
        const bool hl = xx > 0, hr = xx + 4 < Ws;
        n[0] = hl ? p[-Ws - 1] : real(0); m[0] = hl ? p[-1] : real(0); o[0] = hl ? p[Ws - 1] : real(0);
        n[5] = hr ? p[-Ws + 4] : real(0); m[5] = hr ? p[4] : real(0); o[5] = hr ? p[Ws + 4] : real(0);
        word = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          unsigned int code = kNoCode;
          if (xx + k >= 1 && xx + k <= Ws - 2 && g4[k] > real(0) && ((mk >> (8 * k)) & 0xffu) != 0) {
            const real gx = (n[k + 2] + real(2) * m[k + 2] + o[k + 2]) - (n[k] + real(2) * m[k] + o[k]);
            const real gy = (o[k] + real(2) * o[k + 1] + o[k + 2]) - (n[k] + real(2) * n[k + 1] + n[k + 2]);
            v[0] += rabs(gx); v[1] += rabs(gy); v[2] += real(1);
            code = (unsigned int)(sgn_code((double)gx) | (sgn_code((double)gy) << 2));
          }
          word |= code << (8 * k);
        }
      }
      *reinterpret_cast<unsigned int*>(codes + r) = word;
    }
    block_reduce<real, 3>(v, scratch);
    if (threadIdx.x < 3) part[((size_t)s * gridDim.x + blockIdx.x) * 3 + threadIdx.x] = (double)v[threadIdx.x];
    return;
  }
#pragma unroll 4
  for (int j = 0; j < kSP; ++j) {
    const unsigned int r = (blockIdx.x * kSP + j) * kST + threadIdx.x;
    if (r >= total) break;
    const unsigned int b = r / plane, q = r - b * plane;
    const int yy = (int)(q / Ws), xx = (int)(q - (unsigned int)yy * Ws);
    uint8_t code = kNoCode;
    // mask resized by nearest: pixel (yy 2^s, xx 2^s); Sobel output (H-2, W-2) <-> centres 1..H-2 (:158-165)
    if (has_out && yy >= 1 && yy <= Hs - 2 && xx >= 1 && xx <= Ws - 2 &&
        sg_valid<real>(gt, mask, (size_t)b * c.H * c.W + (size_t)(yy << s) * c.W + (xx << s))) {
      const real* p = D + c.off[s] + (size_t)b * plane + (size_t)yy * Ws + xx;
      const real nw = p[-Ws - 1], n = p[-Ws], ne = p[-Ws + 1], w = p[-1], e = p[1], sw = p[Ws - 1], so = p[Ws], se = p[Ws + 1];
      const real gx = (ne + real(2) * e + se) - (nw + real(2) * w + sw);
      const real gy = (sw + real(2) * so + se) - (nw + real(2) * n + ne);
      v[0] += rabs(gx); v[1] += rabs(gy); v[2] += real(1);
      code = (uint8_t)(sgn_code((double)gx) | (sgn_code((double)gy) << 2));
    }
    codes[c.off[s] + r] = code;
  }
  block_reduce<real, 3>(v, scratch);
  if (threadIdx.x < 3) part[((size_t)s * gridDim.x + blockIdx.x) * 3 + threadIdx.x] = (double)v[threadIdx.x];
}

// one CTA of kFT threads: every level's partials are added thread-strided, then by a fixed shuffle tree and a fixed-order
// sum over the warps (run-to-run identical); thread 0 combines the levels (:166-174)
template <typename real>
__global__ void __launch_bounds__(kFT) sobel_finish_kernel(SobelCfg c, unsigned int stride, const double* part, double* state, real* out,
                                                               double* coef) {
  __shared__ double wsum[kFT / 32][kMaxGS * 3];
  __shared__ double sums[kMaxGS][3];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int s = 0; s < c.S; ++s) {
    double a[3] = {0.0, 0.0, 0.0};
    for (unsigned int i = threadIdx.x; i < c.nblk[s]; i += kFT)
      for (int k = 0; k < 3; ++k) a[k] += part[((size_t)s * stride + i) * 3 + k];
    for (int k = 0; k < 3; ++k) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a[k] += __shfl_down_sync(0xffffffffu, a[k], o);
      if (lane == 0) wsum[warp][s * 3 + k] = a[k];
    }
  }
  __syncthreads();
  if ((int)threadIdx.x < c.S * 3) {
    double t = 0.0;
    for (int w = 0; w < kFT / 32; ++w) t += wsum[w][threadIdx.x];
    sums[threadIdx.x / 3][threadIdx.x % 3] = t;
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
  // d loss / d (one Sobel response sign) per level, times the bilinear weight of a level-0 pixel in D_s (0: level unused);
  // kept behind the sign codes for the backward kernel
  for (int l = 0; l < kMaxGS; ++l) {
    const double n = l < c.S ? sums[l][2] : 0.0;
    coef[l] = (used > 0 && n > 0) ? (l ? 0.25 : 1.0) / (n * used) : 0.0;
  }
}

// grid (CTAs of a plane, B): d loss / d pred_inv, gathered.  Every load of a pixel that does not depend on a loaded value
// (gt, pred, mask, the eight level-0 sign codes, the coefficients) is issued before the first test.
template <typename real>
__global__ void __launch_bounds__(kST) sobel_grad_kernel(SobelCfg c, const real* pred, const real* gt, const uint8_t* mask,
                                                         const uint8_t* codes, const double* coef, const real* grad_loss,
                                                         real* d_pred) {
  const unsigned int q = blockIdx.x * kST + threadIdx.x, plane = (unsigned int)c.H * c.W;
  if (q >= plane) return;
  const int b = blockIdx.y;
  const size_t i = (size_t)b * plane + q;
  const real gi = __ldg(gt + i), pi = __ldg(pred + i);
  const int mi = mask ? (int)__ldg(mask + i) : 1;
  const int y = (int)(q / c.W), x = (int)(q - (unsigned int)y * c.W);
  const real gl = __ldg(grad_loss);
  const double cf0 = __ldg(coef);
  // level 0: the Sobel centre that reads this pixel at offset (dy, dx) is (y - dy, x - dx)
  int code0[8];
  {
    const uint8_t* cm = codes + (size_t)b * plane;
    int k = 0;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        if (dy == 0 && dx == 0) continue;
        const int cy = y - dy, cx = x - dx;
        const bool in = cy >= 0 && cy < c.H && cx >= 0 && cx < c.W;
        code0[k++] = in ? (int)__ldg(cm + (in ? (size_t)cy * c.W + cx : 0)) : 0xff;
      }
  }
  if (!(gi > real(0)) || mi == 0 || !(pi >= real(1e-6))) { d_pred[i] = real(0); return; }
  double g = 0.0;
  {
    int acc = 0, k = 0;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        if (dy == 0 && dx == 0) continue;
        const int code = code0[k++];
        const int sx = (code & 3) - 1, sy = ((code >> 2) & 3) - 1;
        const int t = sx * dx * (2 - (dy < 0 ? -dy : dy)) + sy * dy * (2 - (dx < 0 ? -dx : dx));
        acc += code == 0xff ? 0 : t;
      }
    g = (double)acc * cf0;
  }
  for (int s = 1; s < c.S; ++s) {
    const int half = 1 << (s - 1), m = (1 << s) - 1, ry = y & m, rx = x & m;
    if ((ry != half - 1 && ry != half) || (rx != half - 1 && rx != half)) continue;
    const int yy = y >> s, xx = x >> s, Hs = c.Hs[s], Ws = c.Ws[s];
    if (yy >= Hs || xx >= Ws) continue;
    const double cf = __ldg(coef + s);
    if (cf == 0.0) continue;
    const uint8_t* cm = codes + c.off[s] + (size_t)b * Hs * Ws;
    int acc = 0;
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        if (dy == 0 && dx == 0) continue;
        const int cy = yy - dy, cx = xx - dx;
        const bool in = cy >= 0 && cy < Hs && cx >= 0 && cx < Ws;
        const int code = in ? (int)__ldg(cm + (in ? (size_t)cy * Ws + cx : 0)) : 0xff;
        const int sx = (code & 3) - 1, sy = ((code >> 2) & 3) - 1;
        const int t = sx * dx * (2 - (dy < 0 ? -dy : dy)) + sy * dy * (2 - (dx < 0 ? -dx : dx));
        acc += code == 0xff ? 0 : t;
      }
    g += (double)acc * cf;
  }
  const real depth = real(1) / pi;
  d_pred[i] = (real)(-g) * depth * depth * gl;
}

// the eight Sobel centres around pixel (yy, xx) of one level: sum of sign x weight (kNoCode decodes to zero signs)
__device__ __forceinline__ int sobel_taps(const uint8_t* cm, int Hs, int Ws, int yy, int xx) {
  int acc = 0;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      if (dy == 0 && dx == 0) continue;
      const int cy = yy - dy, cx = xx - dx;          // the centre that reads (yy, xx) at offset (dy, dx)
      const bool in = cy >= 0 && cy < Hs && cx >= 0 && cx < Ws;
      const int code = in ? (int)__ldg(cm + (in ? cy * Ws + cx : 0)) : kNoCode;
      const int sx = (code & 3) - 1, sy = ((code >> 2) & 3) - 1;
      acc += sx * dx * (2 - (dy < 0 ? -dy : dy)) + sy * dy * (2 - (dx < 0 ? -dx : dx));
    }
  return acc;
}

// W a multiple of 4: one thread per quad of pixels of a row.  Level 0 in separable form -- per code row rho = y - dy the
// six codes at x - 1 .. x + 4 (one aligned word + two bytes) give A(x) = sx(x - 1) - sx(x + 1) and
// B(x) = sy(x - 1) + 2 sy(x) + sy(x + 1); the pixel's sum is (2 - |dy|) A + dy B over the three rows.  Levels >= 1: each
// D_s pixel the quad feeds is evaluated once (two at level 1, at most one above) and shared by the pixels under it.
template <typename real>
__global__ void __launch_bounds__(kST) sobel_grad4_kernel(SobelCfg c, const real* pred, const real* gt, const uint8_t* mask,
                                                          const uint8_t* codes, const double* coef, const real* grad_loss,
                                                          real* d_pred) {
  const unsigned int wq = (unsigned int)c.W >> 2, per = wq * c.H;
  const unsigned int t = blockIdx.x * kST + threadIdx.x;
  if (t >= per * c.B) return;
  const unsigned int b = t / per, q = t - b * per;
  const int y = (int)(q / wq), x = (int)(q - (unsigned int)y * wq) * 4;
  const size_t plane = (size_t)c.H * c.W, i = (size_t)b * plane + (size_t)y * c.W + x;
  real g4[4], p4[4];
  ld4(gt + i, g4);
  ld4(pred + i, p4);
  const unsigned int mk = mask ? __ldg(reinterpret_cast<const unsigned int*>(mask + i)) : 0x01010101u;
  const real gl = __ldg(grad_loss);
  const uint8_t* cm = codes + (size_t)b * plane;
  int acc[4] = {0, 0, 0, 0};
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy) {
    const int rho = y - dy;
    if (rho < 0 || rho >= c.H) continue;
    const uint8_t* row = cm + (size_t)rho * c.W + x;
    const unsigned int w = __ldg(reinterpret_cast<const unsigned int*>(row));
    int cd[6];
    cd[0] = x > 0 ? (int)__ldg(row - 1) : kNoCode;
    cd[5] = x + 4 < c.W ? (int)__ldg(row + 4) : kNoCode;
#pragma unroll
    for (int k = 0; k < 4; ++k) cd[k + 1] = (int)((w >> (8 * k)) & 0xffu);
    int sx[6], sy[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) { sx[k] = (cd[k] & 3) - 1; sy[k] = ((cd[k] >> 2) & 3) - 1; }
#pragma unroll
    for (int k = 0; k < 4; ++k)
      acc[k] += (2 - (dy < 0 ? -dy : dy)) * (sx[k] - sx[k + 2]) + dy * (sy[k] + 2 * sy[k + 1] + sy[k + 2]);
  }
  bool ok[4];
  bool any = false;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    ok[k] = g4[k] > real(0) && ((mk >> (8 * k)) & 0xffu) != 0 && p4[k] >= real(1e-6);
    any |= ok[k];
  }
  real out[4] = {real(0), real(0), real(0), real(0)};
  if (any) {
    const double cf0 = __ldg(coef);
    double g[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) g[k] = (double)acc[k] * cf0;
    for (int s = 1; s < c.S; ++s) {
      const int half = 1 << (s - 1), m = (1 << s) - 1, ry = y & m;
      if (ry != half - 1 && ry != half) continue;
      const int yy = y >> s, Hs = c.Hs[s], Ws = c.Ws[s];
      if (yy >= Hs) continue;
      const double cf = __ldg(coef + s);
      if (cf == 0.0) continue;
      const uint8_t* cs = codes + c.off[s] + (size_t)b * Hs * Ws;
      if (s == 1) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int xx = (x >> 1) + j;
          if (xx >= Ws) continue;
          const double a = (double)sobel_taps(cs, Hs, Ws, yy, xx) * cf;
          g[2 * j] += a;
          g[2 * j + 1] += a;
        }
      } else if (s == 2) {
        const int xx = x >> 2;
        if (xx >= Ws) continue;
        const double a = (double)sobel_taps(cs, Hs, Ws, yy, xx) * cf;
        g[1] += a;
        g[2] += a;
      } else {
        const int rx = x & m;
        if (rx != half - 4 && rx != half) continue;
        const int xx = (rx == half ? x : x + 3) >> s;
        if (xx >= Ws) continue;
        const double a = (double)sobel_taps(cs, Hs, Ws, yy, xx) * cf;
        if (rx == half) g[0] += a; else g[3] += a;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const real depth = real(1) / p4[k];
      out[k] = ok[k] ? (real)(-g[k]) * depth * depth * gl : real(0);
    }
  }
  st4(d_pred + i, out);
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
    c.nblk[s] = (unsigned int)((n + (size_t)kST * kSP - 1) / ((size_t)kST * kSP));
    off += n;
    if (off >= ((size_t)1 << 31)) return false;
  }
  c.off[kMaxGS] = (unsigned int)off;
  for (int s = S; s < kMaxGS; ++s) c.off[s] = (unsigned int)off;
  return true;
}

// level pyramids of D: the 4 x 4 block kernel for levels 0-2 when the rows hold whole quads (pointers 16-byte aligned:
// cudaMalloc'ed / torch tensors of whole images), the per-pixel kernel for the rest
template <typename real>
static void launch_pyramid(const SobelCfg& c, const real* pred, const real* gt, const uint8_t* mask, real* D, cudaStream_t st) {
  const unsigned int total = c.off[kMaxGS];
  const bool aligned = ((uintptr_t)pred | (uintptr_t)gt | (uintptr_t)D) % 16 == 0 && (uintptr_t)mask % 4 == 0;
  if (c.H % 4 == 0 && c.W % 4 == 0 && aligned) {
    const unsigned int blocks = (unsigned int)c.B * (c.H / 4) * (c.W / 4);
    sobel_pyramid4_kernel<real><<<(blocks + kST - 1) / kST, kST, 0, st>>>(c, pred, gt, mask, D);
    const unsigned int first = c.off[c.S < 3 ? c.S : 3];
    if (total > first) {
      count_launches(1);
      sobel_pyramid_kernel<real><<<(total - first + kST - 1) / kST, kST, 0, st>>>(c, first, pred, gt, mask, D);
    }
  } else {
    sobel_pyramid_kernel<real><<<(total + kST - 1) / kST, kST, 0, st>>>(c, 0u, pred, gt, mask, D);
  }
}

}  // namespace psfm

int psfm_fail(int code, const char* fmt, ...);      // psfm_abi.cu
using namespace psfm;

extern "C" {

size_t psfm_sobel_gradient_code_bytes(int B, int H, int W, int scales) {
  SobelCfg c;
  if (!sobel_cfg(B, H, W, scales, c)) return 0;
  return al256s(c.off[kMaxGS]) + 256;     // sign codes, then kMaxGS doubles: the backward's per-level coefficients
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
  // level-0 quads in the response kernel (PSFM_SOBEL_QUADS=0: the per-pixel form, for comparison)
  static const bool quads_on = [] { const char* e = getenv("PSFM_SOBEL_QUADS"); return !(e && e[0] == '0'); }();
  const bool quad0 = quads_on && W % 4 == 0 && ((uintptr_t)gt_inv | (uintptr_t)workspace) % 16 == 0 &&
                     ((uintptr_t)mask | (uintptr_t)codes) % 4 == 0;
  count_launches(3);
  if (dt == PSFM_F64) {
    typedef double real;
    launch_pyramid<real>(c, (const real*)pred_inv, (const real*)gt_inv, (const uint8_t*)mask, (real*)workspace, st);
    if (quad0)
      sobel_kernel<real, true><<<dim3(c.nblk[0], scales), kST, 0, st>>>(c, (const real*)gt_inv, (const uint8_t*)mask, (const real*)workspace, (uint8_t*)codes, part);
    else
      sobel_kernel<real, false><<<dim3(c.nblk[0], scales), kST, 0, st>>>(c, (const real*)gt_inv, (const uint8_t*)mask, (const real*)workspace, (uint8_t*)codes, part);
    sobel_finish_kernel<real><<<1, kFT, 0, st>>>(c, c.nblk[0], part, (double*)state, (real*)loss_out,
                                                 (double*)((char*)codes + al256s(total)));
  } else {
    typedef float real;
    launch_pyramid<real>(c, (const real*)pred_inv, (const real*)gt_inv, (const uint8_t*)mask, (real*)workspace, st);
    if (quad0)
      sobel_kernel<real, true><<<dim3(c.nblk[0], scales), kST, 0, st>>>(c, (const real*)gt_inv, (const uint8_t*)mask, (const real*)workspace, (uint8_t*)codes, part);
    else
      sobel_kernel<real, false><<<dim3(c.nblk[0], scales), kST, 0, st>>>(c, (const real*)gt_inv, (const uint8_t*)mask, (const real*)workspace, (uint8_t*)codes, part);
    sobel_finish_kernel<real><<<1, kFT, 0, st>>>(c, c.nblk[0], part, (double*)state, (real*)loss_out,
                                                 (double*)((char*)codes + al256s(total)));
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
  const double* coef = (const double*)((const char*)codes + al256s(c.off[kMaxGS]));
  count_launches(1);
  const bool quads = W % 4 == 0 && ((uintptr_t)pred_inv | (uintptr_t)gt_inv | (uintptr_t)d_pred) % 16 == 0 &&
                     ((uintptr_t)mask | (uintptr_t)codes) % 4 == 0;
  const unsigned int qblocks = (unsigned int)(((size_t)B * H * (W / 4) + kST - 1) / kST);
  if (quads && dt == PSFM_F64)
    sobel_grad4_kernel<double><<<qblocks, kST, 0, (cudaStream_t)stream>>>(c, (const double*)pred_inv, (const double*)gt_inv, (const uint8_t*)mask,
                                                                         (const uint8_t*)codes, coef, (const double*)grad_loss, (double*)d_pred);
  else if (quads)
    sobel_grad4_kernel<float><<<qblocks, kST, 0, (cudaStream_t)stream>>>(c, (const float*)pred_inv, (const float*)gt_inv, (const uint8_t*)mask,
                                                                        (const uint8_t*)codes, coef, (const float*)grad_loss, (float*)d_pred);
  else if (dt == PSFM_F64)
    sobel_grad_kernel<double><<<grid, kST, 0, (cudaStream_t)stream>>>(c, (const double*)pred_inv, (const double*)gt_inv, (const uint8_t*)mask,
                                                                      (const uint8_t*)codes, coef, (const double*)grad_loss, (double*)d_pred);
  else
    sobel_grad_kernel<float><<<grid, kST, 0, (cudaStream_t)stream>>>(c, (const float*)pred_inv, (const float*)gt_inv, (const uint8_t*)mask,
                                                                     (const uint8_t*)codes, coef, (const float*)grad_loss, (float*)d_pred);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return psfm_fail(PSFM_ERR_CUDA, "psfm_sobel_gradient_loss_backward: %s", cudaGetErrorString(e));
  return PSFM_OK;
}

}  // extern "C"
