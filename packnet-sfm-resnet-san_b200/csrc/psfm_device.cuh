// psfm_device.cuh -- device-side building blocks shared by the photometric-loss kernels.
//
// Arithmetic follows SURVEY.md Appendix A, which restates the reference
// (geometry/camera.py:73-190, geometry/camera_utils.py:16-59, geometry/pose.py:80-86,
//  utils/depth.py:103-120, losses/multiview_photometric_loss.py:15-54,199-267) and the
// semantics of torch's grid_sampler_2d (bilinear, align_corners=True, zeros|border).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/psfm.h"

namespace psfm {

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
template <typename real> __device__ __forceinline__ real rmax(real a, real b) { return a > b ? a : b; }
template <typename real> __device__ __forceinline__ real rmin(real a, real b) { return a < b ? a : b; }
__device__ __forceinline__ float rfloor(float x) { return floorf(x); }
__device__ __forceinline__ double rfloor(double x) { return floor(x); }
__device__ __forceinline__ float rabs(float x) { return fabsf(x); }
__device__ __forceinline__ double rabs(double x) { return fabs(x); }
__device__ __forceinline__ float rexp(float x) { return expf(x); }
__device__ __forceinline__ double rexp(double x) { return exp(x); }
template <typename real> __device__ __forceinline__ real rsign(real x) {
  return x > real(0) ? real(1) : (x < real(0) ? real(-1) : real(0));
}

// image element load with conversion to the arithmetic type (read-only path)
template <typename real, typename I> __device__ __forceinline__ real ldimg(const float* p, I i) { return (real)__ldg(p + i); }
template <typename real, typename I> __device__ __forceinline__ real ldimg(const double* p, I i) { return (real)__ldg(p + i); }
template <typename real, typename I> __device__ __forceinline__ real ldimg(const __nv_bfloat16* p, I i) {
  return (real)__bfloat162float(__ldg(p + i));
}

// bytes: value = byte / 255, correctly rounded (= torchvision to_tensor).  q = x * fl(1/255) is off by an ulp for
// 126 of the 256 bytes; one residual correction makes all 256 exact (checked exhaustively).
template <typename real, typename I> __device__ __forceinline__ real ldimg(const uint8_t* p, I i) {
  const float x = (float)__ldg(p + i), c = 1.0f / 255.0f;
  const float q = x * c;
  return (real)fmaf(fmaf(-q, 255.0f, x), c, q);
}
// mask element: the value itself (a byte mask holds 0 / 1)
template <typename real, typename T, typename I> __device__ __forceinline__ real ldmask(const T* p, I i) { return ldimg<real>(p, i); }
template <typename real, typename I> __device__ __forceinline__ real ldmask(const uint8_t* p, I i) { return (real)__ldg(p + i); }

// ReflectionPad2d(1) source index: -1 -> 1, n -> n-2 (losses/multiview_photometric_loss.py:34).
// Positions further out (only touched by windows that are never evaluated) are clamped.
__device__ __forceinline__ int reflect1(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  if (i < 0) i = 0;
  if (i >= n) i = n - 1;
  return i;
}

// ------------------------------------------------------------------------------------------
// camera pair of one (batch item, scale): lifting matrix of the target camera, intrinsics of
// the context camera, both rescaled to the scale's resolution.
// ------------------------------------------------------------------------------------------
template <typename real> struct CamPair {
  real ki[9];   // reference `Kinv` (camera.py:73-81): K with 4 entries replaced
  real rk[9];   // scaled ref_K
};

// Camera.scaled(sf) (camera.py:84-108) -> scale_intrinsics (camera_utils.py:16-22); the same
// factor is applied to x and y, and sf == 1 leaves K untouched.
template <typename real> __device__ __forceinline__ void scale_K(real* k, real sf) {
  if (sf != real(1)) {
    k[0] *= sf;
    k[4] *= sf;
    k[2] = (k[2] + real(0.5)) * sf - real(0.5);
    k[5] = (k[5] + real(0.5)) * sf - real(0.5);
  }
}

template <typename real>
__device__ __forceinline__ void setup_cam(CamPair<real>& c, const real* K, const real* refK, int b, real sf) {
  real k[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) { k[i] = K[b * 9 + i]; c.rk[i] = refK[b * 9 + i]; }
  scale_K(k, sf);
  scale_K(c.rk, sf);
#pragma unroll
  for (int i = 0; i < 9; ++i) c.ki[i] = k[i];
  c.ki[0] = real(1) / k[0];
  c.ki[4] = real(1) / k[4];
  c.ki[2] = real(-1) * k[2] / k[0];
  c.ki[5] = real(-1) * k[5] / k[4];
}

// fork-only sigmoid -> inverse-depth pre-map (losses/multiview_photometric_loss.py:362-369,
// utils/post_process_depth.py:102-106).  Returns d and d(d)/d(sigma).
template <typename real>
__device__ __forceinline__ real premap_fork(real sig, real lo, real hi, real& dd_dsig) {
  real inv = lo + (hi - lo) * sig;
  real a = inv + real(1e-8);
  real depth = real(1) / a;
  real bq = depth + real(1e-8);
  real d = real(1) / bq;
  dd_dsig = (real(1) / (bq * bq)) * (real(1) / (a * a)) * (hi - lo);
  return d;
}

// ------------------------------------------------------------------------------------------
// geometry of one pixel: back-project, rigid transform, project, un-normalise.
// ------------------------------------------------------------------------------------------
template <typename real> struct PixGeom {
  real D;          // depth = 1 / max(d, 1e-6)
  real ray[3];     // Kinv @ (u, v, 1)
  real Xc[3];      // ray * D
  real p0, p1, p2; // ref_K @ (R Xc + t)
  real Z;          // max(p2, 1e-5)
  real ix, iy;     // sampling position in pixels (before border clipping)
};

template <typename real>
__device__ __forceinline__ void pixel_geometry_depth(const CamPair<real>& c, const real* T /*[12] rows of 3x4*/,
                                                     real u, real v, real depth, int H, int W, PixGeom<real>& g) {
  g.D = depth;
  g.ray[0] = c.ki[0] * u + c.ki[1] * v + c.ki[2];                       // camera.py:135
  g.ray[1] = c.ki[3] * u + c.ki[4] * v + c.ki[5];
  g.ray[2] = c.ki[6] * u + c.ki[7] * v + c.ki[8];
  g.Xc[0] = g.ray[0] * g.D; g.Xc[1] = g.ray[1] * g.D; g.Xc[2] = g.ray[2] * g.D;   // camera.py:137
  real X0 = T[0] * g.Xc[0] + T[1] * g.Xc[1] + T[2] * g.Xc[2] + T[3];    // pose.py:84-85
  real X1 = T[4] * g.Xc[0] + T[5] * g.Xc[1] + T[6] * g.Xc[2] + T[7];
  real X2 = T[8] * g.Xc[0] + T[9] * g.Xc[1] + T[10] * g.Xc[2] + T[11];
  g.p0 = c.rk[0] * X0 + c.rk[1] * X1 + c.rk[2] * X2;                    // camera.py:172
  g.p1 = c.rk[3] * X0 + c.rk[4] * X1 + c.rk[5] * X2;
  g.p2 = c.rk[6] * X0 + c.rk[7] * X1 + c.rk[8] * X2;
  g.Z = rmax(g.p2, real(1e-5));                                         // camera.py:179
  real xn = real(2) * (g.p0 / g.Z) / real(W - 1) - real(1);             // camera.py:180-181
  real yn = real(2) * (g.p1 / g.Z) / real(H - 1) - real(1);
  g.ix = ((xn + real(1)) / real(2)) * real(W - 1);                      // grid_sampler unnormalize, align_corners
  g.iy = ((yn + real(1)) / real(2)) * real(H - 1);
}

// same, starting from an inverse depth: depth = 1 / max(d, 1e-6)  (utils/depth.py:120)
template <typename real>
__device__ __forceinline__ void pixel_geometry(const CamPair<real>& c, const real* T, real u, real v, real d, int H,
                                               int W, PixGeom<real>& g) {
  pixel_geometry_depth<real>(c, T, u, v, real(1) / rmax(d, real(1e-6)), H, W, g);
}

// ------------------------------------------------------------------------------------------
// chain rule from the sampling position back to the inverse depth and the pose entries
// (adjoint of pixel_geometry).  gix, giy: d loss / d ix, d iy.  Returns d loss / d d and adds
// this pixel's contribution to pose_acc[12] (rows of the 3x4 [R|t]).
// ------------------------------------------------------------------------------------------
template <typename real>
__device__ __forceinline__ real geometry_adjoint_depth(const CamPair<real>& c, const real* T, const PixGeom<real>& g,
                                                       real gix, real giy, real* pose_acc) {
  // ix = p0 / Z (the normalise / un-normalise pair cancels), Z = max(p2, 1e-5)
  const real iz = real(1) / g.Z;
  const real dp0 = gix * iz, dp1 = giy * iz;
  const real dp2 = (g.p2 >= real(1e-5)) ? -(gix * g.p0 + giy * g.p1) * iz * iz : real(0);
  real dX[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) dX[i] = c.rk[i] * dp0 + c.rk[3 + i] * dp1 + c.rk[6 + i] * dp2;   // ref_K^T dp
  real dXc[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) dXc[k] = T[k] * dX[0] + T[4 + k] * dX[1] + T[8 + k] * dX[2];     // R^T dX
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    pose_acc[i * 4 + 0] += dX[i] * g.Xc[0];
    pose_acc[i * 4 + 1] += dX[i] * g.Xc[1];
    pose_acc[i * 4 + 2] += dX[i] * g.Xc[2];
    pose_acc[i * 4 + 3] += dX[i];
  }
  return g.ray[0] * dXc[0] + g.ray[1] * dXc[1] + g.ray[2] * dXc[2];                              // d loss / d depth
}

template <typename real>
__device__ __forceinline__ real geometry_adjoint(const CamPair<real>& c, const real* T, const PixGeom<real>& g, real d,
                                                 real gix, real giy, real* pose_acc) {
  const real dD = geometry_adjoint_depth<real>(c, T, g, gix, giy, pose_acc);
  return (d >= real(1e-6)) ? -(g.D * g.D) * dD : real(0);                                        // D = 1/max(d,1e-6)
}

// ------------------------------------------------------------------------------------------
// bilinear sample of a 3-plane image (grid_sample, bilinear, align_corners=True)
//   PAD == 0 zeros : taps outside the image contribute 0 to value and gradient
//   PAD == 1 border: coordinate clipped to [0, size-1]; zero coordinate-gradient where clipped
// `img` points at plane 0 of the batch item, planes are `plane` elements apart.
// When GRAD: gx[c] = d out[c] / d ix, gy[c] = d out[c] / d iy (already zeroed where clipped).
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t, bool GRAD>
__device__ __forceinline__ void bilinear3(const img_t* __restrict__ img, size_t plane, int H, int W, int pad,
                                          real ix, real iy, real* out, real* gx, real* gy) {
  real mx = real(1), my = real(1);
  if (pad == PSFM_PAD_BORDER) {
    if (ix <= real(0)) { ix = real(0); mx = real(0); }
    else if (ix >= real(W - 1)) { ix = real(W - 1); mx = real(0); }
    if (iy <= real(0)) { iy = real(0); my = real(0); }
    else if (iy >= real(H - 1)) { iy = real(H - 1); my = real(0); }
  }
  real fx = rfloor(ix), fy = rfloor(iy);
  real wx1 = ix - fx, wx0 = (fx + real(1)) - ix;
  real wy1 = iy - fy, wy0 = (fy + real(1)) - iy;
  // saturating conversions; anything far outside the image fails the bounds tests below
  fx = rmax(rmin(fx, real(1 << 30)), real(-(1 << 30)));
  fy = rmax(rmin(fy, real(1 << 30)), real(-(1 << 30)));
  int x0 = (int)fx, y0 = (int)fy, x1 = x0 + 1, y1 = y0 + 1;
  bool bx0 = (x0 >= 0) & (x0 < W), bx1 = (x1 >= 0) & (x1 < W);
  bool by0 = (y0 >= 0) & (y0 < H), by1 = (y1 >= 0) & (y1 < H);
  size_t o00 = (size_t)y0 * W + x0;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const img_t* p = img + c * plane;
    real v00 = (bx0 & by0) ? ldimg<real>(p, o00) : real(0);
    real v01 = (bx1 & by0) ? ldimg<real>(p, o00 + 1) : real(0);
    real v10 = (bx0 & by1) ? ldimg<real>(p, o00 + W) : real(0);
    real v11 = (bx1 & by1) ? ldimg<real>(p, o00 + W + 1) : real(0);
    out[c] = v00 * (wx0 * wy0) + v01 * (wx1 * wy0) + v10 * (wx0 * wy1) + v11 * (wx1 * wy1);
    if (GRAD) {
      gx[c] = mx * ((v01 - v00) * wy0 + (v11 - v10) * wy1);
      gy[c] = my * ((v10 - v00) * wx0 + (v11 - v01) * wx1);
    }
  }
}

// ------------------------------------------------------------------------------------------
// SSIM (3x3, means) from the five window means of one channel.
// ------------------------------------------------------------------------------------------
template <typename real>
__device__ __forceinline__ real ssim_from_means(real mx, real my, real exx, real eyy, real exy, real C1, real C2) {
  real sx = exx - mx * mx, sy = eyy - my * my, sxy = exy - mx * my;
  real n = (real(2) * mx * my + C1) * (real(2) * sxy + C2);
  real d = (mx * mx + my * my + C1) * (sx + sy + C2);
  return n / d;
}

// SSIM distance clamp((1-ssim)/2, 0, 1) and its partial derivatives w.r.t. the window means
// (mu_x, E[x^2], E[xy]) of the first argument.
template <typename real>
__device__ __forceinline__ real ssim_dist_grad(real mx, real my, real exx, real eyy, real exy, real C1, real C2,
                                               real& d_mx, real& d_exx, real& d_exy) {
  real sx = exx - mx * mx, sy = eyy - my * my, sxy = exy - mx * my;
  real a1 = real(2) * mx * my + C1, v1 = real(2) * sxy + C2;
  real b1 = mx * mx + my * my + C1, v2 = sx + sy + C2;
  real inv_d = real(1) / (b1 * v2);
  real ssim = (a1 * v1) * inv_d;
  real dist = (real(1) - ssim) / real(2);
  real act = (dist >= real(0) && dist <= real(1)) ? real(-0.5) : real(0);
  d_mx = act * (real(2) * my * (v1 - a1) * inv_d - ssim * real(2) * mx * (v2 - b1) * inv_d);
  d_exx = act * (-ssim * b1 * inv_d);
  d_exy = act * (real(2) * a1 * inv_d);
  return rmin(rmax(dist, real(0)), real(1));
}

// ------------------------------------------------------------------------------------------
// deterministic block reduction of NV values per thread; result valid in thread 0.
// `scratch` must hold NV * (blockDim.x / 32) reals.  Fixed shuffle tree + fixed warp order.
// ------------------------------------------------------------------------------------------
template <typename real, int NV>
__device__ __forceinline__ void block_reduce(real* vals, real* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    real v = vals[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) scratch[warp * NV + i] = v;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    real v = real(0);
    for (int w = 0; w < nw; ++w) v += scratch[w * NV + threadIdx.x];
    scratch[threadIdx.x] = v;     // warp 0 slot i: written by thread i only after it read it
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NV; ++i) vals[i] = scratch[i];
  __syncthreads();
}

}  // namespace psfm
