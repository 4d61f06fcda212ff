// loss_tile.cuh -- shared-memory tile helpers used by the loss kernels (tile and streaming variants).
#pragma once
#include <math_constants.h>
#include "psfm_device.cuh"
#include "loss_params.h"

namespace psfm {

template <typename real> __device__ __forceinline__ real real_inf();
template <> __device__ __forceinline__ float real_inf<float>() { return CUDART_INF_F; }
template <> __device__ __forceinline__ double real_inf<double>() { return CUDART_INF; }

// ------------------------------------------------------------------------------------------
// shared-memory tile fills.  Tiles are addressed in padded coordinates: cell (ly, lx) of a tile
// with halo h holds the value at image position reflect(v0 - h + ly), reflect(u0 - h + lx).
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t, int SH, int SW>
__device__ __forceinline__ void fill_image_tile(real (*dst)[SH][SW], const img_t* img, size_t plane, int H, int W,
                                                int v_org, int u_org) {
  for (int i = threadIdx.x; i < SH * SW; i += blockDim.x) {
    const int ly = i / SW, lx = i - ly * SW;
    const int gv = reflect1(v_org + ly, H), gu = reflect1(u_org + lx, W);
    const size_t o = (size_t)gv * W + gu;
#pragma unroll
    for (int c = 0; c < 3; ++c) dst[c][ly][lx] = ldimg<real>(img + c * plane, o);
  }
}

template <typename real>
__device__ __forceinline__ real load_inv_depth(const LossParams& P, const real* invd, size_t o, real& chain) {
  real d = invd[o];
  chain = real(1);
  if (P.premap) d = premap_fork<real>(d, (real)P.premap_lo, (real)P.premap_hi, chain);
  return d;
}

template <typename real, typename img_t, int SH, int SW>
__device__ __forceinline__ void fill_warped_tile(real (*dst)[SH][SW], const LossParams& P, const CamPair<real>& cam,
                                                 const real* T, const img_t* ctx, const real* invd, size_t plane,
                                                 int H, int W, int v_org, int u_org) {
  for (int i = threadIdx.x; i < SH * SW; i += blockDim.x) {
    const int ly = i / SW, lx = i - ly * SW;
    const int gv = reflect1(v_org + ly, H), gu = reflect1(u_org + lx, W);
    real chain;
    const real d = load_inv_depth<real>(P, invd, (size_t)gv * W + gu, chain);
    PixGeom<real> g;
    pixel_geometry<real>(cam, T, (real)gu, (real)gv, d, H, W, g);
    real w[3];
    bilinear3<real, img_t, false>(ctx, plane, H, W, P.padding, g.ix, g.iy, w, nullptr, nullptr);
    dst[0][ly][lx] = w[0]; dst[1][ly][lx] = w[1]; dst[2][ly][lx] = w[2];
  }
}

// window means of one channel around tile cell (ly, lx) (centre)
template <typename real, int SH, int SW>
__device__ __forceinline__ void window_means(const real (*X)[SW], const real (*Y)[SW], int ly, int lx, real& mx,
                                             real& my, real& exx, real& eyy, real& exy) {
  real sx = 0, sy = 0, sxx = 0, syy = 0, sxy = 0;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      const real x = X[ly + dy][lx + dx], y = Y[ly + dy][lx + dx];
      sx += x; sy += y; sxx += x * x; syy += y * y; sxy += x * y;
    }
  const real r9 = real(1) / real(9);
  mx = sx * r9; my = sy * r9; exx = sxx * r9; eyy = syy * r9; exy = sxy * r9;
}

// photometric value(s) of one pixel: calc_photometric_loss (losses/...:218-245).
// returns the number of channels written to out (1 when ssim_w > 0, else 3)
template <typename real, int SH, int SW>
__device__ __forceinline__ int photometric_pixel(const real (*X)[SH][SW], const real (*Y)[SH][SW], int ly, int lx,
                                                 real ssim_w, real C1, real C2, real* out) {
  if (ssim_w > real(0)) {
    real ssum = 0, lsum = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      real mx, my, exx, eyy, exy;
      window_means<real, SH, SW>(X[c], Y[c], ly, lx, mx, my, exx, eyy, exy);
      const real ssim = ssim_from_means<real>(mx, my, exx, eyy, exy, C1, C2);
      ssum += rmin(rmax((real(1) - ssim) / real(2), real(0)), real(1));
      lsum += rabs(X[c][ly][lx] - Y[c][ly][lx]);
    }
    out[0] = ssim_w * (ssum / real(3)) + (real(1) - ssim_w) * (lsum / real(3));
    return 1;
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) out[c] = rabs(X[c][ly][lx] - Y[c][ly][lx]);
  return 3;
}

}  // namespace psfm
