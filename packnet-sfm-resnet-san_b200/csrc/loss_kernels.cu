// loss_kernels.cu -- sm_100a kernels of the multi-view photometric loss (split path):
//   mean_kernel      per-image mean of the inverse-depth input (smoothness normalisation)
//   loss_fwd_kernel  per scale: view synthesis of every context + SSIM/L1 photometric terms +
//                    identity (automask) terms + per-pixel min/mean + edge-aware smoothness,
//                    block partials in fixed slots
//   loss_finish      fixed-order reduction of the partials -> {loss, photometric, smoothness}
//   loss_bwd_kernel  per scale: recomputes the warp, SSIM adjoint, d/d inv_depth per pixel and
//                    per-block partials of d/d pose
//   pose_finish      fixed-order reduction of the pose partials -> d pose.mat
// Reference semantics: SURVEY.md Appendix A (losses/multiview_photometric_loss.py:199-410,
// geometry/camera.py:111-190, geometry/camera_utils.py:27-59, utils/depth.py:146-198).
#include <cuda_runtime.h>
#include <math_constants.h>
#include "psfm_device.cuh"
#include "loss_params.h"
#include "loss_tile.cuh"
#include "profile.h"

namespace psfm {

// ------------------------------------------------------------------------------------------
// per-image mean of the raw inverse-depth input: inv_depths_normalize (utils/depth.py:146-162)
// grid (B, n), one CTA per image; deterministic strided sum + fixed tree.
// ------------------------------------------------------------------------------------------
template <typename real>
__global__ void __launch_bounds__(1024) mean_kernel(const LossParams P) {
  const int b = blockIdx.x, s = blockIdx.y;
  const ScaleDesc& S = P.sc[s];
  const size_t N = (size_t)S.H * S.W;
  const real* d = (const real*)S.invd + (size_t)b * N;
  real acc = real(0);
  for (size_t i = threadIdx.x; i < N; i += blockDim.x) acc += d[i];
  __shared__ real scratch[32];
  real v[1] = {acc};
  block_reduce<real, 1>(v, scratch);
  if (threadIdx.x == 0) ((real*)P.means)[s * P.B + b] = v[0] / real(N);
}

// ------------------------------------------------------------------------------------------
// forward kernel, one launch per scale.  grid (tiles_x, tiles_y, B), kThreads threads, one
// output pixel per thread.  stats_mode: only accumulate (sum, sum^2) of every candidate group
// for the clip_loss threshold (losses/...:249-253).
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t>
__global__ void __launch_bounds__(kThreads) loss_fwd_kernel(const LossParams P, const int s, const int stats_mode) {
  constexpr int SH = kTileH + 2, SW = kTileW + 2;
  static_assert(kTileW * kTileH == kThreads, "one pixel per thread");
  const ScaleDesc& S = P.sc[s];
  const int H = S.H, W = S.W, b = blockIdx.z;
  const int u0 = blockIdx.x * kTileW, v0 = blockIdx.y * kTileH;
  const int tx = threadIdx.x % kTileW, ty = threadIdx.x / kTileW;
  const int u = u0 + tx, v = v0 + ty;
  const bool valid = (u < W) && (v < H);
  const size_t plane = (size_t)H * W;
  const int tile = blockIdx.y * gridDim.x + blockIdx.x, ntiles = gridDim.x * gridDim.y;

  __shared__ real Ys[3][SH][SW];
  __shared__ real Xs[3][SH][SW];
  __shared__ CamPair<real> cam;
  __shared__ real Ts[12];
  __shared__ real scratch[4 * (kThreads / 32)];

  const img_t* image = (const img_t*)S.image + (size_t)b * 3 * plane;
  const real* invd = (const real*)S.invd + (size_t)b * plane;
  if (threadIdx.x == 0) setup_cam<real>(cam, (const real*)P.K, (const real*)P.refK, b, (real)S.sf);
  fill_image_tile<real, img_t, SH, SW>(Ys, image, plane, H, W, v0 - 1, u0 - 1);

  const real ssim_w = (real)P.ssim_w, C1 = (real)P.C1, C2 = (real)P.C2;
  const int nc = P.ssim_w > 0.0 ? 1 : 3;
  const bool clip = P.clip_loss > 0.0 && !stats_mode;
  const real* thr = (const real*)P.clip_thr + (size_t)s * P.ncand;
  real m = real(1);
  if (S.mask != nullptr && valid) m = ldimg<real>((const img_t*)S.mask + (size_t)b * plane, (size_t)v * W + u);

  real best = real_inf<real>(), acc = real(0);
  int sel = 0, k = 0;      // k: candidate group index
  const int ngroups = P.automask ? 2 * P.C : P.C;

  for (int j = 0; j < P.C; ++j) {
    const img_t* ctx = (const img_t*)S.ctx[j] + (size_t)b * 3 * plane;
    for (int pass = 0; pass < (P.automask ? 2 : 1); ++pass, ++k) {
      __syncthreads();   // previous readers of Xs / Ts done; cam + Ys visible on first trip
      if (pass == 0) {
        if (threadIdx.x < 12) {
          const int r = threadIdx.x / 4, c = threadIdx.x % 4;
          Ts[threadIdx.x] = ((const real*)P.pose[j])[(size_t)b * 16 + r * 4 + c];
        }
        __syncthreads();
        fill_warped_tile<real, img_t, SH, SW>(Xs, P, cam, Ts, ctx, invd, plane, H, W, v0 - 1, u0 - 1);
      } else {
        fill_image_tile<real, img_t, SH, SW>(Xs, ctx, plane, H, W, v0 - 1, u0 - 1);
      }
      __syncthreads();
      real val[3];
      photometric_pixel<real, SH, SW>(Xs, Ys, ty + 1, tx + 1, ssim_w, C1, C2, val);
      if (stats_mode) {
        real st[2] = {real(0), real(0)};
        if (valid)
          for (int c = 0; c < nc; ++c) { st[0] += val[c]; st[1] += val[c] * val[c]; }
        block_reduce<real, 2>(st, scratch);
        if (threadIdx.x == 0) {
          double* o = P.clip_part + S.clip_off + ((((size_t)b * ntiles) + tile) * ngroups + k) * 2;
          o[0] = (double)st[0]; o[1] = (double)st[1];
        }
        continue;
      }
      for (int c = 0; c < nc; ++c) {
        real x = val[c];
        if (clip) x = rmin(x, thr[k]);
        x *= m;
        acc += x;
        if (x < best) { best = x; sel = k * nc + c; }
      }
    }
  }
  if (stats_mode) return;

  real sums[kLossVals] = {real(0), real(0), real(0)};
  if (valid) {
    S.sel[(size_t)b * plane + (size_t)v * W + u] = (uint8_t)sel;
    sums[0] = (P.reduce_op == PSFM_REDUCE_MIN) ? best : acc / real(ngroups * nc);
    if (P.smooth_w > 0.0) {
      // calc_smoothness (utils/depth.py:165-198): mean-normalised inverse depth, forward differences
      const real mean = rmax(((const real*)P.means)[s * P.B + b], real(1e-6));
      const size_t o = (size_t)v * W + u;
      const real d0 = invd[o] / mean;
      const int ly = ty + 1, lx = tx + 1;
      if (u + 1 < W) {
        const real gx = d0 - invd[o + 1] / mean;
        real ig = 0;
#pragma unroll
        for (int c = 0; c < 3; ++c) ig += rabs(Ys[c][ly][lx] - Ys[c][ly][lx + 1]);
        sums[1] = rabs(gx * rexp(-(ig / real(3))));
      }
      if (v + 1 < H) {
        const real gy = d0 - invd[o + W] / mean;
        real ig = 0;
#pragma unroll
        for (int c = 0; c < 3; ++c) ig += rabs(Ys[c][ly][lx] - Ys[c][ly + 1][lx]);
        sums[2] = rabs(gy * rexp(-(ig / real(3))));
      }
    }
  }
  __syncthreads();
  block_reduce<real, kLossVals>(sums, scratch);
  if (threadIdx.x == 0) {
    double* o = P.loss_part + S.part_off + ((size_t)b * ntiles + tile) * kLossVals;
#pragma unroll
    for (int i = 0; i < kLossVals; ++i) o[i] = (double)sums[i];
  }
}

// ------------------------------------------------------------------------------------------
// fixed-order sum of `count` doubles by one CTA (strided per-thread sums + fixed tree)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double cta_sum(const double* p, size_t count, size_t stride, double* sh) {
  double a = 0.0;
  for (size_t i = threadIdx.x; i < count; i += blockDim.x) a += p[i * stride];
  sh[threadIdx.x] = a;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  const double r = sh[0];
  __syncthreads();
  return r;
}

// clip thresholds: mean + clip_loss * std (unbiased) of every candidate group (losses/...:249-253)
template <typename real>
__global__ void __launch_bounds__(256) clip_finish_kernel(const LossParams P) {
  __shared__ double sh[256];
  const int ngroups = P.automask ? 2 * P.C : P.C;
  const int nc = P.ssim_w > 0.0 ? 1 : 3;
  for (int s = 0; s < P.n; ++s) {
    const ScaleDesc& S = P.sc[s];
    const size_t nblk = (size_t)P.B * S.tiles_x * S.tiles_y;
    const double N = (double)P.B * S.H * S.W * nc;
    for (int k = 0; k < ngroups; ++k) {
      const double* base = P.clip_part + S.clip_off + (size_t)k * 2;
      const double s1 = cta_sum(base, nblk, (size_t)ngroups * 2, sh);
      const double s2 = cta_sum(base + 1, nblk, (size_t)ngroups * 2, sh);
      if (threadIdx.x == 0) {
        const double mean = s1 / N;
        double var = (s2 - s1 * s1 / N) / (N - 1.0);
        if (var < 0.0) var = 0.0;
        ((real*)P.clip_thr)[s * P.ncand + k] = (real)((real)mean + (real)P.clip_loss * (real)sqrt(var));
      }
    }
  }
}

// loss finish: {loss, photometric_loss, smoothness_loss} and the per-image smoothness sums
// the backward needs.  reduce_photometric_loss (:269-297), calc_smoothness_loss (:301-327).
template <typename real>
__global__ void __launch_bounds__(256) loss_finish_kernel(const LossParams P) {
  __shared__ double sh[256];
  double photo = 0.0, smooth = 0.0;
  for (int s = 0; s < P.n; ++s) {
    const ScaleDesc& S = P.sc[s];
    const size_t ntiles = (size_t)S.tiles_x * S.tiles_y;
    double ph = 0.0;
    for (int b = 0; b < P.B; ++b) {
      const double* base = P.loss_part + S.part_off + (size_t)b * ntiles * kLossVals;
      ph += cta_sum(base, ntiles, kLossVals, sh);
      const double sx = cta_sum(base + 1, ntiles, kLossVals, sh);
      const double sy = cta_sum(base + 2, ntiles, kLossVals, sh);
      smooth += sx * S.coef_x + sy * S.coef_y;
      if (threadIdx.x == 0) {
        ((real*)P.smooth_img)[(s * P.B + b) * 2 + 0] = (real)sx;
        ((real*)P.smooth_img)[(s * P.B + b) * 2 + 1] = (real)sy;
      }
    }
    photo += ph * S.inv_count;
  }
  if (threadIdx.x == 0) {
    real* out = (real*)P.loss_out;
    out[0] = (real)((real)photo + (real)smooth);
    out[1] = (real)photo;
    out[2] = (real)smooth;
    *(real*)P.one = real(1);
  }
}

// pose finish: d pose[j].mat[b] rows 0..2 = sum over scales and tiles of the block partials.
// grid (B, C), 32 threads: thread t < 12 owns one matrix entry, fixed summation order.
template <typename real>
__global__ void __launch_bounds__(32) pose_finish_kernel(const LossParams P, const int scale_by_grad) {
  const int b = blockIdx.x, j = blockIdx.y, t = threadIdx.x;
  real* out = (real*)P.d_pose[j] + (size_t)b * 16;
  if (t < 12) {
    double a = 0.0;
    for (int s = 0; s < P.n; ++s) {
      const ScaleDesc& S = P.sc[s];
      const size_t ntiles = (size_t)S.tiles_x * S.tiles_y;
      const real* base = (const real*)P.pose_part + S.pose_off + (((size_t)b * ntiles) * P.C + j) * 12 + t;
      for (size_t i = 0; i < ntiles; ++i) a += (double)base[i * P.C * 12];
    }
    (void)scale_by_grad;
    out[(t / 4) * 4 + (t % 4)] = (real)a;
  } else if (t < 16) {
    out[12 + (t - 12)] = real(0);
  }
}

// ------------------------------------------------------------------------------------------
// backward kernel, one launch per scale.  grid (tiles_x, tiles_y, B).  Recomputes the warp on
// the tile + 2-pixel halo, evaluates the SSIM adjoint on the tile + 1-pixel halo for the
// pixels whose arg-min is this context (io->sel), gathers it on the tile, and pushes the
// result through the bilinear sampler and the projection.
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t>
__global__ void __launch_bounds__(kThreads) loss_bwd_kernel(const LossParams P, const int s) {
  constexpr int SH2 = kTileH + 4, SW2 = kTileW + 4, SH1 = kTileH + 2, SW1 = kTileW + 2;
  const ScaleDesc& S = P.sc[s];
  const int H = S.H, W = S.W, b = blockIdx.z;
  const int u0 = blockIdx.x * kTileW, v0 = blockIdx.y * kTileH;
  const int tx = threadIdx.x % kTileW, ty = threadIdx.x / kTileW;
  const int u = u0 + tx, v = v0 + ty;
  const bool valid = (u < W) && (v < H);
  const size_t plane = (size_t)H * W;
  const int tile = blockIdx.y * gridDim.x + blockIdx.x, ntiles = gridDim.x * gridDim.y;

  __shared__ real Ys[3][SH2][SW2];
  __shared__ real Xs[3][SH2][SW2];
  __shared__ real Fs[3][SH1][SW1];
  __shared__ real Gs[SH1][SW1];
  __shared__ uint8_t Sel[SH1][SW1];
  __shared__ CamPair<real> cam;
  __shared__ real Ts[12];
  __shared__ real scratch[12 * (kThreads / 32)];

  const img_t* image = (const img_t*)S.image + (size_t)b * 3 * plane;
  const real* invd = (const real*)S.invd + (size_t)b * plane;
  const real g_up = *(const real*)P.grad_loss;
  if (threadIdx.x == 0) setup_cam<real>(cam, (const real*)P.K, (const real*)P.refK, b, (real)S.sf);
  fill_image_tile<real, img_t, SH2, SW2>(Ys, image, plane, H, W, v0 - 2, u0 - 2);
  for (int i = threadIdx.x; i < SH1 * SW1; i += blockDim.x) {
    const int ly = i / SW1, lx = i - ly * SW1;
    const int pv = v0 - 1 + ly, pu = u0 - 1 + lx;
    const bool real_p = (pv >= 0) & (pv < H) & (pu >= 0) & (pu < W);
    (&Sel[0][0])[i] = real_p ? S.sel[(size_t)b * plane + (size_t)pv * W + pu] : (uint8_t)255;
  }

  const real ssim_w = (real)P.ssim_w, C1 = (real)P.C1, C2 = (real)P.C2;
  const int nc = P.ssim_w > 0.0 ? 1 : 3;
  const int ngroups = P.automask ? 2 * P.C : P.C;
  const bool clip = P.clip_loss > 0.0;
  const real* thr = (const real*)P.clip_thr + (size_t)s * P.ncand;
  const real base_w = g_up * (real)S.inv_count;
  const real mean_w = real(1) / real(ngroups * nc);
  const img_t* maskp = S.mask ? (const img_t*)S.mask + (size_t)b * plane : nullptr;
  // reflection multiplicities of the 3x3 adjoint at this pixel (ReflectionPad2d(1))
  real mv[3] = {real(1), real(1), real(1)}, mu[3] = {real(1), real(1), real(1)};
  if (v == 1) mv[0] += real(1);
  if (v == H - 2) mv[2] += real(1);
  if (u == 1) mu[0] += real(1);
  if (u == W - 2) mu[2] += real(1);

  real d_acc = real(0);          // d loss / d (raw inverse-depth input) of this pixel
  real dchain = real(1);
  const real d_q = valid ? load_inv_depth<real>(P, invd, (size_t)v * W + u, dchain) : real(1);

  for (int j = 0; j < P.C; ++j) {
    const int k = P.automask ? 2 * j : j;          // candidate group of the warped term
    const img_t* ctx = (const img_t*)S.ctx[j] + (size_t)b * 3 * plane;
    __syncthreads();
    if (threadIdx.x < 12) Ts[threadIdx.x] = ((const real*)P.pose[j])[(size_t)b * 16 + (threadIdx.x / 4) * 4 + threadIdx.x % 4];
    __syncthreads();
    fill_warped_tile<real, img_t, SH2, SW2>(Xs, P, cam, Ts, ctx, invd, plane, H, W, v0 - 2, u0 - 2);
    __syncthreads();

    real dLdx[3] = {real(0), real(0), real(0)};
    if (nc == 1) {
      // weight of the photometric value of every pixel of the tile + 1 halo
      for (int i = threadIdx.x; i < SH1 * SW1; i += blockDim.x) {
        const int ly = i / SW1, lx = i - ly * SW1;
        const uint8_t sl = Sel[ly][lx];
        real w = real(0);
        if (sl != 255) {
          w = (P.reduce_op == PSFM_REDUCE_MIN) ? (sl == k ? real(1) : real(0)) : mean_w;
          if (w != real(0) && maskp) w *= ldimg<real>(maskp, (size_t)(v0 - 1 + ly) * W + (u0 - 1 + lx));
          if (w != real(0) && clip) {
            real val[3];
            photometric_pixel<real, SH2, SW2>(Xs, Ys, ly + 1, lx + 1, ssim_w, C1, C2, val);
            if (val[0] > thr[k]) w = real(0);
          }
        }
        Gs[ly][lx] = w * base_w;
      }
      __syncthreads();
      for (int c = 0; c < 3; ++c) {
        for (int i = threadIdx.x; i < SH1 * SW1; i += blockDim.x) {
          const int ly = i / SW1, lx = i - ly * SW1;
          const real w = Gs[ly][lx] * ssim_w / real(3);
          real fa = real(0), fb = real(0), fc = real(0);
          if (w != real(0)) {
            real mx, my, exx, eyy, exy, dmx, dexx, dexy;
            window_means<real, SH2, SW2>(Xs[c], Ys[c], ly + 1, lx + 1, mx, my, exx, eyy, exy);
            ssim_dist_grad<real>(mx, my, exx, eyy, exy, C1, C2, dmx, dexx, dexy);
            const real w9 = w / real(9);
            fa = w9 * dmx; fb = w9 * real(2) * dexx; fc = w9 * dexy;
          }
          Fs[0][ly][lx] = fa; Fs[1][ly][lx] = fb; Fs[2][ly][lx] = fc;
        }
        __syncthreads();
        if (valid) {
          real sa = 0, sb = 0, sc = 0;
#pragma unroll
          for (int dy = 0; dy < 3; ++dy)
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) {
              const real mlt = mv[dy] * mu[dx];
              sa += mlt * Fs[0][ty + dy][tx + dx];
              sb += mlt * Fs[1][ty + dy][tx + dx];
              sc += mlt * Fs[2][ty + dy][tx + dx];
            }
          const real xq = Xs[c][ty + 2][tx + 2], yq = Ys[c][ty + 2][tx + 2];
          dLdx[c] = sa + xq * sb + yq * sc + Gs[ty + 1][tx + 1] * ((real(1) - ssim_w) / real(3)) * rsign(xq - yq);
        }
        __syncthreads();
      }
    } else if (valid) {
      // L1-only photometric loss: three candidates per group, no spatial coupling
      const uint8_t sl = Sel[ty + 1][tx + 1];
      real m = maskp ? ldimg<real>(maskp, (size_t)v * W + u) : real(1);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const real xq = Xs[c][ty + 2][tx + 2], yq = Ys[c][ty + 2][tx + 2];
        real w = (P.reduce_op == PSFM_REDUCE_MIN) ? (sl == k * 3 + c ? real(1) : real(0)) : mean_w;
        if (clip && rabs(xq - yq) > thr[k]) w = real(0);
        dLdx[c] = w * m * base_w * rsign(xq - yq);
      }
    }

    real pose_acc[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) pose_acc[i] = real(0);
    if (valid && (dLdx[0] != real(0) || dLdx[1] != real(0) || dLdx[2] != real(0))) {
      PixGeom<real> g;
      pixel_geometry<real>(cam, Ts, (real)u, (real)v, d_q, H, W, g);
      real w[3], gx[3], gy[3];
      bilinear3<real, img_t, true>(ctx, plane, H, W, P.padding, g.ix, g.iy, w, gx, gy);
      const real gix = dLdx[0] * gx[0] + dLdx[1] * gx[1] + dLdx[2] * gx[2];
      const real giy = dLdx[0] * gy[0] + dLdx[1] * gy[1] + dLdx[2] * gy[2];
      d_acc += geometry_adjoint<real>(cam, Ts, g, d_q, gix, giy, pose_acc) * dchain;
    }
    block_reduce<real, 12>(pose_acc, scratch);
    if (threadIdx.x < 12) {
      real* o = (real*)P.pose_part + S.pose_off + ((((size_t)b * ntiles) + tile) * P.C + j) * 12;
      o[threadIdx.x] = pose_acc[threadIdx.x];
    }
  }

  if (valid) {
    if (P.smooth_w > 0.0) {
      // adjoint of calc_smoothness_loss; the per-image mean couples all pixels
      const real mraw = ((const real*)P.means)[s * P.B + b];
      const real mean = rmax(mraw, real(1e-6));
      const size_t o = (size_t)v * W + u;
      const int ly = ty + 2, lx = tx + 2;
      const real d0 = invd[o] / mean;
      real q = real(0);
      auto edge = [&](real dn, int ny, int nx) {   // sign(g) * w of the edge between (ly,lx) and (ny,nx)
        real ig = 0;
#pragma unroll
        for (int c = 0; c < 3; ++c) ig += rabs(Ys[c][ly][lx] - Ys[c][ny][nx]);
        return rsign(dn) * rexp(-(ig / real(3)));
      };
      if (u + 1 < W) q += (real)S.coef_x * edge(d0 - invd[o + 1] / mean, ly, lx + 1);
      if (u >= 1) q -= (real)S.coef_x * edge(invd[o - 1] / mean - d0, ly, lx - 1);
      if (v + 1 < H) q += (real)S.coef_y * edge(d0 - invd[o + W] / mean, ly + 1, lx);
      if (v >= 1) q -= (real)S.coef_y * edge(invd[o - W] / mean - d0, ly - 1, lx);
      real gs = q / mean;
      if (mraw >= real(1e-6)) {
        const real* si = (const real*)P.smooth_img + (s * P.B + b) * 2;
        gs -= ((real)S.coef_x * si[0] + (real)S.coef_y * si[1]) / (real(plane) * mean);
      }
      d_acc += g_up * gs;
    }
    ((real*)S.d_invd)[(size_t)b * plane + (size_t)v * W + u] = d_acc;
  }
}

// ------------------------------------------------------------------------------------------
// host side: parameter block, workspace carving, launch sequences
// ------------------------------------------------------------------------------------------
static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

struct WorkspaceLayout {
  size_t means, smooth_img, clip_thr, one, loss_part, pose_part, clip_part, total;
};

static WorkspaceLayout carve(const psfm_loss_cfg* cfg, LossParams& P) {
  WorkspaceLayout L;
  const size_t rs = cfg->real_dtype == PSFM_F64 ? 8 : 4;
  const int ngroups = cfg->automask ? 2 * cfg->C : cfg->C;
  size_t off = 0, lp = 0, pp = 0, cp = 0;
  for (int s = 0; s < cfg->n; ++s) {
    ScaleDesc& S = P.sc[s];
    S.tiles_x = (cfg->Ws[s] + kTileW - 1) / kTileW;
    S.tiles_y = (cfg->Hs[s] + kTileH - 1) / kTileH;
    const size_t nblk = (size_t)cfg->B * S.tiles_x * S.tiles_y;
    S.part_off = lp; lp += nblk * kLossVals;
    S.pose_off = pp; pp += nblk * cfg->C * 12;
    S.clip_off = cp; cp += nblk * ngroups * 2;
  }
  L.means = off;      off = align_up(off + (size_t)cfg->n * cfg->B * rs);
  L.smooth_img = off; off = align_up(off + (size_t)cfg->n * cfg->B * 2 * rs);
  L.clip_thr = off;   off = align_up(off + (size_t)cfg->n * ngroups * 3 * rs);
  L.one = off;        off = align_up(off + rs);
  L.loss_part = off;  off = align_up(off + lp * sizeof(double));
  L.pose_part = off;  off = align_up(off + pp * rs);
  L.clip_part = off;  off = align_up(off + (cfg->clip_loss > 0.0 ? cp * sizeof(double) : 0));
  L.total = off;
  return L;
}

size_t stream_workspace_bytes(const psfm_loss_cfg* cfg, const psfm_loss_io* io);        // loss_sweep.cu
bool stream_eligible(const psfm_loss_cfg* cfg);
cudaError_t loss_stream_dispatch(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws_split, void* ws_stream,
                                 void* loss_out, cudaStream_t st);

static size_t split_workspace_bytes(const psfm_loss_cfg* cfg) {
  LossParams P;
  return carve(cfg, P).total;
}

// [ tile-kernel (split) region | streaming-kernel region ]
// io == nullptr: by configuration alone (every scale its own image set); with io: for the image sets that exist
size_t loss_workspace_bytes(const psfm_loss_cfg* cfg, const psfm_loss_io* io) {
  return split_workspace_bytes(cfg) + (stream_eligible(cfg) ? stream_workspace_bytes(cfg, io) : 0);
}

void build_params(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, LossParams& P) {
  P.B = cfg->B; P.C = cfg->C; P.n = cfg->n; P.H = cfg->H; P.W = cfg->W;
  P.padding = cfg->padding; P.reduce_op = cfg->reduce_op; P.automask = cfg->automask; P.premap = cfg->premap; P.camera = cfg->camera_model;
  const int ngroups = cfg->automask ? 2 * cfg->C : cfg->C;
  P.ncand = ngroups;    // thresholds are per candidate group
  P.ssim_w = cfg->ssim_w; P.C1 = cfg->C1; P.C2 = cfg->C2; P.clip_loss = cfg->clip_loss; P.smooth_w = cfg->smooth_w;
  const double dmax = cfg->max_depth > 1e-6 ? cfg->max_depth : 1e-6, dmin = cfg->min_depth > 1e-6 ? cfg->min_depth : 1e-6;
  P.premap_lo = 1.0 / dmax; P.premap_hi = 1.0 / dmin;
  const WorkspaceLayout L = carve(cfg, P);
  char* w = (char*)ws;
  P.one = w + L.one;
  P.means = w + L.means; P.smooth_img = w + L.smooth_img; P.clip_thr = w + L.clip_thr;
  P.loss_part = (double*)(w + L.loss_part); P.pose_part = w + L.pose_part; P.clip_part = (double*)(w + L.clip_part);
  for (int s = 0; s < cfg->n; ++s) {
    ScaleDesc& S = P.sc[s];
    S.image = io->image[s]; S.invd = io->inv_depth[s]; S.mask = io->mask[s]; S.sel = io->sel[s]; S.dbg = io->dbg[s];
    S.d_invd = io->d_inv_depth[s];
    for (int j = 0; j < cfg->C; ++j) S.ctx[j] = io->ctx[s][j];
    S.H = cfg->Hs[s]; S.W = cfg->Ws[s];
    S.Hd = cfg->Hd[s] > 0 ? cfg->Hd[s] : S.H; S.Wd = cfg->Wd[s] > 0 ? cfg->Wd[s] : S.W;
    S.rs = (float)S.Hd / (float)S.H; S.cs = (float)S.Wd / (float)S.W;
    S.d_full = S.d_invd;
    S.sf = (double)S.W / (double)cfg->W;
    const double px = (double)cfg->B * S.H * S.W;
    S.inv_count = 1.0 / ((double)cfg->n * px);
    const double sc = cfg->smooth_w / ((double)cfg->n * (double)(1 << s));
    S.coef_x = sc / ((double)cfg->B * S.H * (S.W - 1));
    S.coef_y = sc / ((double)cfg->B * (S.H - 1) * S.W);
  }
  P.K = io->K; P.refK = io->ref_K;
  for (int j = 0; j < cfg->C; ++j) { P.pose[j] = io->pose[j]; P.d_pose[j] = io->d_pose[j]; }
  P.grad_loss = nullptr; P.loss_out = nullptr;
}

template <typename real, typename img_t>
static cudaError_t run_forward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, void* loss_out,
                               cudaStream_t st) {
  LossParams P;
  build_params(cfg, io, ws, P);
  P.loss_out = loss_out;
  if (cfg->smooth_w > 0.0) {
    ProfScope ps(st, kTagMean, 1);
    mean_kernel<real><<<dim3(cfg->B, cfg->n), 1024, 0, st>>>(P);
  }
  if (cfg->clip_loss > 0.0) {
    ProfScope ps(st, kTagOther, cfg->n + 1);
    for (int s = 0; s < cfg->n; ++s)
      loss_fwd_kernel<real, img_t><<<dim3(P.sc[s].tiles_x, P.sc[s].tiles_y, cfg->B), kThreads, 0, st>>>(P, s, 1);
    clip_finish_kernel<real><<<1, 256, 0, st>>>(P);
  }
  {
    ProfScope ps(st, kTagFwd, cfg->n);
    for (int s = 0; s < cfg->n; ++s)
      loss_fwd_kernel<real, img_t><<<dim3(P.sc[s].tiles_x, P.sc[s].tiles_y, cfg->B), kThreads, 0, st>>>(P, s, 0);
  }
  {
    ProfScope ps(st, kTagFinish, 1);
    loss_finish_kernel<real><<<1, 256, 0, st>>>(P);
  }
  return cudaGetLastError();
}

template <typename real, typename img_t>
static cudaError_t run_backward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, const void* grad_loss,
                                cudaStream_t st) {
  LossParams P;
  build_params(cfg, io, ws, P);
  P.grad_loss = grad_loss;
  {
    ProfScope ps(st, kTagBwd, cfg->n);
    for (int s = 0; s < cfg->n; ++s)
      loss_bwd_kernel<real, img_t><<<dim3(P.sc[s].tiles_x, P.sc[s].tiles_y, cfg->B), kThreads, 0, st>>>(P, s);
  }
  {
    ProfScope ps(st, kTagFinish, 1);
    pose_finish_kernel<real><<<dim3(cfg->B, cfg->C), 32, 0, st>>>(P, 0);
  }
  return cudaGetLastError();
}

// single-pass entry point for configurations the streaming kernel does not cover: forward followed by
// backward with upstream gradient 1 (needs io->sel)
template <typename real, typename img_t>
static cudaError_t run_fused(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, void* loss_out, cudaStream_t st) {
  cudaError_t e = run_forward<real, img_t>(cfg, io, ws, loss_out, st);
  if (e != cudaSuccess) return e;
  LossParams P;
  build_params(cfg, io, ws, P);
  return run_backward<real, img_t>(cfg, io, ws, P.one, st);
}

cudaError_t loss_fused_dispatch(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, void* loss_out,
                                cudaStream_t st) {
  if (stream_eligible(cfg))
    return loss_stream_dispatch(cfg, io, ws, (char*)ws + split_workspace_bytes(cfg), loss_out, st);
  if (cfg->real_dtype == PSFM_F64) return run_fused<double, double>(cfg, io, ws, loss_out, st);
  if (cfg->img_dtype == PSFM_BF16) return run_fused<float, __nv_bfloat16>(cfg, io, ws, loss_out, st);
  return run_fused<float, float>(cfg, io, ws, loss_out, st);
}

cudaError_t loss_forward_dispatch(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, void* loss_out,
                                  cudaStream_t st) {
  if (cfg->real_dtype == PSFM_F64) return run_forward<double, double>(cfg, io, ws, loss_out, st);
  if (cfg->img_dtype == PSFM_BF16) return run_forward<float, __nv_bfloat16>(cfg, io, ws, loss_out, st);
  return run_forward<float, float>(cfg, io, ws, loss_out, st);
}

cudaError_t loss_backward_dispatch(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, const void* grad_loss,
                                   cudaStream_t st) {
  if (cfg->real_dtype == PSFM_F64) return run_backward<double, double>(cfg, io, ws, grad_loss, st);
  if (cfg->img_dtype == PSFM_BF16) return run_backward<float, __nv_bfloat16>(cfg, io, ws, grad_loss, st);
  return run_backward<float, float>(cfg, io, ws, grad_loss, st);
}

}  // namespace psfm
