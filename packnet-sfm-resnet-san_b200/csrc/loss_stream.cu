// loss_stream.cu -- single-pass ("fused") photometric loss: loss AND gradients in one sweep.
//
//   prep_kernel      identity (automask) candidates: per-pixel min + arg-min over the unwarped
//                    contexts, once per resolution (shared by all full-resolution scales);
//                    raw smoothness sums and inverse-depth sums per image and scale
//   stats_kernel     per (scale, image): clamped mean, smoothness sums, gradient coupling constant
//   stream_kernel    the hot kernel.  One CTA owns a column strip x row chunk of one image and
//                    ALL scales of equal resolution; warp groups are specialised by (scale,
//                    context).  The CTA marches down the rows; every thread owns one column and
//                    keeps the 3-row sliding sums of the SSIM window and of its adjoint in
//                    registers, so each warped value is produced once (no recompute, no halo
//                    re-gather besides 2 columns / 2 rows per strip / chunk) and feeds both the
//                    loss and d/d inv_depth, d/d pose.
//   final_kernel     fixed-order reduction of the CTA partials -> loss triple and d pose.mat
//
// Semantics: SURVEY.md Appendix A (reference losses/multiview_photometric_loss.py:199-410).
// Restrictions of this path (others use the tile kernels of loss_kernels.cu): photometric_reduce_op
// = 'min', ssim_loss_weight > 0, clip_loss = 0.
#include <cuda_runtime.h>
#include "psfm_device.cuh"
#include "loss_params.h"
#include "loss_tile.cuh"
#include "profile.h"

namespace psfm {

constexpr int kMaxGroups = 16;       // (scale, context) warp groups per CTA
constexpr int kStreamThreads = 512;  // upper bound of G * L
constexpr int kPrepVals = 3;         // sum d, sum |dx d| wx, sum |dy d| wy

struct StreamSet {                   // scales of equal resolution sharing the same images
  int ns;
  int scale[PSFM_MAX_SCALES];
  int H, W;
  int L;                             // lanes (columns incl. 2+2 halo) per group
  int chunk_h;                       // output rows per CTA
  int nstrips, nchunks;
  int ptiles_x, ptiles_y;            // prep kernel tiling
  void* idval;                       // real [B,H,W] min identity photometric value (automask)
  uint8_t* idsel;                    // u8   [B,H,W] its candidate index (2j+1)
};

struct StreamParams {
  LossParams P;
  StreamSet set;
  // per scale (indexed by the scale id)
  double* prep_part[PSFM_MAX_SCALES];   // [B][ptiles][kPrepVals]
  void* stats[PSFM_MAX_SCALES];         // real [B][4]: clamped mean, coupling constant, SX, SY
  double* loss_part[PSFM_MAX_SCALES];   // [B * nstrips * nchunks] photometric sums
  void* pose_part[PSFM_MAX_SCALES];     // real [B][nctas][C][12]
};

// ------------------------------------------------------------------------------------------
// prep kernel: grid (ptiles_x, ptiles_y, B), kThreads threads, 32x8 pixel tiles.
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t>
__global__ void __launch_bounds__(kThreads) prep_kernel(const StreamParams SP) {
  constexpr int SH = kTileH + 2, SW = kTileW + 2;
  const LossParams& P = SP.P;
  const StreamSet& S = SP.set;
  const ScaleDesc& S0 = P.sc[S.scale[0]];
  const int H = S.H, W = S.W, b = blockIdx.z;
  const int u0 = blockIdx.x * kTileW, v0 = blockIdx.y * kTileH;
  const int tx = threadIdx.x % kTileW, ty = threadIdx.x / kTileW;
  const int u = u0 + tx, v = v0 + ty;
  const bool valid = (u < W) && (v < H);
  const size_t plane = (size_t)H * W;
  const int tile = blockIdx.y * gridDim.x + blockIdx.x, ntiles = gridDim.x * gridDim.y;

  __shared__ real Ys[3][SH][SW];
  __shared__ real Xs[3][SH][SW];
  __shared__ real scratch[kPrepVals * (kThreads / 32)];

  const img_t* image = (const img_t*)S0.image + (size_t)b * 3 * plane;
  fill_image_tile<real, img_t, SH, SW>(Ys, image, plane, H, W, v0 - 1, u0 - 1);
  const real ssim_w = (real)P.ssim_w, C1 = (real)P.C1, C2 = (real)P.C2;

  if (P.automask) {
    real best = real_inf<real>();
    int sel = 1;
    for (int j = 0; j < P.C; ++j) {
      const img_t* ctx = (const img_t*)S0.ctx[j] + (size_t)b * 3 * plane;
      __syncthreads();
      fill_image_tile<real, img_t, SH, SW>(Xs, ctx, plane, H, W, v0 - 1, u0 - 1);
      __syncthreads();
      real val[3];
      photometric_pixel<real, SH, SW>(Xs, Ys, ty + 1, tx + 1, ssim_w, C1, C2, val);
      if (val[0] < best) { best = val[0]; sel = 2 * j + 1; }
    }
    if (valid) {
      ((real*)S.idval)[(size_t)b * plane + (size_t)v * W + u] = best;
      S.idsel[(size_t)b * plane + (size_t)v * W + u] = (uint8_t)sel;
    }
  } else {
    __syncthreads();
  }

  if (P.smooth_w > 0.0) {
    // edge weights depend on the image only; raw sums are normalised by the mean afterwards
    real wx = real(0), wy = real(0);
    const int ly = ty + 1, lx = tx + 1;
    if (valid && u + 1 < W) {
      real ig = 0;
#pragma unroll
      for (int c = 0; c < 3; ++c) ig += rabs(Ys[c][ly][lx] - Ys[c][ly][lx + 1]);
      wx = rexp(-(ig / real(3)));
    }
    if (valid && v + 1 < H) {
      real ig = 0;
#pragma unroll
      for (int c = 0; c < 3; ++c) ig += rabs(Ys[c][ly][lx] - Ys[c][ly + 1][lx]);
      wy = rexp(-(ig / real(3)));
    }
    for (int k = 0; k < S.ns; ++k) {
      const int s = S.scale[k];
      const real* invd = (const real*)P.sc[s].invd + (size_t)b * plane;
      real sums[kPrepVals] = {real(0), real(0), real(0)};
      if (valid) {
        const size_t o = (size_t)v * W + u;
        const real d0 = invd[o];
        sums[0] = d0;
        if (u + 1 < W) sums[1] = rabs(d0 - invd[o + 1]) * wx;
        if (v + 1 < H) sums[2] = rabs(d0 - invd[o + W]) * wy;
      }
      block_reduce<real, kPrepVals>(sums, scratch);
      if (threadIdx.x == 0) {
        double* o = SP.prep_part[s] + ((size_t)b * ntiles + tile) * kPrepVals;
#pragma unroll
        for (int i = 0; i < kPrepVals; ++i) o[i] = (double)sums[i];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// stats kernel: grid (B, ns), 128 threads.  stats[s][b] = {m', c, SX, SY}:
//   m' = max(mean d, 1e-6);  SX = sum|gx wx| = raw_x / m', SY likewise (utils/depth.py:146-198);
//   c  = (coef_x SX + coef_y SY) / (N m') when mean >= 1e-6 (gradient through the mean), else 0.
// ------------------------------------------------------------------------------------------
template <typename real>
__global__ void __launch_bounds__(128) stats_kernel(const StreamParams SP) {
  const LossParams& P = SP.P;
  const StreamSet& S = SP.set;
  const int b = blockIdx.x, s = S.scale[blockIdx.y];
  const int ntiles = S.ptiles_x * S.ptiles_y;
  __shared__ double sh[3][128];
  double a[3] = {0.0, 0.0, 0.0};
  const double* base = SP.prep_part[s] + (size_t)b * ntiles * kPrepVals;
  for (int i = threadIdx.x; i < ntiles; i += blockDim.x)
    for (int k = 0; k < 3; ++k) a[k] += base[(size_t)i * kPrepVals + k];
  for (int k = 0; k < 3; ++k) sh[k][threadIdx.x] = a[k];
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o)
      for (int k = 0; k < 3; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double N = (double)S.H * S.W;
    const real mean = (real)(sh[0][0] / N);
    const real m = rmax(mean, real(1e-6));
    const real sx = (real)sh[1][0] / m, sy = (real)sh[2][0] / m;
    real* o = (real*)SP.stats[s] + (size_t)b * 4;
    o[0] = m;
    o[1] = (mean >= real(1e-6)) ? ((real)P.sc[s].coef_x * sx + (real)P.sc[s].coef_y * sy) / ((real)N * m) : real(0);
    o[2] = sx;
    o[3] = sy;
  }
}

// ------------------------------------------------------------------------------------------
// stream kernel
// ------------------------------------------------------------------------------------------
template <typename real> struct StreamSmem {
  real* Yring;   // [8][3][L]   target rows (ring over padded rows)
  real* Ysum;    // [6][L]      window sums of y and y^2 for the row being evaluated
  real* Wx;      // [2][L]      smoothness weights exp(-mean|dx I|)
  real* Wy;      // [4][L]
  real* Idv;     // [L]         min identity photometric value of the row being evaluated
  real* Msk;     // [L]
  int* Ids;      // [L]
  real* Xrow;    // [G][3][L]   warped values of the newest row
  real* Photo;   // [G][L]
  real* Frow;    // [G][9][L]   weighted SSIM adjoint coefficients of the evaluated row
  real* Dpart;   // [G][L]      d/d inv_depth contributions of the contexts j > 0
  real* Drow;    // [ns][4][L]  raw inverse depths (ring) for the smoothness neighbours
  real* Priv;    // [G][3][10][L] per-thread ring: x[3], dx/dix[3], dx/diy[3], raw d
  real* Ts;      // [C][12]
  real* St;      // [ns][2]     clamped mean, coupling constant
  real* Red;     // [32][12]    final reductions
  CamPair<real>* cam;
};

template <typename real>
__host__ __device__ inline size_t stream_smem_reals(int G, int ns, int C, int L) {
  return (size_t)8 * 3 * L + 6 * L + 2 * L + 4 * L + L + L + L /*Ids as real-sized*/ + (size_t)G * 3 * L + (size_t)G * L +
         (size_t)G * 9 * L + (size_t)G * L + (size_t)ns * 4 * L + (size_t)G * 3 * 10 * L + (size_t)C * 12 + ns * 2 +
         32 * 12 + 18 + 8;
}

template <typename real, typename img_t, int L>
__global__ void __launch_bounds__(kStreamThreads, 1) stream_kernel(const StreamParams SP) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const LossParams& P = SP.P;
  const StreamSet& S = SP.set;
  const int C = P.C, ns = S.ns, G = ns * C;
  const int H = S.H, W = S.W, b = blockIdx.z;
  const int c0 = blockIdx.x * (L - 4);
  const int r0 = blockIdx.y * S.chunk_h;
  const int r1 = (r0 + S.chunk_h < H) ? r0 + S.chunk_h : H;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int g = tid / L, l = tid - g * L;
  const int si = g / C, j = g - si * C, s = S.scale[si];
  const int u = c0 - 2 + l, su = reflect1(u, W);
  const bool col_real = (u >= 0) && (u < W);
  const bool col_inner = (l >= 2) && (l < L - 2) && col_real;
  const int lm = l > 0 ? l - 1 : 0, lp = l < L - 1 ? l + 1 : L - 1;
  const size_t plane = (size_t)H * W;
  const int cta = (blockIdx.y * gridDim.x + blockIdx.x), nctas = gridDim.x * gridDim.y;

  StreamSmem<real> M;
  {
    real* p = (real*)smem_raw;
    M.Yring = p; p += 8 * 3 * L;
    M.Ysum = p; p += 6 * L;
    M.Wx = p; p += 2 * L;
    M.Wy = p; p += 4 * L;
    M.Idv = p; p += L;
    M.Msk = p; p += L;
    M.Ids = (int*)p; p += L;
    M.Xrow = p; p += (size_t)G * 3 * L;
    M.Photo = p; p += (size_t)G * L;
    M.Frow = p; p += (size_t)G * 9 * L;
    M.Dpart = p; p += (size_t)G * L;
    M.Drow = p; p += (size_t)ns * 4 * L;
    M.Priv = p; p += (size_t)G * 3 * 10 * L;
    M.Ts = p; p += C * 12;
    M.St = p; p += ns * 2;
    M.Red = p; p += 32 * 12;
    M.cam = (CamPair<real>*)p;
  }
  const ScaleDesc& SD = P.sc[s];
  const img_t* image = (const img_t*)SD.image + (size_t)b * 3 * plane;
  const img_t* ctx = (const img_t*)SD.ctx[j] + (size_t)b * 3 * plane;
  const real* invd = (const real*)SD.invd + (size_t)b * plane;
  const img_t* maskp = SD.mask ? (const img_t*)SD.mask + (size_t)b * plane : nullptr;
  const real* idval = (const real*)S.idval + (size_t)b * plane;
  const uint8_t* idsel = S.idsel + (size_t)b * plane;

  // ---- prologue: cameras, poses, statistics, first target rows ------------------------------------
  if (tid == 0) setup_cam<real>(*M.cam, (const real*)P.K, (const real*)P.refK, b, (real)SD.sf);
  for (int i = tid; i < C * 12; i += nthr) {
    const int jj = i / 12, e = i - jj * 12;
    M.Ts[i] = ((const real*)P.pose[jj])[(size_t)b * 16 + (e / 4) * 4 + (e % 4)];
  }
  for (int i = tid; i < ns * 2; i += nthr) {
    const int k = i / 2;
    M.St[i] = (P.smooth_w > 0.0) ? ((const real*)SP.stats[S.scale[k]])[(size_t)b * 4 + (i & 1)] : real(1);
  }
  auto yslot = [&](int row) { return ((row + 16) & 7) * 3 * L; };
  auto load_y_row = [&](int row) {
    const int sv = reflect1(row, H);
    for (int i = tid; i < 3 * L; i += nthr) {
      const int c = i / L, ll = i - c * L;
      M.Yring[yslot(row) + i] = ldimg<real>(image + c * plane, (size_t)sv * W + reflect1(c0 - 2 + ll, W));
    }
  };
  const int t0 = r0 - 2;
  load_y_row(t0 - 2); load_y_row(t0 - 1); load_y_row(t0);

  const real ssim_w = (real)P.ssim_w, C1 = (real)P.C1, C2 = (real)P.C2;
  const real inv_count = (real)SD.inv_count;
  const real mu0 = (u == 1) ? real(2) : real(1), mu2 = (u == W - 2) ? real(2) : real(1);
  const int mycand = P.automask ? 2 * j : j;
  real* Tj = M.Ts + j * 12;
  real* priv = M.Priv + (size_t)g * 3 * 10 * L + l;            // [slot][k] at priv[(slot*10 + k) * L]
  real* xrow = M.Xrow + (size_t)g * 3 * L;
  real* frow = M.Frow + (size_t)g * 9 * L;
  real* drow = M.Drow + (size_t)si * 4 * L;

  real h[3][9];      // ring over rows: per channel {sum x, sum x^2, sum x y} of the 3 columns
  real hF[3][9];     // ring over rows: horizontal sums of the adjoint coefficient rows
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int k = 0; k < 9; ++k) { h[a][k] = real(0); hF[a][k] = real(0); }
  real pose_acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) pose_acc[k] = real(0);
  real loss_acc = real(0), w_cur = real(0), w_old = real(0);
  real d1 = real(0), d2 = real(0), d3 = real(0), d0 = real(0);   // raw inverse depths of rows t, t-1, t-2, t-3
  real dsum_prev = real(0);
  int prev_q = -1;
  __syncthreads();

  auto finalize_row = [&]() {           // group j == 0 adds the other contexts' parts and stores the row
    if (j == 0 && prev_q >= 0) {
      real tot = dsum_prev;
      for (int jj = 1; jj < C; ++jj) tot += M.Dpart[(size_t)(si * C + jj) * L + l];
      if (col_inner) ((real*)SD.d_invd)[(size_t)b * plane + (size_t)prev_q * W + u] = tot;
    }
    prev_q = -1;
  };

  auto step = [&](const int t, const int PH) {
    const int p = t - 1, q = t - 2;
    const int sl_t = PH, sl_p = (PH + 2) % 3, sl_q = (PH + 1) % 3;
    // ---------------- phase (a): warp row t, cooperative loads, row statistics -------------------
    {
      const int sv = reflect1(t, H);
      d3 = d2; d2 = d1; d1 = d0;
      d0 = invd[(size_t)sv * W + su];
      real chain, dd = d0;
      if (P.premap) dd = premap_fork<real>(d0, (real)P.premap_lo, (real)P.premap_hi, chain);
      PixGeom<real> gm;
      pixel_geometry<real>(*M.cam, Tj, (real)su, (real)sv, dd, H, W, gm);
      real x[3], gx[3], gy[3];
      bilinear3<real, img_t, true>(ctx, plane, H, W, P.padding, gm.ix, gm.iy, x, gx, gy);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        xrow[c * L + l] = x[c];
        priv[(sl_t * 10 + c) * L] = x[c];
        priv[(sl_t * 10 + 3 + c) * L] = gx[c];
        priv[(sl_t * 10 + 6 + c) * L] = gy[c];
      }
      priv[(sl_t * 10 + 9) * L] = d0;
      if (j == 0) drow[((t + 16) & 3) * L + l] = d0;
    }
    load_y_row(t + 1);
    {
      const bool p_row_real = (p >= 0) && (p < H);
      for (int i = tid; i < L; i += nthr) {
        const int uu = c0 - 2 + i;
        const bool ok = p_row_real && uu >= 0 && uu < W;
        if (P.automask) {
          M.Idv[i] = ok ? idval[(size_t)p * W + uu] : real(0);
          M.Ids[i] = ok ? (int)idsel[(size_t)p * W + uu] : 1;
        }
        if (maskp) M.Msk[i] = ok ? ldimg<real>(maskp, (size_t)p * W + uu) : real(0);
      }
      // window sums of y, y^2 at row p (rows t-2..t, columns ll-1..ll+1)
      for (int i = tid; i < 6 * L; i += nthr) {
        const int k = i / L, ll = i - k * L, c = k % 3;
        const int a = ll > 0 ? ll - 1 : 0, e = ll < L - 1 ? ll + 1 : L - 1;
        real acc = real(0);
#pragma unroll
        for (int dr = 0; dr < 3; ++dr) {
          const real* yr = M.Yring + yslot(t - 2 + dr) + c * L;
          const real ya = yr[a], yb = yr[ll], yc = yr[e];
          acc += (k < 3) ? (ya + yb + yc) : (ya * ya + yb * yb + yc * yc);
        }
        M.Ysum[i] = acc;
      }
      // smoothness weights of row q (need rows q, q+1 = t-2, t-1)
      if (P.smooth_w > 0.0) {
        for (int i = tid; i < 2 * L; i += nthr) {
          const int k = i / L, ll = i - k * L;
          const int e = ll < L - 1 ? ll + 1 : L - 1;
          const real* y0 = M.Yring + yslot(q);
          const real* y1 = M.Yring + yslot(q + 1);
          real ig = real(0);
#pragma unroll
          for (int c = 0; c < 3; ++c)
            ig += (k == 0) ? rabs(y0[c * L + ll] - y0[c * L + e]) : rabs(y0[c * L + ll] - y1[c * L + ll]);
          if (k == 0) M.Wx[((q + 16) & 1) * L + ll] = rexp(-(ig / real(3)));
          else M.Wy[((q + 16) & 3) * L + ll] = rexp(-(ig / real(3)));
        }
      }
    }
    __syncthreads();                                                               // sync #1
    // ---------------- phase (b): finish the previous output row, SSIM at row p --------------------
    finalize_row();
    real Ft[9];
    {
      const real* yt = M.Yring + yslot(t);
      const real* yp = M.Yring + yslot(p);
      real dist_sum = real(0), l1_sum = real(0);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const real xa = xrow[c * L + lm], xb = xrow[c * L + l], xc = xrow[c * L + lp];
        const real ya = yt[c * L + lm], yb = yt[c * L + l], yc = yt[c * L + lp];
        h[sl_t][c * 3 + 0] = xa + xb + xc;
        h[sl_t][c * 3 + 1] = xa * xa + xb * xb + xc * xc;
        h[sl_t][c * 3 + 2] = xa * ya + xb * yb + xc * yc;
        const real r9 = real(1) / real(9);
        const real mx = (h[0][c * 3] + h[1][c * 3] + h[2][c * 3]) * r9;
        const real exx = (h[0][c * 3 + 1] + h[1][c * 3 + 1] + h[2][c * 3 + 1]) * r9;
        const real exy = (h[0][c * 3 + 2] + h[1][c * 3 + 2] + h[2][c * 3 + 2]) * r9;
        const real my = M.Ysum[c * L + l] * r9, eyy = M.Ysum[(3 + c) * L + l] * r9;
        real dmx, dexx, dexy;
        dist_sum += ssim_dist_grad<real>(mx, my, exx, eyy, exy, C1, C2, dmx, dexx, dexy);
        Ft[c * 3 + 0] = dmx * r9; Ft[c * 3 + 1] = real(2) * dexx * r9; Ft[c * 3 + 2] = dexy * r9;
        l1_sum += rabs(priv[(sl_p * 10 + c) * L] - yp[c * L + l]);
      }
      M.Photo[(size_t)g * L + l] = ssim_w * (dist_sum / real(3)) + (real(1) - ssim_w) * (l1_sum / real(3));
    }
    __syncthreads();                                                               // sync #2
    // ---------------- phase (c): arg-min over the candidates of row p, adjoint coefficients -------
    {
      const real m = maskp ? M.Msk[l] : real(1);
      real best = real_inf<real>();
      int sel = 0;
      for (int jj = 0; jj < C; ++jj) {
        const real v = M.Photo[(size_t)(si * C + jj) * L + l] * m;
        if (v < best) { best = v; sel = P.automask ? 2 * jj : jj; }
      }
      if (P.automask) {
        const real vi = M.Idv[l] * m;
        const int ki = M.Ids[l];
        if (vi < best || (vi == best && ki < sel)) { best = vi; sel = ki; }
      }
      const bool p_real = (p >= 0) && (p < H) && col_real;
      w_old = w_cur;
      w_cur = (p_real && sel == mycand) ? inv_count * m : real(0);
      const real wf = w_cur * ssim_w / real(3);
#pragma unroll
      for (int k = 0; k < 9; ++k) frow[k * L + l] = wf * Ft[k];
      if (j == 0 && p_real && col_inner && p >= r0 && p < r1) loss_acc += best;
    }
    __syncthreads();                                                               // sync #3
    // ---------------- phase (d): adjoint at row q, sampler + geometry adjoint, smoothness ---------
    {
#pragma unroll
      for (int k = 0; k < 9; ++k)
        hF[sl_p][k] = mu0 * frow[k * L + lm] + frow[k * L + l] + mu2 * frow[k * L + lp];
      if (q >= r0 && q < r1) {
        const real mv0 = (q == 1) ? real(2) : real(1), mv2 = (q == H - 2) ? real(2) : real(1);
        const real* yq = M.Yring + yslot(q);
        real gix = real(0), giy = real(0);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const real sa = mv0 * hF[sl_t][c * 3 + 0] + hF[sl_q][c * 3 + 0] + mv2 * hF[sl_p][c * 3 + 0];
          const real sb = mv0 * hF[sl_t][c * 3 + 1] + hF[sl_q][c * 3 + 1] + mv2 * hF[sl_p][c * 3 + 1];
          const real sc = mv0 * hF[sl_t][c * 3 + 2] + hF[sl_q][c * 3 + 2] + mv2 * hF[sl_p][c * 3 + 2];
          const real xq = priv[(sl_q * 10 + c) * L], yv = yq[c * L + l];
          const real dl = sa + xq * sb + yv * sc + w_old * ((real(1) - ssim_w) / real(3)) * rsign(xq - yv);
          gix += dl * priv[(sl_q * 10 + 3 + c) * L];
          giy += dl * priv[(sl_q * 10 + 6 + c) * L];
        }
        real dphoto = real(0);
        if (col_inner && (gix != real(0) || giy != real(0))) {
          real chain = real(1), dd = d2;
          if (P.premap) dd = premap_fork<real>(d2, (real)P.premap_lo, (real)P.premap_hi, chain);
          PixGeom<real> gm;
          pixel_geometry<real>(*M.cam, Tj, (real)u, (real)q, dd, H, W, gm);
          dphoto = geometry_adjoint<real>(*M.cam, Tj, gm, dd, gix, giy, pose_acc) * chain;
        }
        if (j == 0) {
          real gs = real(0);
          if (P.smooth_w > 0.0 && col_inner) {
            const real mclamp = M.St[si * 2], cb = M.St[si * 2 + 1];
            const real* dq = drow + ((q + 16) & 3) * L;
            const real* wxq = M.Wx + ((q + 16) & 1) * L;
            real qq = real(0);
            if (u + 1 < W) qq += (real)SD.coef_x * rsign(d2 - dq[lp]) * wxq[l];
            if (u >= 1) qq -= (real)SD.coef_x * rsign(dq[lm] - d2) * wxq[lm];
            if (q + 1 < H) qq += (real)SD.coef_y * rsign(d2 - d1) * M.Wy[((q + 16) & 3) * L + l];
            if (q >= 1) qq -= (real)SD.coef_y * rsign(d3 - d2) * M.Wy[((q + 15) & 3) * L + l];
            gs = qq / mclamp - cb;
          }
          dsum_prev = dphoto + gs;
          prev_q = q;
        } else {
          M.Dpart[(size_t)g * L + l] = dphoto;
        }
      }
    }
  };

  for (int tb = t0; tb <= r1 + 1; tb += 3) {
    step(tb, 0);
    step(tb + 1, 1);
    step(tb + 2, 2);
  }
  __syncthreads();
  finalize_row();

  // ---- epilogue: deterministic reductions of the loss and pose partials ---------------------------
  const int lane = tid & 31, warp = tid >> 5, wpg = L / 32;
#pragma unroll
  for (int k = 0; k < 12; ++k) {
    real v = pose_acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) M.Red[warp * 12 + k] = v;
  }
  __syncthreads();
  if (l < 12) {
    real v = real(0);
    for (int w = 0; w < wpg; ++w) v += M.Red[(g * wpg + w) * 12 + l];
    ((real*)SP.pose_part[s])[(((size_t)b * nctas + cta) * C + j) * 12 + l] = v;
  }
  __syncthreads();
  {
    real v = loss_acc;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) M.Red[warp] = v;
  }
  __syncthreads();
  if (j == 0 && l == 0) {
    real v = real(0);
    for (int w = 0; w < wpg; ++w) v += M.Red[g * wpg + w];
    SP.loss_part[s][(size_t)b * nctas + cta] = (double)v;
  }
}

// ------------------------------------------------------------------------------------------
// final kernel: block 0 -> loss triple; block 1 + b * C + j -> d pose[j].mat[b].
// Works over all sets (called once, after every set's stream kernel).
// ------------------------------------------------------------------------------------------
struct FinalParams {
  int B, C, n;
  int nctas[PSFM_MAX_SCALES];
  double inv_count[PSFM_MAX_SCALES], coef_x[PSFM_MAX_SCALES], coef_y[PSFM_MAX_SCALES];
  const double* loss_part[PSFM_MAX_SCALES];
  const void* pose_part[PSFM_MAX_SCALES];
  const void* stats[PSFM_MAX_SCALES];
  void* d_pose[PSFM_MAX_CTX];
  void* loss_out;
  void* one;
  int smooth;
};

template <typename real>
__global__ void __launch_bounds__(256) final_kernel(const FinalParams F) {
  __shared__ double sh[256];
  const int tid = threadIdx.x;
  if (blockIdx.x == 0) {
    double photo = 0.0, smooth = 0.0;
    for (int s = 0; s < F.n; ++s) {
      double a = 0.0;
      const size_t cnt = (size_t)F.B * F.nctas[s];
      for (size_t i = tid; i < cnt; i += blockDim.x) a += F.loss_part[s][i];
      sh[tid] = a;
      __syncthreads();
      for (int o = 128; o > 0; o >>= 1) {
        if (tid < o) sh[tid] += sh[tid + o];
        __syncthreads();
      }
      photo += sh[0] * F.inv_count[s];
      __syncthreads();
      if (F.smooth && tid == 0)
        for (int b = 0; b < F.B; ++b) {
          const real* st = (const real*)F.stats[s] + (size_t)b * 4;
          smooth += (double)st[2] * F.coef_x[s] + (double)st[3] * F.coef_y[s];
        }
    }
    if (tid == 0) {
      real* out = (real*)F.loss_out;
      out[0] = (real)((real)photo + (real)smooth);
      out[1] = (real)photo;
      out[2] = (real)smooth;
      *(real*)F.one = real(1);
    }
    return;
  }
  const int bj = blockIdx.x - 1, b = bj / F.C, j = bj - b * F.C;
  // 12 entries x up to 21 partial lanes, fixed order: lane k of entry e sums ctas k, k+21, ...
  const int e = tid % 12, k = tid / 12;      // 252 active threads
  double a = 0.0;
  if (tid < 252) {
    for (int s = 0; s < F.n; ++s) {
      const real* base = (const real*)F.pose_part[s] + (((size_t)b * F.nctas[s]) * F.C + j) * 12 + e;
      for (int i = k; i < F.nctas[s]; i += 21) a += (double)base[(size_t)i * F.C * 12];
    }
  }
  sh[tid] = a;
  __syncthreads();
  if (tid < 12) {
    double v = 0.0;
    for (int kk = 0; kk < 21; ++kk) v += sh[kk * 12 + tid];
    ((real*)F.d_pose[j])[(size_t)b * 16 + (tid / 4) * 4 + (tid % 4)] = (real)v;
  } else if (tid < 16) {
    ((real*)F.d_pose[j])[(size_t)b * 16 + tid] = real(0);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static inline size_t al(size_t x) { return (x + 255) / 256 * 256; }

bool stream_eligible(const psfm_loss_cfg* cfg) {
  if (cfg->reduce_op != PSFM_REDUCE_MIN || !(cfg->ssim_w > 0.0) || cfg->clip_loss > 0.0) return false;
  if (cfg->C > kMaxGroups) return false;
  for (int s = 0; s < cfg->n; ++s)
    if (cfg->Hs[s] < 4 || cfg->Ws[s] < 4) return false;
  return true;
}

struct StreamPlan {
  int nsets;
  StreamSet sets[PSFM_MAX_SCALES];
  size_t off_idval[PSFM_MAX_SCALES], off_idsel[PSFM_MAX_SCALES];          // per set
  size_t off_prep[PSFM_MAX_SCALES], off_stats[PSFM_MAX_SCALES], off_loss[PSFM_MAX_SCALES],
      off_pose[PSFM_MAX_SCALES];                                          // per scale
  int nctas[PSFM_MAX_SCALES];
  size_t off_one, total;
};

static int pick_chunk(int H, int nstrips, int B) {
  // rows per CTA: keep >= ~4 waves of CTAs on 148 SMs while amortising the 4 warm-up rows
  int best = H;
  for (int ch = 96; ch >= 12; ch -= 4) {
    const long ctas = (long)nstrips * ((H + ch - 1) / ch) * B;
    best = ch;
    if (ctas >= 4 * 148) break;
  }
  return best < H ? best : H;
}

static void make_plan(const psfm_loss_cfg* cfg, const psfm_loss_io* io, StreamPlan& pl) {
  const size_t rs = cfg->real_dtype == PSFM_F64 ? 8 : 4;
  pl.nsets = 0;
  bool used[PSFM_MAX_SCALES] = {false, false, false, false};
  const int maxg = kStreamThreads / 32;
  for (int s = 0; s < cfg->n; ++s) {
    if (used[s]) continue;
    StreamSet& S = pl.sets[pl.nsets++];
    S.ns = 0; S.H = cfg->Hs[s]; S.W = cfg->Ws[s];
    for (int k = s; k < cfg->n; ++k) {
      if (used[k] || cfg->Hs[k] != S.H || cfg->Ws[k] != S.W) continue;
      if (io && (io->image[k] != io->image[s] || io->mask[k] != io->mask[s])) continue;
      bool same = true;
      if (io) for (int j = 0; j < cfg->C; ++j) same = same && io->ctx[k][j] == io->ctx[s][j];
      if (!same || (S.ns + 1) * cfg->C > maxg) continue;
      S.scale[S.ns++] = k;
      used[k] = true;
    }
    const int G = S.ns * cfg->C;
    S.L = (G * 64 <= kStreamThreads && S.W > 40) ? 64 : 32;
    S.nstrips = (S.W + S.L - 5) / (S.L - 4);
    S.chunk_h = pick_chunk(S.H, S.nstrips, cfg->B);
    if (S.chunk_h < 12) S.chunk_h = S.H < 12 ? S.H : 12;
    S.nchunks = (S.H + S.chunk_h - 1) / S.chunk_h;
    S.ptiles_x = (S.W + kTileW - 1) / kTileW;
    S.ptiles_y = (S.H + kTileH - 1) / kTileH;
  }
  // workspace layout does not depend on the grouping: every scale owns worst-case sized regions
  size_t off = 0;
  pl.off_one = off; off = al(off + rs);
  size_t idv[PSFM_MAX_SCALES], ids[PSFM_MAX_SCALES];
  for (int s = 0; s < cfg->n; ++s) {
    const size_t px = (size_t)cfg->B * cfg->Hs[s] * cfg->Ws[s];
    const size_t worst = (size_t)((cfg->Ws[s] + 27) / 28) * ((cfg->Hs[s] + 11) / 12);
    const size_t pt = (size_t)((cfg->Ws[s] + kTileW - 1) / kTileW) * ((cfg->Hs[s] + kTileH - 1) / kTileH);
    idv[s] = off; off = al(off + px * rs);
    ids[s] = off; off = al(off + px);
    pl.off_prep[s] = off;  off = al(off + (size_t)cfg->B * pt * kPrepVals * sizeof(double));
    pl.off_stats[s] = off; off = al(off + (size_t)cfg->B * 4 * rs);
    pl.off_loss[s] = off;  off = al(off + (size_t)cfg->B * worst * sizeof(double));
    pl.off_pose[s] = off;  off = al(off + (size_t)cfg->B * worst * cfg->C * 12 * rs);
  }
  for (int i = 0; i < pl.nsets; ++i) {
    const StreamSet& S = pl.sets[i];
    pl.off_idval[i] = idv[S.scale[0]];
    pl.off_idsel[i] = ids[S.scale[0]];
    for (int k = 0; k < S.ns; ++k) pl.nctas[S.scale[k]] = S.nstrips * S.nchunks;
  }
  pl.total = off;
}

size_t stream_workspace_bytes(const psfm_loss_cfg* cfg) {
  StreamPlan pl;
  make_plan(cfg, nullptr, pl);
  return pl.total;
}

void build_params(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, LossParams& P);   // loss_kernels.cu

template <typename real, typename img_t>
static cudaError_t run_stream(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws_split, void* ws_stream,
                              void* loss_out, cudaStream_t st) {
  StreamPlan pl;
  make_plan(cfg, io, pl);
  StreamParams SP;
  build_params(cfg, io, ws_split, SP.P);
  char* w = (char*)ws_stream;
  for (int s = 0; s < cfg->n; ++s) {
    SP.prep_part[s] = (double*)(w + pl.off_prep[s]);
    SP.stats[s] = w + pl.off_stats[s];
    SP.loss_part[s] = (double*)(w + pl.off_loss[s]);
    SP.pose_part[s] = w + pl.off_pose[s];
  }
  for (int i = 0; i < pl.nsets; ++i) {
    SP.set = pl.sets[i];
    SP.set.idval = w + pl.off_idval[i];
    SP.set.idsel = (uint8_t*)(w + pl.off_idsel[i]);
    const StreamSet& S = SP.set;
    if (cfg->automask || cfg->smooth_w > 0.0) {
      ProfScope ps(st, kTagFwd, cfg->smooth_w > 0.0 ? 2 : 1);
      prep_kernel<real, img_t><<<dim3(S.ptiles_x, S.ptiles_y, cfg->B), kThreads, 0, st>>>(SP);
      if (cfg->smooth_w > 0.0) stats_kernel<real><<<dim3(cfg->B, S.ns), 128, 0, st>>>(SP);
    }
    const int G = S.ns * cfg->C;
    const size_t smem = stream_smem_reals<real>(G, S.ns, cfg->C, S.L) * sizeof(real);
    dim3 grid(S.nstrips, S.nchunks, cfg->B);
    ProfScope ps(st, kTagFused, 1);
    if (S.L == 64) {
      auto kern = stream_kernel<real, img_t, 64>;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      kern<<<grid, G * 64, smem, st>>>(SP);
    } else {
      auto kern = stream_kernel<real, img_t, 32>;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      kern<<<grid, G * 32, smem, st>>>(SP);
    }
  }
  FinalParams F;
  F.B = cfg->B; F.C = cfg->C; F.n = cfg->n; F.smooth = cfg->smooth_w > 0.0;
  for (int s = 0; s < cfg->n; ++s) {
    F.nctas[s] = pl.nctas[s];
    F.inv_count[s] = SP.P.sc[s].inv_count; F.coef_x[s] = SP.P.sc[s].coef_x; F.coef_y[s] = SP.P.sc[s].coef_y;
    F.loss_part[s] = SP.loss_part[s]; F.pose_part[s] = SP.pose_part[s]; F.stats[s] = SP.stats[s];
  }
  for (int j = 0; j < cfg->C; ++j) F.d_pose[j] = io->d_pose[j];
  F.loss_out = loss_out;
  F.one = w + pl.off_one;
  {
    ProfScope ps(st, kTagFinish, 1);
    final_kernel<real><<<1 + cfg->B * cfg->C, 256, 0, st>>>(F);
  }
  return cudaGetLastError();
}

cudaError_t loss_stream_dispatch(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws_split, void* ws_stream,
                                 void* loss_out, cudaStream_t st) {
  if (cfg->real_dtype == PSFM_F64) return run_stream<double, double>(cfg, io, ws_split, ws_stream, loss_out, st);
  if (cfg->img_dtype == PSFM_BF16) return run_stream<float, __nv_bfloat16>(cfg, io, ws_split, ws_stream, loss_out, st);
  return run_stream<float, float>(cfg, io, ws_split, ws_stream, loss_out, st);
}

}  // namespace psfm
