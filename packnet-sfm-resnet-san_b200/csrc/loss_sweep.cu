// loss_sweep.cu -- single-pass photometric loss: loss AND gradients in one sweep over the inputs.
//
//   prep_kernel    once per resolution (shared by all full-resolution scales): every context image as a
//                  channel-interleaved, zero-bordered 4-vector image in the arithmetic type, so that one
//                  bilinear tap is ONE 16-byte load and taps outside the image need no bounds test (zeros
//                  padding reads the border); per-pixel records with
//                  the target pixel, its window statistics (mean, variance + C2), the mask, the
//                  automask candidate (min / arg-min of the identity photometric terms), the edge
//                  weights of the smoothness term and the raw per-image sums its normalisation needs.
//   stats_kernel   per (scale, image): clamped mean, smoothness sums, gradient coupling constant.
//   sweep_kernel   the hot kernel.  One warp = 32 adjacent columns of one (image, scale, context);
//                  it marches down the rows with the 3-row SSIM window and its adjoint in register
//                  rings and exchanges horizontal neighbours by shuffles, so every warped value is
//                  gathered once and feeds the loss, d/d inv_depth and d/d pose.  The C warps of a
//                  CTA (one per context) meet once per row to take the per-pixel min.
//   final_kernel   fixed-order reduction of the CTA partials -> loss triple and d pose.mat.
//
// Semantics: SURVEY.md Appendix A (reference losses/multiview_photometric_loss.py:199-410).
// Restrictions (other settings run on the tile kernels of loss_kernels.cu): photometric_reduce_op
// = 'min', ssim_loss_weight > 0, clip_loss = 0.
#include <cuda_runtime.h>
#include <cstdlib>
#include "sweep_device.cuh"
#include "loss_params.h"
#include "loss_tile.cuh"
#include "profile.h"

#include "sweep_common.cuh"

namespace psfm {


template <typename real, typename img_t, int CT, bool AUTOMASK, bool FIRST>
__global__ void __launch_bounds__(32 * kPrepWarps) prep_kernel(const SweepParams SP, const int j0) {
  typedef P2<real> pr;
  typedef Real4<real> vec4;
  const LossParams& P = SP.P;
  const SweepSet& S = SP.set;
  const int lane = threadIdx.x & 31;
  const int punits = S.pstrips * S.pchunks;
  const int H = S.H, W = S.W, b = blockIdx.y;
  const int nprep = (punits + kPrepWarps - 1) / kPrepWarps;
  if ((int)blockIdx.x >= nprep) {
    // trailing blocks: zero frame of this pass's packed context images, one warp per padded row (a frame row is
    // cleared entirely, an image row only in its 2 + 3 frame columns)
    const int r = ((int)blockIdx.x - nprep) * kPrepWarps + (threadIdx.x >> 5);
    if (r >= S.Hp * CT) return;
    const int jj = r / S.Hp, yp = r - jj * S.Hp;
    vec4* row = (vec4*)S.ctxp[j0 + jj] + ((size_t)b * S.Hp + yp) * S.Wp;
    if (yp < kPadLo || yp >= kPadLo + H) {
      for (int x = lane; x < S.Wp; x += 32) st4(row + x, real(0), real(0), real(0), real(0));
    } else if (lane < kPadLo + kPadHi) {
      st4(row + (lane < kPadLo ? lane : W + lane), real(0), real(0), real(0), real(0));
    }
    return;
  }
  const int unit = blockIdx.x * kPrepWarps + (threadIdx.x >> 5);
  if (unit >= punits) return;
  const int strip = unit % S.pstrips, chunk = unit / S.pstrips;
  const int plane = H * W;
  const int u = strip * kPrepInner - 1 + lane, su = reflect1(u, W);
  const int r0 = chunk * S.pchunk_h, r1 = min(r0 + S.pchunk_h, H);
  const bool inner = (lane >= 1) && (lane <= kPrepInner) && (u < W);
  constexpr bool first = FIRST;
  const ScaleDesc& S0 = P.sc[S.scale[0]];
  const img_t* y0p = (const img_t*)S0.image + (size_t)b * 3 * plane;
  const img_t *y1p = y0p + plane, *y2p = y0p + 2 * plane;
  asm volatile("" : "+l"(y0p), "+l"(y1p), "+l"(y2p));      // opaque: one IMAD.WIDE per load off a 32-bit offset
  const img_t* xp0[CT];
  vec4* cpk[CT];
#pragma unroll
  for (int jj = 0; jj < CT; ++jj) {
    xp0[jj] = (const img_t*)S0.ctx[j0 + jj] + (size_t)b * 3 * plane;
    cpk[jj] = (vec4*)S.ctxp[j0 + jj] + (size_t)b * S.Hp * S.Wp + kPadLo * S.Wp + kPadLo;
  }
  const int Wh = S.Wh, rplane = H * 2 * Wh * 3;
  vec4* ra = (vec4*)S.ra + (size_t)b * rplane;
  vec4* rb = (vec4*)S.rb + (size_t)b * rplane;
  vec4* rc = (vec4*)S.rc + (size_t)b * rplane;
  real* wxy = (real*)S.wxy + (size_t)b * plane * 2;
  const img_t* maskp = S0.mask ? (const img_t*)S0.mask + (size_t)b * plane : nullptr;
  const real C1 = (real)P.C1, C2 = (real)P.C2;
  const real kd = (real)P.ssim_w / real(3), kl = (real(1) - (real)P.ssim_w) / real(3);
  const real r3 = real(1) / real(3), r9 = real(1) / real(9);
  const bool smooth = first && P.smooth_w > 0.0;

  real sums[PSFM_MAX_SCALES][kPrepVals];
#pragma unroll
  for (int k = 0; k < PSFM_MAX_SCALES; ++k)
#pragma unroll
    for (int i = 0; i < kPrepVals; ++i) sums[k][i] = real(0);

  // horizontal sums of the two previous rows: y: {sum y, sum y^2}; context: {sum x, sum x^2, sum x y}
  pr hyA[2][2], hxA[CT][2][3], ypA(real(0)), ycpA(real(0)), xpA[CT];
  real hyB[2][2], hxB[CT][2][3], ypB = real(0), ycpB = real(0), xpB[CT];
  real dp[PSFM_MAX_SCALES];
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    hyA[a][0] = pr(real(0)); hyA[a][1] = pr(real(0)); hyB[a][0] = real(0); hyB[a][1] = real(0);
#pragma unroll
    for (int jj = 0; jj < CT; ++jj)
#pragma unroll
      for (int m = 0; m < 3; ++m) { hxA[jj][a][m] = pr(real(0)); hxB[jj][a][m] = real(0); }
  }
#pragma unroll
  for (int jj = 0; jj < CT; ++jj) { xpA[jj] = pr(real(0)); xpB[jj] = real(0); }
#pragma unroll
  for (int k = 0; k < PSFM_MAX_SCALES; ++k) dp[k] = real(0);

  // inverse-depth maps: at (H, W), or smaller and nearest-upsampled here (ScaleDesc::Hd, Wd)
  const real* dptr[PSFM_MAX_SCALES];
  int dcol[PSFM_MAX_SCALES];
  bool any_ups = false;
#pragma unroll
  for (int k = 0; k < PSFM_MAX_SCALES; ++k) {
    const ScaleDesc& Sk = P.sc[S.scale[k < S.ns ? k : 0]];
    dptr[k] = (const real*)Sk.invd + (size_t)b * Sk.Hd * Sk.Wd;
    dcol[k] = nearest_src(su, Sk.cs, Sk.Wd, Sk.Wd != W);
    any_ups = any_ups || Sk.Hd != H || Sk.Wd != W;
  }
  const int Wp = S.Wp, ns = S.ns;
  // software pipeline: the loads of row t + 1 are issued before row t is processed
  // All loads of a row share one scoreboard, and a wait on it waits for everything in flight: the values of
  // row t must therefore be waited for BEFORE the loads of row t + 1 are issued, or the prefetch is void
  // (measured: 1.2 long-scoreboard stalls per issued instruction).  `dep` (always 0, unknown to the compiler)
  // makes the new addresses depend on a value of the previous batch.
  real yn[3], xn[CT][3], dn[PSFM_MAX_SCALES] = {real(0), real(0), real(0), real(0)}, idn = real_inf<real>();
  const int never = H >> 31;
  auto load_row = [&](int t, int dep) {
    const int o = reflect1(t, H) * W + su + dep;
    if (AUTOMASK && !FIRST) idn = rb[rec_ofs(min(max(t - 1, 0), H - 1), su, Wh) + dep].w;
    yn[0] = ldimg<real>(y0p, o); yn[1] = ldimg<real>(y1p, o); yn[2] = ldimg<real>(y2p, o);
#pragma unroll
    for (int jj = 0; jj < CT; ++jj) {
      xn[jj][0] = ldimg<real>(xp0[jj], o); xn[jj][1] = ldimg<real>(xp0[jj] + plane, o);
      xn[jj][2] = ldimg<real>(xp0[jj] + 2 * plane, o);
    }
    if (smooth) {
#pragma unroll
      for (int k = 0; k < PSFM_MAX_SCALES; ++k)
        if (k < ns) {
          int ok = o;
          if (any_ups) {
            const ScaleDesc& Sk = P.sc[S.scale[k]];
            ok = nearest_src(reflect1(t, H), Sk.rs, Sk.Hd, Sk.Hd != H) * Sk.Wd + dcol[k] + dep;
          }
          dn[k] = __ldg(dptr[k] + ok);
        }
    }
  };
  load_row(r0 - 1, 0);
  // two steps per trip: the 2-row rings are indexed by the step's parity, so nothing is rotated
  auto step = [&](const int t, auto phc) {
    constexpr int PH = decltype(phc)::value;      // slot of row t - 2 (overwritten by row t), 1 - PH: row t - 1
    const int p = t - 1;
    const bool row_ok = (p >= r0) && (p < r1);
    const bool emit = inner && row_ok;
    const int op = p * W + u;
    const pr yA(yn[0], yn[1]);
    const real yB = yn[2];
    pr xA[CT];
    real xB[CT], dcur[PSFM_MAX_SCALES];
#pragma unroll
    for (int jj = 0; jj < CT; ++jj) { xA[jj] = pr(xn[jj][0], xn[jj][1]); xB[jj] = xn[jj][2]; }
#pragma unroll
    for (int k = 0; k < PSFM_MAX_SCALES; ++k) dcur[k] = dn[k];
    real idv = (AUTOMASK && !FIRST) ? idn : real_inf<real>(), idk = real(0);
    load_row(t + 1, value_bits(yn[0]) & never);
    // horizontal sums of row t
    const pr yaA = shfl_up1(yA), ycA = shfl_dn1(yA);
    const real yaB = shfl_up1(yB), ycB = shfl_dn1(yB);
    const pr hynA0 = (yaA + yA) + ycA, hynA1 = vfma(ycA, ycA, vfma(yA, yA, yaA * yaA));
    const real hynB0 = (yaB + yB) + ycB, hynB1 = vfma(ycB, ycB, vfma(yB, yB, yaB * yaB));
    pr hxnA[CT][3];
    real hxnB[CT][3];
    if (AUTOMASK) {
#pragma unroll
      for (int jj = 0; jj < CT; ++jj) {
        const pr zA = xA[jj] * yA;
        const real zB = xB[jj] * yB;
        const pr xaA = shfl_up1(xA[jj]), xcA = shfl_dn1(xA[jj]), zaA = shfl_up1(zA), zcA = shfl_dn1(zA);
        const real xaB = shfl_up1(xB[jj]), xcB = shfl_dn1(xB[jj]), zaB = shfl_up1(zB), zcB = shfl_dn1(zB);
        hxnA[jj][0] = (xaA + xA[jj]) + xcA;
        hxnA[jj][1] = vfma(xcA, xcA, vfma(xA[jj], xA[jj], xaA * xaA));
        hxnA[jj][2] = (zaA + zA) + zcA;
        hxnB[jj][0] = (xaB + xB[jj]) + xcB;
        hxnB[jj][1] = vfma(xcB, xcB, vfma(xB[jj], xB[jj], xaB * xaB));
        hxnB[jj][2] = (zaB + zB) + zcB;
      }
    }
    real wx = real(0), wy = real(0);
    if (row_ok) {
      // window statistics of the target at row p (reference :15-54): mean, variance + C2
      const pr myA = ((hyA[PH][0] + hyA[1 - PH][0]) + hynA0) * pr(r9);
      const real myB = ((hyB[PH][0] + hyB[1 - PH][0]) + hynB0) * r9;
      const pr qyA = vfma((hyA[PH][1] + hyA[1 - PH][1]) + hynA1, pr(r9), pr(C2)) - myA * myA;
      const real qyB = vfma((hyB[PH][1] + hyB[1 - PH][1]) + hynB1, r9, C2) - myB * myB;
      if (AUTOMASK) {
        bool changed = false;
#pragma unroll
        for (int jj = 0; jj < CT; ++jj) {
          SsimSt<pr> stA;
          SsimSt<real> stB;
          const pr dA = ssim_fwd_v<real, pr>((hxA[jj][PH][0] + hxA[jj][1 - PH][0]) + hxnA[jj][0],
                                             (hxA[jj][PH][1] + hxA[jj][1 - PH][1]) + hxnA[jj][1],
                                             (hxA[jj][PH][2] + hxA[jj][1 - PH][2]) + hxnA[jj][2], myA, qyA, C1, C2, stA);
          const real dB = ssim_fwd_v<real, real>((hxB[jj][PH][0] + hxB[jj][1 - PH][0]) + hxnB[jj][0],
                                                 (hxB[jj][PH][1] + hxB[jj][1 - PH][1]) + hxnB[jj][1],
                                                 (hxB[jj][PH][2] + hxB[jj][1 - PH][2]) + hxnB[jj][2], myB, qyB, C1, C2, stB);
          const real ds = vclamp01_sum(dA) + vclamp01_sum(dB);
          const pr lA = xpA[jj] - ypA;
          const real ls = (rabs(lA.lo()) + rabs(lA.hi())) + rabs(xpB[jj] - ypB);
          const real val = kd * ds + kl * ls;
          // candidate order [.., id_j, ..]: the first minimum wins (reference :269-297, torch.min)
          if ((first && jj == 0) || val < idv) { idv = val; idk = (real)(2 * (j0 + jj) + 1); changed = true; }
        }
        const int orc = rec_ofs(p, u, Wh);
        if (!first && changed && emit) { rb[orc].w = idv; rc[orc].w = idk; }
      }
      if (emit) {
#pragma unroll
        for (int jj = 0; jj < CT; ++jj) st4(cpk[jj] + p * Wp + u, xpA[jj].lo(), xpA[jj].hi(), xpB[jj], real(0));
        if (first) {
          const int orc = rec_ofs(p, u, Wh);
          st4(ra + orc, ypA.lo(), ypA.hi(), ypB, maskp ? ldmask<real>(maskp, op) : real(1));
          st4(rb + orc, myA.lo(), myA.hi(), myB, idv);
          st4(rc + orc, qyA.lo(), qyA.hi(), qyB, idk);
        }
        if (smooth) {
          const pr ex = ypA - ycpA, ey = ypA - yA;
          if (u + 1 < W) wx = rexp(-(((rabs(ex.lo()) + rabs(ex.hi())) + rabs(ypB - ycpB)) * r3));
          if (p + 1 < H) wy = rexp(-(((rabs(ey.lo()) + rabs(ey.hi())) + rabs(ypB - yB)) * r3));
          wxy[(size_t)op * 2] = wx;
          wxy[(size_t)op * 2 + 1] = wy;
        }
      }
    }
    if (smooth) {
      // raw smoothness sums of row p for every scale of the set (all lanes take part in the
      // right-neighbour shuffle)
#pragma unroll
      for (int k = 0; k < PSFM_MAX_SCALES; ++k) {
        if (k < ns) {
          const real dt = dcur[k];
          const real dr = shfl_dn1(dp[k]);
          if (emit) {
            sums[k][0] += dp[k];
            if (u + 1 < W) sums[k][1] += rabs(dp[k] - dr) * wx;
            if (p + 1 < H) sums[k][2] += rabs(dp[k] - dt) * wy;
          }
          dp[k] = dt;
        }
      }
    }
    // row t replaces row t - 2 in the rings
    hyA[PH][0] = hynA0; hyA[PH][1] = hynA1; hyB[PH][0] = hynB0; hyB[PH][1] = hynB1;
    if (AUTOMASK) {
#pragma unroll
      for (int jj = 0; jj < CT; ++jj)
#pragma unroll
        for (int m = 0; m < 3; ++m) { hxA[jj][PH][m] = hxnA[jj][m]; hxB[jj][PH][m] = hxnB[jj][m]; }
    }
    ypA = yA; ypB = yB; ycpA = ycA; ycpB = ycB;
#pragma unroll
    for (int jj = 0; jj < CT; ++jj) { xpA[jj] = xA[jj]; xpB[jj] = xB[jj]; }
  };
  for (int t = r0 - 1; t <= r1; t += 2) {
    step(t, IntC<0>());
    step(t + 1, IntC<1>());
  }
  if (smooth) {
#pragma unroll
    for (int k = 0; k < PSFM_MAX_SCALES; ++k) {
      if (k < S.ns) {
#pragma unroll
        for (int i = 0; i < kPrepVals; ++i) {
          const real v = warp_allsum<real>(sums[k][i]);
          if (lane == 0) SP.prep_part[S.scale[k]][((size_t)b * punits + unit) * kPrepVals + i] = (double)v;
        }
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
__global__ void __launch_bounds__(128) stats_kernel(const SweepParams SP) {
  const LossParams& P = SP.P;
  const SweepSet& S = SP.set;
  const int b = blockIdx.x, s = S.scale[blockIdx.y];
  const int punits = S.pstrips * S.pchunks;
  __shared__ double sh[3][4];
  double a[3] = {0.0, 0.0, 0.0};
  const double* base = SP.prep_part[s] + (size_t)b * punits * kPrepVals;
  for (int i = threadIdx.x; i < punits; i += blockDim.x)
    for (int k = 0; k < 3; ++k) a[k] += base[(size_t)i * kPrepVals + k];
  // fixed-order reduction: butterfly inside each warp, then the four warp sums in order
  for (int k = 0; k < 3; ++k) {
    a[k] = warp_allsum<double>(a[k]);
    if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = a[k];
  }
  __syncthreads();
  if (threadIdx.x == 0)
    for (int k = 0; k < 3; ++k) sh[k][0] = ((sh[k][0] + sh[k][1]) + sh[k][2]) + sh[k][3];
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
// fisheye ray kernel: grid (ceil(H W / 256), B).  FisheyeCamera.reconstruct, frame 'c' (geometry/camera.py:242-305) up
// to the depth factor: x_d = (u - ux) / s, y_d = (v - uy) / div, theta = r_d (the reference's placeholder inverse of the
// distortion polynomial), ray = (tan(r_d) / r_d) (x_d, y_d), 1.  Depth-independent, shared by the scales of a set and
// by all contexts, so it is evaluated once per pixel here instead of per (scale, context) in the sweep.
// ------------------------------------------------------------------------------------------
template <typename real>
__global__ void __launch_bounds__(256) fisheye_ray_kernel(const SweepParams SP) {
  const LossParams& P = SP.P;
  const SweepSet& S = SP.set;
  const int b = blockIdx.y, H = S.H, W = S.W;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= H * W) return;
  const int v = i / W, u = i - v * W;
  const real eps = (real)2.220446049250313e-16;                 // sys.float_info.epsilon (camera.py:287)
  // per-scale intrinsics as warp_ref_image builds them: only the principal point is rescaled
  // (losses/multiview_photometric_loss.py:158-183)
  const real sw = (real)P.sc[S.scale[0]].sf, sh = (real)((double)H / (double)P.H);
  const real* ti = (const real*)P.K + (size_t)b * 11;
  const real xd = ((real)u - ((ti[9] + real(0.5)) * sw - real(0.5))) / ti[7];
  const real yd = ((real)v - ((ti[10] + real(0.5)) * sh - real(0.5))) / ti[8];
  const real rd = rroot(xd * xd + yd * yd);
  const real sc = rtan(rd) / (rd < eps ? eps : rd);
  real* out = (real*)S.ray + ((size_t)b * H * W + i) * 2;
  out[0] = sc * xd;
  out[1] = sc * yd;
}

// ------------------------------------------------------------------------------------------
// sweep kernel: grid (nstrips * nchunks, ns, B), 32 * C threads (warp j = context j).
// Lane l is column c0 - 2 + l (2 halo columns per side: SSIM needs +-1, its adjoint +-1 more).
// Step t finishes the gather of row t (issued one step earlier), starts the gather of row t + 1,
// evaluates SSIM / the candidate min at row p = t - 1 and the adjoint (d/d inv_depth, d/d pose)
// at row q = t - 2.
// ------------------------------------------------------------------------------------------

#ifndef PSFM_SWEEP_MINB
#define PSFM_SWEEP_MINB 1            // resident CTAs per SM the 2-context fp32 instantiation is compiled for
#endif
// VAR: 0 = plain, 1 = inverse depths upsampled on load (all scales of the set), 2 = rare settings (fork pre-map;
// upsampling decided per scale at run time)
// CAM: 0 = pinhole cameras (K, ref_K 3x3), 1 = fisheye (VADAS) cameras (geometry/camera.py:194-394): K, ref_K are
// rows of 11 values (k0..k6, s, div, ux, uy)
// DBG: test instantiation that also records the discrete decisions (io->dbg decision words, io->sel arg-min maps)
template <typename real, int PAD, int VAR, int CT, int CAM = 0, bool DBG = false>
#ifdef PSFM_SWEEP_MAXREG
__global__ void __maxnreg__(PSFM_SWEEP_MAXREG) sweep_kernel(const SweepParams SP) {
#else
__global__ void __launch_bounds__(CT ? 32 * CT : 256, (CT == 2 && sizeof(real) == 4) ? PSFM_SWEEP_MINB : 1) sweep_kernel(const SweepParams SP) {
#endif
  typedef Real4<real> vec4;
  constexpr int kRing = CAM ? 15 : ((PAD == PSFM_PAD_BORDER) ? 11 : 10);     // per-row values parked in shared memory
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const LossParams& P = SP.P;
  const SweepSet& S = SP.set;
  const SweepConsts<real>& K = sweep_consts<real>(SP);
  const int C = CT ? CT : P.C;
  const int lane = threadIdx.x & 31, j = threadIdx.x >> 5;
  const int unit = blockIdx.x, nunits = gridDim.x;
  const int strip = unit % S.nstrips, chunk = unit / S.nstrips;
  const int si = blockIdx.y, s = S.scale[si], b = blockIdx.z;
  const int H = S.H, W = S.W, plane = H * W;
  const int u = strip * kInner - 2 + lane, su = reflect1(u, W);
  const bool col_real = (u >= 0) && (u < W);
  const bool col_inner = (lane >= 2) && (lane < 30) && col_real;
  const int r0 = chunk * S.chunk_h, r1 = min(r0 + S.chunk_h, H);

  real* ring = (real*)smem_raw + (size_t)j * 4 * kRing * 32 + lane;        // [4 slots][kRing][32]
  real* photo_s = (real*)smem_raw + (size_t)C * 4 * kRing * 32;            // [2][C][32]
  real* dpart_s = photo_s + 2 * C * 32;                                    // [2][C][32]
  real* fc = dpart_s + 2 * C * 32 + j * 24;                                // fisheye: [C][24] per-context constants

  const ScaleDesc& SD = P.sc[s];
  const vec4* ctxp = (const vec4*)S.ctxp[j] + (size_t)b * S.Hp * S.Wp;
  const int Wh = S.Wh, rplane = H * 2 * Wh * 3;
  const vec4* ra = (const vec4*)S.ra + (size_t)b * rplane;
  const vec4* rb = (const vec4*)S.rb + (size_t)b * rplane;
  const vec4* rcp = (const vec4*)S.rc + (size_t)b * rplane;
  asm volatile("" : "+l"(ctxp), "+l"(ra), "+l"(rb), "+l"(rcp));   // opaque: one IMAD.WIDE per load off a 32-bit offset
  const bool ups = VAR == 1 || (VAR == 2 && (SD.Hd != H || SD.Wd != W));   // inverse depths nearest-upsampled on load
  const real* invd = (const real*)SD.invd + (size_t)b * SD.Hd * SD.Wd;
  const int Wd = SD.Wd, sud = nearest_src(su, SD.cs, SD.Wd, ups);
  auto dofs = [&](int row) { return nearest_src(reflect1(row, H), SD.rs, SD.Hd, ups) * Wd + sud; };
  const real* wxy = (const real*)S.wxy + (size_t)b * plane * 2;
  const int Wp = S.Wp;
  const bool premap = VAR == 2 && P.premap;
  const bool smooth = P.smooth_w > 0.0;

  // ---- per-thread geometry constants: p = (A0 + A1 v) D + m with A = ref_K R Kinv (u, v, 1) ---------
  real A0[3], A1[3], mt[3];
  const real* rayp = (const real*)S.ray + (size_t)b * plane * 2;   // fisheye: rays of the target camera (fisheye_ray_kernel)
  real rxc = real(0), ryc = real(0), rxn = real(0), ryn = real(0);  // ray of the row being started / prefetched
  if (CAM) {
    // per-scale intrinsics of the context camera as warp_ref_image builds them: only the principal point is rescaled
    // (losses/multiview_photometric_loss.py:158-183)
    const real sw = K.sf[s], sh = (real)((double)H / (double)P.H);
    const real* ri = (const real*)P.refK + (size_t)b * 11;
    const real* T = (const real*)P.pose[j] + (size_t)b * 16;
    if (lane < 12) fc[lane] = T[lane];
    else if (lane < 21) fc[lane] = ri[lane - 12];                       // k0..k6, s, div of the context camera
    else if (lane == 21) fc[21] = (ri[9] + real(0.5)) * sw - real(0.5);
    else if (lane == 22) fc[22] = (ri[10] + real(0.5)) * sh - real(0.5);
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 3; ++i) { A0[i] = real(0); A1[i] = real(0); mt[i] = real(0); }
  } else {
    SweepCam<real> cam;
    sweep_cam<real>(cam, P, b, j, K.sf[s]);
    real M[9];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
        M[i * 3 + c] = cam.rk[i * 3] * cam.T[c] + cam.rk[i * 3 + 1] * cam.T[4 + c] + cam.rk[i * 3 + 2] * cam.T[8 + c];
      mt[i] = cam.rk[i * 3] * cam.T[3] + cam.rk[i * 3 + 1] * cam.T[7] + cam.rk[i * 3 + 2] * cam.T[11];
    }
    const real fu = (real)su;
    const real b0 = cam.ki[0] * fu + cam.ki[2], b1 = cam.ki[3] * fu + cam.ki[5], b2 = cam.ki[6] * fu + cam.ki[8];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      A0[i] = M[i * 3] * b0 + M[i * 3 + 1] * b1 + M[i * 3 + 2] * b2;
      A1[i] = M[i * 3] * cam.ki[1] + M[i * 3 + 1] * cam.ki[4] + M[i * 3 + 2] * cam.ki[7];
    }
  }
  const real mu0 = (u == 1) ? real(2) : real(1), mu2 = (u == W - 2) ? real(2) : real(1);
  const int mycand = P.automask ? 2 * j : j;
  real inv_m = real(1), cb = real(0);
  if (smooth) {
    const real* st = (const real*)SP.stats[s] + (size_t)b * 4;
    inv_m = real(1) / st[0];
    cb = st[1];
  }

  // window state, by row phase; per-channel values live as a (channel 0, channel 1) pair + a scalar channel 2
  // (pair_math.cuh); index m = 0: sums of x, 1: of x^2, 2: of x y (h) / their gradient coefficients (hF)
  typedef P2<real> pr;
  pr hA[3][3], hFA[3][3], xrA[3], yrA[3];
  real hB[3][3], hFB[3][3], xrB[3], yrB[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int m = 0; m < 3; ++m) { hA[a][m] = pr(real(0)); hFA[a][m] = pr(real(0)); hB[a][m] = real(0); hFB[a][m] = real(0); }
    xrA[a] = pr(real(0)); yrA[a] = pr(real(0)); xrB[a] = real(0); yrB[a] = real(0);
  }
  const pr mu0P(mu0), mu2P(mu2);
#pragma unroll
  for (int k = 0; k < 4 * kRing; ++k) ring[k * 32] = real(0);
  real S0a[3] = {real(0), real(0), real(0)}, Sva[3] = {real(0), real(0), real(0)}, S1a[3] = {real(0), real(0), real(0)};
  real Sua[3] = {real(0), real(0), real(0)};                       // fisheye only
  real loss_acc = real(0), w_cur = real(0), w_old = real(0);
  real d0 = real(0), d1 = real(0), d2 = real(0), d3 = real(0);   // raw inverse depths of rows t, t-1, t-2, t-3
  real wy_prev = real(0);                                          // edge weight wy of row q - 1
  real dsum_prev = real(0);
  int prev_q = -1;
  real* d_out = (real*)SD.d_full + (size_t)b * plane;

  // debug instantiation: floor of the sampling position after the padding mode's clipping (include/psfm.h, io->dbg)
  auto record_decision = [&](int row, real ix, real iy) {
    if (!SD.dbg || row < r0 || row >= r1 || !col_inner) return;
    if (PAD == PSFM_PAD_BORDER) {
      ix = !(ix > real(0)) ? real(0) : rmin(ix, real(W - 1));
      iy = !(iy > real(0)) ? real(0) : rmin(iy, real(H - 1));
    } else {
      ix = rmin(rmax(ix, real(-2)), real(W + 1));
      iy = rmin(rmax(iy, real(-2)), real(H + 1));
    }
    const int x0 = (int)rfloor(ix), y0 = (int)rfloor(iy);
    SD.dbg[((size_t)j * P.B + b) * plane + (size_t)row * W + u] =
        (uint32_t)min(max(x0 + 8, 0), 4095) | ((uint32_t)min(max(y0 + 8, 0), 2047) << 12);
  };
  // gather pipeline: geometry of `row`, the 4 tap loads go in flight, consumed one step later
  vec4 tap[4];
  real twx = real(0), twy = real(0);
  auto start_row = [&](int row, real draw) {
    real dd = draw;
    if (premap) { real ch; dd = premap_fork<real>(draw, K.premap_lo, K.premap_hi, ch); }
    const real D = rrcp(rmax(dd, real(1e-6)));                    // utils/depth.py:120
    const real fv = (real)reflect1(row, H);
    real* rs = ring + (row & 3) * kRing * 32;
    if (CAM) {
      const real eps = (real)2.220446049250313e-16;               // sys.float_info.epsilon (camera.py:287,355,373)
      const real rx = rxc, ry = ryc;                               // FisheyeCamera.reconstruct: fisheye_ray_kernel
      // context camera frame (pose.py:80-86)
      const real X0 = (fc[0] * rx + fc[1] * ry + fc[2]) * D + fc[3];
      const real X1 = (fc[4] * rx + fc[5] * ry + fc[6]) * D + fc[7];
      const real X2 = (fc[8] * rx + fc[9] * ry + fc[10]) * D + fc[11];
      // FisheyeCamera.project, frame 'c' (camera.py:352-383)
      const real iz = rrcp(rmax(X2, eps));
      const real xn = X0 * iz, yn = X1 * iz;
      const real r2 = xn * xn + yn * yn, r = rroot(r2);
      const real th = ratan(r);
      real poly = fc[18], dpoly = real(6) * fc[18];
#pragma unroll
      for (int i = 5; i >= 1; --i) { poly = poly * th + fc[12 + i]; dpoly = dpoly * th + (real)i * fc[12 + i]; }
      poly = poly * th + fc[12];
      const bool tiny = r < eps;
      const real ir = rrcp(tiny ? eps : r);
      const real g = poly * ir;
      const real gp = dpoly * rrcp(real(1) + r2);
      const real h = tiny ? real(0) : (gp - g) * ir * ir;         // (d g / d r) / r
      const real ix = fc[19] * (g * xn) + fc[21], iy = fc[20] * (g * yn) + fc[22];
      real mx, my;
      sample4_issue<real, vec4, PAD>(ctxp, H, W, Wp, kPadLo, ix, iy, twx, twy, mx, my, tap);
      if (DBG) record_decision(row, ix, iy);
      // d (ix, iy) / d X, rows scaled by the border-clamp masks
      const real hxy = h * xn * yn;
      real a00 = fc[19] * (g + h * xn * xn), a01 = fc[19] * hxy, a10 = fc[20] * hxy, a11 = fc[20] * (g + h * yn * yn);
      if (PAD == PSFM_PAD_BORDER) { a00 *= mx; a01 *= mx; a10 *= my; a11 *= my; }
      const real zk = (X2 >= eps) ? -iz : real(0);
      rs[6 * 32] = D;
      rs[7 * 32] = rx;
      rs[8 * 32] = ry;
      rs[9 * 32] = a00 * iz; rs[10 * 32] = a01 * iz; rs[11 * 32] = (a00 * xn + a01 * yn) * zk;
      rs[12 * 32] = a10 * iz; rs[13 * 32] = a11 * iz; rs[14 * 32] = (a10 * xn + a11 * yn) * zk;
      return;
    }
    const real p0 = (A0[0] + A1[0] * fv) * D + mt[0];
    const real p1 = (A0[1] + A1[1] * fv) * D + mt[1];
    const real p2 = (A0[2] + A1[2] * fv) * D + mt[2];
    const real iz = rrcp(rmax(p2, real(1e-5)));                   // camera.py:179
    const real ix = p0 * iz, iy = p1 * iz;
    real mx, my;
    sample4_issue<real, vec4, PAD>(ctxp, H, W, Wp, kPadLo, ix, iy, twx, twy, mx, my, tap);
    if (DBG) record_decision(row, ix, iy);
    rs[6 * 32] = (p2 >= real(1e-5)) ? D : -D;                     // sign flags the clamp of p2
    rs[7 * 32] = ix;
    rs[8 * 32] = iy;
    if (PAD == PSFM_PAD_BORDER) { rs[9 * 32] = iz * mx; rs[10 * 32] = iz * my; }
    else rs[9 * 32] = iz;
  };
  // Consumes the taps in the same step that issued them (at its end): only the sample and its gradients
  // are carried into the next step.  (Carrying the tap vectors across the loop edge made ptxas copy them
  // right behind the loads in one of the three unrolled phases -- a full memory latency per three rows.)
  pr xcarA(real(0));
  real xcarB = real(0);
  auto finish_row = [&](int row) {
    // channels 0, 1 as a pair; channel 2 rides with the (always zero) 4th component through the first two
    // subtractions, which also keeps that component's register allocated until the load has landed (a dead
    // component is re-used at once, and that write-after-write then waits on the loads' scoreboard: measured
    // 1.3 long-scoreboard stalls per issued instruction)
    pr gxA, gyA;
    bilerp_v<pr>(pr(tap[0].x, tap[0].y), pr(tap[1].x, tap[1].y), pr(tap[2].x, tap[2].y), pr(tap[3].x, tap[3].y), pr(twx),
                 pr(twy), xcarA, gxA, gyA);
    const pr dt = pr(tap[1].z, tap[1].w) - pr(tap[0].z, tap[0].w), db = pr(tap[3].z, tap[3].w) - pr(tap[2].z, tap[2].w);
    const real dxt = dt.lo() + dt.hi(), dxb = db.lo() + db.hi();     // the high lanes are 0 - 0: consumed on purpose
    const real top = vfma(dxt, twx, tap[0].z), bot = vfma(dxb, twx, tap[2].z);
    const real gyB = bot - top;
    xcarB = vfma(gyB, twy, top);
    const real gxB = vfma(dxb - dxt, twy, dxt);
    real* rs = ring + (row & 3) * kRing * 32;
    rs[0 * 32] = gxA.lo(); rs[1 * 32] = gxA.hi(); rs[2 * 32] = gxB;
    rs[3 * 32] = gyA.lo(); rs[4 * 32] = gyA.hi(); rs[5 * 32] = gyB;
  };
  // prefetched operands of the next steps: target pixel (+ mask) record, statistics records
  vec4 ran, rbn, rcn;
  real mcur = real(1);                                             // mask value of row t - 1
  real wxn = real(0), wyn = real(0);                               // edge weights of row q + 1 (context 0 only)
  if (j == 0 && smooth) load_pair(wxy + (min(max(r0 - 4, 0), H - 1) * W + su) * 2, wxn, wyn);
  real dnx = __ldg(invd + dofs(r0 - 2));
  real dpre = __ldg(invd + dofs(r0 - 1));
  if (CAM) {
    load_pair(rayp + (reflect1(r0 - 2, H) * W + su) * 2, rxc, ryc);
    load_pair(rayp + (reflect1(r0 - 1, H) * W + su) * 2, rxn, ryn);
  }
  {
    ran = ld4(ra + rec_ofs(reflect1(r0 - 2, H), su, Wh));
    const int orec = rec_ofs(min(max(r0 - 3, 0), H - 1), su, Wh);
    rbn = ld4(rb + orec);
    rcn = ld4(rcp + orec);
  }
  start_row(r0 - 2, dnx);
  finish_row(r0 - 2);

  auto finalize_row = [&](int par) {      // context 0 adds the other contexts' parts and stores the row
    if (j == 0 && prev_q >= 0) {
      real tot = dsum_prev;
#pragma unroll
      for (int jj = 1; jj < 8; ++jj)
        if (jj < C) tot += dpart_s[(par * C + jj) * 32 + lane];
      d_out[prev_q * W + u] = tot;
    }
    prev_q = -1;
  };

  const int t0 = r0 - 2;
  auto step = [&](const int t, auto phc) {
    constexpr int PH = decltype(phc)::value, Q = (PH + 1) % 3, P1 = (PH + 2) % 3;
    const int p = t - 1, q = t - 2;
    const int par = t & 1;
    // ---------------- operands prefetched one step ago; issue the prefetch of the next step -----------
    const bool p_real = (p >= 0) && (p < H) && col_real;
    real wxq = real(0), wyq = real(0), m = real(1);
    const pr yA(ran.x, ran.y), myA(rbn.x, rbn.y), qyA(rcn.x, rcn.y);
    const real yB = ran.z, myB = rbn.z, qyB = rcn.z, idmin = rbn.w, idcand = rcn.w;
    {
      m = p_real ? mcur : real(0);                                 // 1 without a mask (every record component is consumed)
      mcur = ran.w;
      d3 = d2; d2 = d1; d1 = d0;
      d0 = dnx;
      dnx = dpre;
      wxq = wxn; wyq = wyn;
      if (CAM) { rxc = rxn; ryc = ryn; }
    }
    // ---------------- prefetch of the next steps' operands, gather of row t + 1 ----------------------------
    {
      dpre = __ldg(invd + dofs(t + 2));
      ran = ld4(ra + rec_ofs(reflect1(t + 1, H), su, Wh));
      const int orec = rec_ofs(min(max(p + 1, 0), H - 1), su, Wh);
      rbn = ld4(rb + orec);
      rcn = ld4(rcp + orec);
      if (j == 0 && smooth) load_pair(wxy + (min(max(q + 1, 0), H - 1) * W + su) * 2, wxn, wyn);
      if (CAM) load_pair(rayp + (reflect1(t + 2, H) * W + su) * 2, rxn, ryn);
      start_row(t + 1, dnx);
    }
    // ---------------- horizontal window sums of row t ------------------------------------------------------
    {
      const pr xA = xcarA, zA = xA * yA;
      const real xB = xcarB, zB = xB * yB;
      xrA[PH] = xA; yrA[PH] = yA; xrB[PH] = xB; yrB[PH] = yB;
      const pr xaA = shfl_up1(xA), xcA = shfl_dn1(xA), zaA = shfl_up1(zA), zcA = shfl_dn1(zA);
      const real xaB = shfl_up1(xB), xcB = shfl_dn1(xB), zaB = shfl_up1(zB), zcB = shfl_dn1(zB);
      hA[PH][0] = (xaA + xA) + xcA;
      hA[PH][1] = vfma(xcA, xcA, vfma(xA, xA, xaA * xaA));
      hA[PH][2] = (zaA + zA) + zcA;
      hB[PH][0] = (xaB + xB) + xcB;
      hB[PH][1] = vfma(xcB, xcB, vfma(xB, xB, xaB * xaB));
      hB[PH][2] = (zaB + zB) + zcB;
    }
    // ---------------- SSIM + L1 at row p, candidate min across the contexts --------------------------
    pr FA[3];
    real FB[3];
    {
      SsimSt<pr> stA;
      SsimSt<real> stB;
      const pr dA = ssim_fwd_v<real, pr>((hA[0][0] + hA[1][0]) + hA[2][0], (hA[0][1] + hA[1][1]) + hA[2][1],
                                         (hA[0][2] + hA[1][2]) + hA[2][2], myA, qyA, K.C1, K.C2, stA);
      const real dB = ssim_fwd_v<real, real>((hB[0][0] + hB[1][0]) + hB[2][0], (hB[0][1] + hB[1][1]) + hB[2][1],
                                             (hB[0][2] + hB[1][2]) + hB[2][2], myB, qyB, K.C1, K.C2, stB);
      const real ds = vclamp01_sum(dA) + vclamp01_sum(dB);
      const pr lA = xrA[P1] - yrA[P1];
      const real ls = (rabs(lA.lo()) + rabs(lA.hi())) + rabs(xrB[P1] - yrB[P1]);
      const real photo = K.ssim_w3 * ds + K.k1 * ls;
      if (DBG) {       // L1 signs (bits 23-28) and SSIM clamp-range flags (bits 29-31) of this context at row p
        if (SD.dbg && p >= r0 && p < r1 && p_real && col_inner) {
          const real eB = xrB[P1] - yrB[P1];
          auto in01 = [](real d) { return d >= real(0) && d <= real(1); };
          auto sg = [](real e) { return e > real(0) ? 2u : (e < real(0) ? 0u : 1u); };
          const uint32_t bits = sg(lA.lo()) | (sg(lA.hi()) << 2) | (sg(eB) << 4) |
                                (in01(dA.lo()) ? 64u : 0u) | (in01(dA.hi()) ? 128u : 0u) | (in01(dB) ? 256u : 0u);
          SD.dbg[((size_t)j * P.B + b) * plane + (size_t)p * W + u] |= bits << 23;
        }
      }
      real best = photo * m;
      int sel = mycand;
      if (C > 1) {
        // the contexts meet once per row (a split arrive / wait on an mbarrier measured slower than bar.sync)
        photo_s[(par * C + j) * 32 + lane] = photo;
        __syncthreads();
        best = real_inf<real>();
        sel = 0;
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
          if (jj < C) {
            const real v = photo_s[(par * C + jj) * 32 + lane] * m;
            if (v < best) { best = v; sel = P.automask ? 2 * jj : jj; }
          }
        }
      }
      if (p_real) {                                                // without automask the record holds +inf
        const real vi = idmin * m;
        const int ki = (int)idcand;
        if (vi < best || (vi == best && ki < sel)) { best = vi; sel = ki; }
      }
      w_old = w_cur;
      w_cur = (p_real && sel == mycand) ? K.inv_count * m : real(0);
      if (j == 0 && p_real && col_inner && p >= r0 && p < r1) loss_acc += best;
      if (DBG) {
        if (j == 0 && p_real && col_inner && p >= r0 && p < r1 && SD.sel) SD.sel[(size_t)b * plane + (size_t)p * W + u] = (uint8_t)sel;
      }
      const real wf = w_cur * K.ssim_w3;
      ssim_grad_v<real, pr>(stA, myA, wf, FA);
      ssim_grad_v<real, real>(stB, myB, wf, FB);
    }
    finalize_row(par ^ 1);
    // ---------------- adjoint at row q ----------------------------------------------------------------
    {
#pragma unroll
      for (int m = 0; m < 3; ++m) {
        hFA[P1][m] = vfma(mu2P, shfl_dn1(FA[m]), vfma(mu0P, shfl_up1(FA[m]), FA[m]));
        hFB[P1][m] = vfma(mu2, shfl_dn1(FB[m]), vfma(mu0, shfl_up1(FB[m]), FB[m]));
      }
      const bool q_ok = (q >= r0) && (q < r1) && col_inner;
      const real mv0 = (q == 1) ? real(2) : real(1), mv2 = (q == H - 2) ? real(2) : real(1);
      const pr mv0P(mv0), mv2P(mv2);
      const real* rq = ring + (q & 3) * kRing * 32;
      const real wk = w_old * K.k1;
      pr sA[3];
      real sB[3];
#pragma unroll
      for (int m = 0; m < 3; ++m) {
        sA[m] = vfma(mv2P, hFA[P1][m], vfma(mv0P, hFA[PH][m], hFA[Q][m]));
        sB[m] = vfma(mv2, hFB[P1][m], vfma(mv0, hFB[PH][m], hFB[Q][m]));
      }
      const pr xqA = xrA[Q], yqA = yrA[Q], eA = xqA - yqA;
      const real xqB = xrB[Q], yqB = yrB[Q];
      const pr dlA = vfma(yqA, sA[2], vfma(xqA, sA[1], sA[0])) + pr(signed_by(wk, eA.lo()), signed_by(wk, eA.hi()));
      const real dlB = vfma(yqB, sB[2], vfma(xqB, sB[1], sB[0])) + signed_by(wk, xqB - yqB);
      real gix = dlA.lo() * rq[0 * 32] + dlA.hi() * rq[1 * 32] + dlB * rq[2 * 32];
      real giy = dlA.lo() * rq[3 * 32] + dlA.hi() * rq[4 * 32] + dlB * rq[5 * 32];
      gix = q_ok ? gix : real(0);
      giy = q_ok ? giy : real(0);
      real D, dD, dp0, dp1, dp2;
      const real fq = (real)q;
      if (CAM) {
        D = rq[6 * 32];
        const real rx = rq[7 * 32], ry = rq[8 * 32];
        dp0 = gix * rq[9 * 32] + giy * rq[12 * 32];
        dp1 = gix * rq[10 * 32] + giy * rq[13 * 32];
        dp2 = gix * rq[11 * 32] + giy * rq[14 * 32];
        dD = (fc[0] * rx + fc[1] * ry + fc[2]) * dp0 + (fc[4] * rx + fc[5] * ry + fc[6]) * dp1 +
             (fc[8] * rx + fc[9] * ry + fc[10]) * dp2;
        const real e0 = dp0 * D, e1 = dp1 * D, e2 = dp2 * D;
        Sua[0] += e0 * rx; Sua[1] += e1 * rx; Sua[2] += e2 * rx;
        Sva[0] += e0 * ry; Sva[1] += e1 * ry; Sva[2] += e2 * ry;
        S0a[0] += e0; S0a[1] += e1; S0a[2] += e2;
        S1a[0] += dp0; S1a[1] += dp1; S1a[2] += dp2;
      } else {
        const real Ds = rq[6 * 32], ixq = rq[7 * 32], iyq = rq[8 * 32];
        D = rabs(Ds);
        dp0 = gix * rq[9 * 32];
        dp1 = giy * rq[(PAD == PSFM_PAD_BORDER ? 10 : 9) * 32];
        dp2 = (Ds > real(0)) ? -(dp0 * ixq + dp1 * iyq) : real(0);
        dD = (A0[0] + A1[0] * fq) * dp0 + (A0[1] + A1[1] * fq) * dp1 + (A0[2] + A1[2] * fq) * dp2;
      }
      real chain = real(1), dd = d2;
      if (premap) dd = premap_fork<real>(d2, K.premap_lo, K.premap_hi, chain);
      const real dphoto = (q_ok && dd >= real(1e-6)) ? -(D * D) * dD * chain : real(0);
      if (!CAM) {
        const real e0 = dp0 * D, e1 = dp1 * D, e2 = dp2 * D;
        S0a[0] += e0; S0a[1] += e1; S0a[2] += e2;
        Sva[0] += e0 * fq; Sva[1] += e1 * fq; Sva[2] += e2 * fq;
        S1a[0] += dp0; S1a[1] += dp1; S1a[2] += dp2;
      }
      if (j == 0) {
        real gs = real(0);
        if (smooth) {
          const real dl = shfl_up1(d2), dr = shfl_dn1(d2);
          const real wxl = shfl_up1(wxq);
          const real cx = K.coef_x[s], cy = K.coef_y[s];
          real qq = real(0);
          if (u + 1 < W) qq += signed_by(cx * wxq, d2 - dr);
          if (u >= 1) qq -= signed_by(cx * wxl, dl - d2);
          if (q + 1 < H) qq += signed_by(cy * wyq, d2 - d1);
          if (q >= 1) qq -= signed_by(cy * wy_prev, d3 - d2);
          gs = qq * inv_m - cb;
          wy_prev = wyq;
        }
        if (q_ok) {
          if (C > 1) { dsum_prev = dphoto + gs; prev_q = q; }
          else d_out[q * W + u] = dphoto + gs;
        }
      } else {
        dpart_s[(par * C + j) * 32 + lane] = dphoto;
      }
    }
    // ---------------- finish the gather of row t + 1: its loads had the whole step to land -----------------
    finish_row(t + 1);
  };

  for (int tb = t0; tb <= r1 + 1; tb += 3) {
    step(tb, IntC<0>());
    step(tb + 1, IntC<1>());
    step(tb + 2, IntC<2>());
  }
  if (C > 1) {
    __syncthreads();
    const int last_t = t0 + ((r1 + 1 - t0) / 3) * 3 + 2;
    finalize_row(last_t & 1);
  }

  // ---- epilogue: pose partials (converted from the (u, v, 1) basis) and the loss partial -----------
  if (CAM) {
    // d loss / d [R|t] directly: rows i, columns (ray x, ray y, 1 | translation)
    real out = real(0);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const real g0 = warp_allsum<real>(Sua[i]), g1 = warp_allsum<real>(Sva[i]), g2 = warp_allsum<real>(S0a[i]),
                 g3 = warp_allsum<real>(S1a[i]);
      if (lane == i * 4) out = g0;
      if (lane == i * 4 + 1) out = g1;
      if (lane == i * 4 + 2) out = g2;
      if (lane == i * 4 + 3) out = g3;
    }
    if (lane < 12) ((real*)SP.pose_part[s])[(((size_t)b * nunits + unit) * C + j) * 12 + lane] = out;
  } else {
    SweepCam<real> cam;
    sweep_cam<real>(cam, P, b, j, K.sf[s]);
    const real fu = (real)su;
    real G[12];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const real eu = fu * S0a[i], ev = Sva[i], e1 = S0a[i];
#pragma unroll
      for (int k = 0; k < 3; ++k) G[i * 4 + k] = cam.ki[k * 3] * eu + cam.ki[k * 3 + 1] * ev + cam.ki[k * 3 + 2] * e1;
      G[i * 4 + 3] = S1a[i];
    }
#pragma unroll
    for (int e = 0; e < 12; ++e) G[e] = warp_allsum<real>(G[e]);
    // d loss / d [R|t] = ref_K^T G
    real out = real(0);
#pragma unroll
    for (int e = 0; e < 12; ++e) {
      const int i = e / 4, k = e % 4;
      const real v = cam.rk[i] * G[k] + cam.rk[3 + i] * G[4 + k] + cam.rk[6 + i] * G[8 + k];
      if (lane == e) out = v;
    }
    if (lane < 12) ((real*)SP.pose_part[s])[(((size_t)b * nunits + unit) * C + j) * 12 + lane] = out;
  }
  const real lsum = warp_allsum<real>(loss_acc);
  if (j == 0 && lane == 0) SP.loss_part[s][(size_t)b * nunits + unit] = (double)lsum;
}

// ------------------------------------------------------------------------------------------
// pool kernel: adjoint of the in-kernel nearest upsampling.  grid (ceil(Hd Wd / 256), B), 256 threads; one
// thread sums, in a fixed order, the full-resolution gradients of the pixels that read its map element.
// ------------------------------------------------------------------------------------------
// power-of-two factors (the decoder pyramid): block = one map row x 256 full-resolution columns; coalesced column sums
// over the fy rows, then fx neighbours from shared memory, both in a fixed order
template <typename real>
__global__ void __launch_bounds__(256) pool_pow2_kernel(int H, int W, int Hd, int Wd, int fy, int fx,
                                                        const real* __restrict__ full, real* __restrict__ out) {
  __shared__ real col[256];
  const int vs = blockIdx.y, b = blockIdx.z, c = blockIdx.x * 256 + threadIdx.x;
  real acc = real(0);
  if (c < W) {
    const real* src = full + ((size_t)b * H + (size_t)vs * fy) * W + c;
    for (int r = 0; r < fy; ++r) acc += src[(size_t)r * W];
  }
  col[threadIdx.x] = acc;
  __syncthreads();
  const int us = (blockIdx.x * 256) / fx + threadIdx.x;
  if ((int)threadIdx.x < 256 / fx && us < Wd) {
    real a = real(0);
    for (int k = 0; k < fx; ++k) a += col[threadIdx.x * fx + k];
    out[((size_t)b * Hd + vs) * Wd + us] = a;
  }
}

template <typename real>
__global__ void __launch_bounds__(256) pool_kernel(int H, int W, int Hd, int Wd, float rs, float cs,
                                                   const real* __restrict__ full, real* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= Hd * Wd) return;
  const int vs = i / Wd, us = i - vs * Wd, b = blockIdx.y;
  const int v0 = max(0, (int)((float)vs / rs) - 1), v1 = min(H - 1, (int)((float)(vs + 1) / rs) + 1);
  const int u0 = max(0, (int)((float)us / cs) - 1), u1 = min(W - 1, (int)((float)(us + 1) / cs) + 1);
  const real* src = full + (size_t)b * H * W;
  real acc = real(0);
  for (int v = v0; v <= v1; ++v) {
    if (nearest_src(v, rs, Hd, true) != vs) continue;
    for (int u = u0; u <= u1; ++u)
      if (nearest_src(u, cs, Wd, true) == us) acc += src[(size_t)v * W + u];
  }
  out[(size_t)b * Hd * Wd + i] = acc;
}

// ------------------------------------------------------------------------------------------
// final kernel: block 0 -> loss triple; block 1 + b * C + j -> d pose[j].mat[b].
// Works over all sets (called once, after every set's sweep kernel).
// ------------------------------------------------------------------------------------------
struct FinalParams {
  int B, C, n;
  int nunits[PSFM_MAX_SCALES];
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
    // one pass over all scales: per-thread partial sums in a fixed order, one butterfly per warp, eight warp sums
    double photo = 0.0, smooth = 0.0;
    // the scales' partial lists are read side by side (independent loads in flight; the order of every sum is fixed)
    double a4[PSFM_MAX_SCALES] = {0.0, 0.0, 0.0, 0.0};
    size_t cmax = 0;
    for (int s = 0; s < F.n; ++s) cmax = max(cmax, (size_t)F.B * F.nunits[s]);
    for (size_t i = tid; i < cmax; i += blockDim.x) {
#pragma unroll
      for (int s = 0; s < PSFM_MAX_SCALES; ++s)
        if (s < F.n && i < (size_t)F.B * F.nunits[s]) a4[s] += F.loss_part[s][i];
    }
    for (int s = 0; s < F.n; ++s) photo += a4[s] * F.inv_count[s];
    if (F.smooth)
      for (int i = tid; i < F.n * F.B; i += blockDim.x) {
        const int s = i / F.B, b = i - s * F.B;
        const real* st = (const real*)F.stats[s] + (size_t)b * 4;
        smooth += (double)st[2] * F.coef_x[s] + (double)st[3] * F.coef_y[s];
      }
    photo = warp_allsum<double>(photo);
    smooth = warp_allsum<double>(smooth);
    if ((tid & 31) == 0) { sh[tid >> 5] = photo; sh[8 + (tid >> 5)] = smooth; }
    __syncthreads();
    if (tid == 0) {
      double ph = 0.0, sm = 0.0;
      for (int w = 0; w < 8; ++w) { ph += sh[w]; sm += sh[8 + w]; }
      real* out = (real*)F.loss_out;
      out[0] = (real)((real)ph + (real)sm);
      out[1] = (real)ph;
      out[2] = (real)sm;
      *(real*)F.one = real(1);
    }
    return;
  }
  const int bj = blockIdx.x - 1, b = bj / F.C, j = bj - b * F.C;
  // 12 entries x 21 partial lanes, fixed order: lane k of entry e sums units k, k+21, ...
  const int e = tid % 12, k = tid / 12;      // 252 active threads
  double a = 0.0;
  if (tid < 252) {
    double a4[PSFM_MAX_SCALES] = {0.0, 0.0, 0.0, 0.0};
    int umax = 0;
    for (int s = 0; s < F.n; ++s) umax = max(umax, F.nunits[s]);
    for (int i = k; i < umax; i += 21) {
#pragma unroll
      for (int s = 0; s < PSFM_MAX_SCALES; ++s)
        if (s < F.n && i < F.nunits[s])
          a4[s] += (double)((const real*)F.pose_part[s])[(((size_t)b * F.nunits[s] + i) * F.C + j) * 12 + e];
    }
    for (int s = 0; s < F.n; ++s) a += a4[s];
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
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

constexpr int kMinChunk = 16;        // smallest row chunk of either kernel (bounds the partial buffers)

bool stream_eligible(const psfm_loss_cfg* cfg) {
  if (cfg->reduce_op != PSFM_REDUCE_MIN || !(cfg->ssim_w > 0.0) || cfg->clip_loss > 0.0) return false;
  if (cfg->C > 8) return false;
  for (int s = 0; s < cfg->n; ++s)
    if (cfg->Hs[s] < 4 || cfg->Ws[s] < 4) return false;
  return true;
}

struct SweepPlan {
  int nsets;
  SweepSet sets[PSFM_MAX_SCALES];
  size_t off_rec[PSFM_MAX_SCALES], off_wxy[PSFM_MAX_SCALES], off_ctxp[PSFM_MAX_SCALES], off_dfull[PSFM_MAX_SCALES], off_ray[PSFM_MAX_SCALES];                                         // per scale (a set uses its first scale's)
  size_t off_prep[PSFM_MAX_SCALES], off_stats[PSFM_MAX_SCALES], off_loss[PSFM_MAX_SCALES],
      off_pose[PSFM_MAX_SCALES];                                          // per scale
  int nunits[PSFM_MAX_SCALES];
  size_t off_one, total;
};

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

// Rows per sweep unit.  A unit of h rows runs h + 4 steps (2 warm-up rows of the window, 2 of the adjoint),
// rounded up to the 3-step unrolling; the grid runs in waves of (148 SMs x resident CTAs).  Pick the
// height that minimises waves x steps (measured at config 2: 48 and 64 rows 242 us, 32 rows 259 us,
// 96 rows 261 us, 24 rows 279 us -- the model orders them the same way).
static int pick_sweep_chunk(int H, long ctas_per_chunk, int C, bool f64) {
  static const int cand[] = {16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512, 1 << 30};
  const long resident = 148L * ((f64 ? 4 : 8) / (C < 1 ? 1 : (C > 4 ? 4 : C)) > 0 ? (f64 ? 4 : 8) / (C < 1 ? 1 : (C > 4 ? 4 : C)) : 1);
  int best = H;
  double best_cost = 1e300;
  for (int c : cand) {
    const int ch = c < H ? c : H;
    if (ch < kMinChunk && ch < H) continue;
    const long chunks = cdiv(H, ch);
    const long waves = (chunks * ctas_per_chunk + resident - 1) / resident;
    const double cost = (double)waves * (double)((ch + 4 + 2) / 3 * 3);
    if (cost < best_cost * 0.999) { best_cost = cost; best = ch; }
    if (c >= H) break;
  }
  return best;
}

// Rows per unit of the two-pixels-per-lane sweep (loss_sweep2.cu): h rows run h + 4 steps, any height is allowed;
// 148 SMs x (8 / C) CTAs are resident (255 registers per thread).
static int pick_sweep2_chunk(int H, long ctas_per_chunk, int C, bool f64) {
  const long resident = 148L * (f64 ? 1 : (C <= 1 ? 8 : (C == 2 ? 4 : 2)));
  int best = H;
  double best_cost = 1e300;
  for (int ch = (H < kMinChunk ? H : kMinChunk); ch <= H; ++ch) {
    const long chunks = cdiv(H, ch);
    if (ch != cdiv(H, (int)chunks)) continue;            // the balanced height for this number of chunks
    const long waves = (chunks * ctas_per_chunk + resident - 1) / resident;
    const double cost = (double)waves * (double)(ch + 4);
    if (cost < best_cost * 0.999) { best_cost = cost; best = ch; }
  }
  return best;
}

// Rows per prep unit (one warp): h rows run h + 2 steps, rounded up to the 2-step unrolling; 148 SMs x 2 CTAs x
// 4 warps are resident.  Measured at config 2: 32 rows 63 us, 16 rows 68 us, 24 / 48 rows 78-79 us.
static int pick_prep_chunk(int H, long warps_per_chunk) {
  static const int cand[] = {16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512, 1 << 30};
  const long resident = 148L * 2 * kPrepWarps;
  int best = H;
  double best_cost = 1e300;
  for (int c : cand) {
    const int ch = c < H ? c : H;
    if (ch < kMinChunk && ch < H) continue;
    const long waves = ((long)cdiv(H, ch) * warps_per_chunk + resident - 1) / resident;
    const double cost = (double)waves * (double)((ch + 2 + 1) / 2 * 2);
    if (cost < best_cost * 0.999) { best_cost = cost; best = ch; }
    if (c >= H) break;
  }
  return best;
}

static bool use_sweep2(const psfm_loss_cfg* cfg) { return sweep2_eligible(cfg) && env_int("PSFM_SWEEP2", 0) != 0; }

static void make_plan(const psfm_loss_cfg* cfg, const psfm_loss_io* io, SweepPlan& pl) {
  const size_t rs = cfg->real_dtype == PSFM_F64 ? 8 : 4;
  pl.nsets = 0;
  bool used[PSFM_MAX_SCALES] = {false, false, false, false};
  for (int s = 0; s < cfg->n; ++s) {
    if (used[s]) continue;
    SweepSet& S = pl.sets[pl.nsets++];
    S.ns = 0; S.H = cfg->Hs[s]; S.W = cfg->Ws[s];
    for (int k = s; k < cfg->n; ++k) {
      if (used[k] || cfg->Hs[k] != S.H || cfg->Ws[k] != S.W) continue;
      if (io && (io->image[k] != io->image[s] || io->mask[k] != io->mask[s])) continue;
      bool same = true;
      if (io) for (int j = 0; j < cfg->C; ++j) same = same && io->ctx[k][j] == io->ctx[s][j];
      if (!same) continue;
      S.scale[S.ns++] = k;
      used[k] = true;
    }
    S.Hp = S.H + kPadLo + kPadHi; S.Wp = S.W + kPadLo + kPadHi;
    S.Wh = (S.W + 1) / 2;
    S.nstrips = cdiv(S.W, kInner);
    const int forced = env_int("PSFM_SWEEP_CHUNK", 0);
    S.chunk_h = forced > 0 ? (forced < S.H ? forced : S.H)
                           : pick_sweep_chunk(S.H, (long)S.nstrips * S.ns * cfg->B, cfg->C, cfg->real_dtype == PSFM_F64);
    if (S.chunk_h < kMinChunk) S.chunk_h = S.H < kMinChunk ? S.H : kMinChunk;
    S.nchunks = cdiv(S.H, S.chunk_h);
    S.nstrips2 = cdiv(S.W, kInner2);
    const int forced2 = env_int("PSFM_SWEEP2_CHUNK", 0);
    S.chunk_h2 = forced2 > 0 ? (forced2 < S.H ? forced2 : S.H)
                             : pick_sweep2_chunk(S.H, (long)S.nstrips2 * S.ns * cfg->B, cfg->C, cfg->real_dtype == PSFM_F64);
    if (S.chunk_h2 < kMinChunk) S.chunk_h2 = S.H < kMinChunk ? S.H : kMinChunk;
    S.nchunks2 = cdiv(S.H, S.chunk_h2);
    S.pstrips = cdiv(S.W, kPrepInner);
    const int pforced = env_int("PSFM_PREP_CHUNK", 0);
    S.pchunk_h = pforced > 0 ? (pforced < S.H ? pforced : S.H) : pick_prep_chunk(S.H, (long)S.pstrips * cfg->B);
    if (S.pchunk_h < kMinChunk) S.pchunk_h = S.H < kMinChunk ? S.H : kMinChunk;
    S.pchunks = cdiv(S.H, S.pchunk_h);
  }
  // workspace layout: the image-sized regions (records, packed contexts, edge weights, rays) exist once per SET --
  // one set in the shipped upsample_depth_maps mode, where all scales share the images; without `io` (size query by
  // configuration alone) every scale is assumed to be its own set
  bool heavy[PSFM_MAX_SCALES];
  for (int s = 0; s < cfg->n; ++s) heavy[s] = io == nullptr;
  for (int i = 0; i < pl.nsets && io; ++i) heavy[pl.sets[i].scale[0]] = true;
  size_t off = 0;
  pl.off_one = off; off = al(off + rs);
  for (int s = 0; s < cfg->n; ++s) {
    const int Hs = cfg->Hs[s], Ws = cfg->Ws[s];
    const size_t px = (size_t)cfg->B * Hs * Ws;
    const int minc = Hs < kMinChunk ? Hs : kMinChunk;
    const size_t worst = (size_t)cdiv(Ws, kInner) * cdiv(Hs, minc);
    const size_t pworst = (size_t)cdiv(Ws, kPrepInner) * cdiv(Hs, minc);
    pl.off_rec[s] = off;   if (heavy[s]) off = al(off + (size_t)cfg->B * Hs * 2 * ((Ws + 1) / 2) * 12 * rs);   // three real4 parts
    pl.off_ctxp[s] = off;  if (heavy[s]) off = al(off + (size_t)cfg->C * cfg->B * (Hs + kPadLo + kPadHi) * (Ws + kPadLo + kPadHi) * 4 * rs);
    pl.off_wxy[s] = off;   if (heavy[s]) off = al(off + px * 2 * rs);
    pl.off_ray[s] = off;   if (heavy[s] && cfg->camera_model == PSFM_CAM_FISHEYE) off = al(off + px * 2 * rs);
    const bool ups = cfg->Hd[s] > 0 && (cfg->Hd[s] != Hs || cfg->Wd[s] != Ws);
    pl.off_dfull[s] = off; if (ups) off = al(off + px * rs);            // full-resolution d inv_depth before pooling
    pl.off_prep[s] = off;  off = al(off + (size_t)cfg->B * pworst * kPrepVals * sizeof(double));
    pl.off_stats[s] = off; off = al(off + (size_t)cfg->B * 4 * rs);
    pl.off_loss[s] = off;  off = al(off + (size_t)cfg->B * worst * sizeof(double));
    pl.off_pose[s] = off;  off = al(off + (size_t)cfg->B * worst * cfg->C * 12 * rs);
  }
  for (int i = 0; i < pl.nsets; ++i) {
    const SweepSet& S = pl.sets[i];
    for (int k = 0; k < S.ns; ++k) pl.nunits[S.scale[k]] = use_sweep2(cfg) ? S.nstrips2 * S.nchunks2 : S.nstrips * S.nchunks;
  }
  pl.total = off;
}

size_t stream_workspace_bytes(const psfm_loss_cfg* cfg, const psfm_loss_io* io) {
  SweepPlan pl;
  make_plan(cfg, io, pl);
  return pl.total;
}

void build_params(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws, LossParams& P);   // loss_kernels.cu

template <typename real>
static void fill_consts(SweepConsts<real>& c, const LossParams& P, int s0) {
  c.ssim_w3 = (real)P.ssim_w / real(3);
  c.k1 = (real(1) - (real)P.ssim_w) / real(3);
  c.C1 = (real)P.C1; c.C2 = (real)P.C2;
  c.inv_count = (real)P.sc[s0].inv_count;
  c.premap_lo = (real)P.premap_lo; c.premap_hi = (real)P.premap_hi;
  c.pad_ = real(0);
  for (int s = 0; s < PSFM_MAX_SCALES; ++s) {
    c.coef_x[s] = s < P.n ? (real)P.sc[s].coef_x : real(0);
    c.coef_y[s] = s < P.n ? (real)P.sc[s].coef_y : real(0);
    c.sf[s] = s < P.n ? (real)P.sc[s].sf : real(1);
  }
}

template <typename real, int PAD, int VAR, int CT, int CAM = 0, bool DBG = false>
static cudaError_t launch_sweep_ct(const SweepParams& SP, dim3 grid, int threads, size_t smem, cudaStream_t st) {
  auto kern = sweep_kernel<real, PAD, VAR, CT, CAM, DBG>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  kern<<<grid, threads, smem, st>>>(SP);
  return cudaSuccess;
}

// the common context counts get their own instantiation (unrolled candidate min, tighter launch bounds)
template <typename real, int PAD, int VAR>
static cudaError_t launch_sweep(const SweepParams& SP, dim3 grid, int threads, size_t smem, cudaStream_t st) {
  if (VAR != 2) {
    if (SP.P.C == 2) return launch_sweep_ct<real, PAD, VAR, 2>(SP, grid, threads, smem, st);
    if (SP.P.C == 4) return launch_sweep_ct<real, PAD, VAR, 4>(SP, grid, threads, smem, st);
  }
  return launch_sweep_ct<real, PAD, VAR, 0>(SP, grid, threads, smem, st);
}

template <typename real, typename img_t>
static cudaError_t run_sweep(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* ws_split, void* ws_stream,
                             void* loss_out, cudaStream_t st) {
  SweepPlan pl;
  make_plan(cfg, io, pl);
  SweepParams SP;
  build_params(cfg, io, ws_split, SP.P);
  char* w = (char*)ws_stream;
  for (int s = 0; s < cfg->n; ++s) {
    SP.prep_part[s] = (double*)(w + pl.off_prep[s]);
    SP.stats[s] = w + pl.off_stats[s];
    SP.loss_part[s] = (double*)(w + pl.off_loss[s]);
    SP.pose_part[s] = w + pl.off_pose[s];
  }
  for (int s = 0; s < cfg->n; ++s)      // upsampled scales: the sweep writes full-resolution gradients, pooled afterwards
    if (SP.P.sc[s].Hd != SP.P.sc[s].H || SP.P.sc[s].Wd != SP.P.sc[s].W) SP.P.sc[s].d_full = (char*)ws_stream + pl.off_dfull[s];
  for (int i = 0; i < pl.nsets; ++i) {
    SP.set = pl.sets[i];
    const int s0 = SP.set.scale[0];
    {
      const size_t px = (size_t)cfg->B * SP.set.H * 2 * SP.set.Wh * 4 * sizeof(real);
      SP.set.ra = w + pl.off_rec[s0];
      SP.set.rb = w + pl.off_rec[s0] + 4 * sizeof(real);
      SP.set.rc = w + pl.off_rec[s0] + 8 * sizeof(real);
      const size_t cpx = (size_t)cfg->B * SP.set.Hp * SP.set.Wp * 4 * sizeof(real);
      for (int j = 0; j < PSFM_MAX_CTX; ++j) SP.set.ctxp[j] = j < cfg->C ? w + pl.off_ctxp[s0] + j * cpx : nullptr;
    }
    SP.set.wxy = w + pl.off_wxy[s0];
    SP.set.ray = w + pl.off_ray[s0];
    fill_consts(SP.cf, SP.P, s0);
    fill_consts(SP.cd, SP.P, s0);
    const SweepSet& S = SP.set;
    {
      int npass = 0;
      for (int j0 = 0; j0 < cfg->C; j0 += (cfg->C - j0 >= 4 ? 4 : (cfg->C - j0 >= 2 ? 2 : 1))) ++npass;
      ProfScope ps(st, kTagFwd, npass + (cfg->smooth_w > 0.0 ? 1 : 0) + (cfg->camera_model == PSFM_CAM_FISHEYE ? 1 : 0));
      for (int j0 = 0; j0 < cfg->C;) {
        const int ct = cfg->C - j0 >= 4 ? 4 : (cfg->C - j0 >= 2 ? 2 : 1);
        const dim3 pgrid(cdiv(S.pstrips * S.pchunks, kPrepWarps) + cdiv(S.Hp * ct, kPrepWarps), cfg->B);
#define PSFM_PREP(CTV, AM, FI) prep_kernel<real, img_t, CTV, AM, FI><<<pgrid, 32 * kPrepWarps, 0, st>>>(SP, j0)
#define PSFM_PREP_CT(AM, FI) (ct == 4 ? PSFM_PREP(4, AM, FI) : ct == 2 ? PSFM_PREP(2, AM, FI) : PSFM_PREP(1, AM, FI))
        if (cfg->automask) { if (j0 == 0) PSFM_PREP_CT(true, true); else PSFM_PREP_CT(true, false); }
        else { if (j0 == 0) PSFM_PREP_CT(false, true); else PSFM_PREP_CT(false, false); }
#undef PSFM_PREP_CT
#undef PSFM_PREP
        j0 += ct;
      }
      if (cfg->smooth_w > 0.0) stats_kernel<real><<<dim3(cfg->B, S.ns), 128, 0, st>>>(SP);
      if (cfg->camera_model == PSFM_CAM_FISHEYE)
        fisheye_ray_kernel<real><<<dim3(cdiv(S.H * S.W, 256), cfg->B), 256, 0, st>>>(SP);
    }
    const bool debug = io->dbg[s0] != nullptr;
    if (use_sweep2(cfg)) {
      ProfScope ps(st, kTagFused, 1);
      bool ups2 = false;
      for (int k = 0; k < S.ns; ++k)
        if (SP.P.sc[S.scale[k]].d_full != SP.P.sc[S.scale[k]].d_invd) ups2 = true;
      const cudaError_t e2 = launch_sweep2<real>(SP, ups2, debug, st);
      if (e2 != cudaSuccess) return e2;
      continue;
    }
    const bool fisheye = cfg->camera_model == PSFM_CAM_FISHEYE;
    const int ring = fisheye ? 15 : (cfg->padding == PSFM_PAD_BORDER ? 11 : 10);
    const size_t smem = ((size_t)cfg->C * 4 * ring * 32 + (size_t)4 * cfg->C * 32 + (fisheye ? 24 * cfg->C : 0)) * sizeof(real);
    const dim3 grid(S.nstrips * S.nchunks, S.ns, cfg->B);
    const int threads = 32 * cfg->C;
    ProfScope ps(st, kTagFused, 1);
    cudaError_t e;
    int var = cfg->premap ? 2 : 0;
    for (int k = 0; k < S.ns && var == 0; ++k)
      if (SP.P.sc[S.scale[k]].d_full != SP.P.sc[S.scale[k]].d_invd) var = 1;
#define PSFM_SWEEP(PADV) (var == 2 ? launch_sweep<real, PADV, 2>(SP, grid, threads, smem, st) \
                          : var == 1 ? launch_sweep<real, PADV, 1>(SP, grid, threads, smem, st) \
                                     : launch_sweep<real, PADV, 0>(SP, grid, threads, smem, st))
    if (debug) {
      // decision words / arg-min maps (parity tests): fp32, plain inputs, 2 or 4 contexts (fisheye: 2)
      e = cudaErrorNotSupported;
      if constexpr (sizeof(real) == 4) {
        const bool bd = cfg->padding == PSFM_PAD_BORDER;
        if (var == 0 && fisheye && cfg->C == 2)
          e = bd ? launch_sweep_ct<real, PSFM_PAD_BORDER, 0, 2, 1, true>(SP, grid, threads, smem, st)
                 : launch_sweep_ct<real, PSFM_PAD_ZEROS, 0, 2, 1, true>(SP, grid, threads, smem, st);
        else if (var == 0 && !fisheye && cfg->C == 2)
          e = bd ? launch_sweep_ct<real, PSFM_PAD_BORDER, 0, 2, 0, true>(SP, grid, threads, smem, st)
                 : launch_sweep_ct<real, PSFM_PAD_ZEROS, 0, 2, 0, true>(SP, grid, threads, smem, st);
        else if (var == 0 && !fisheye && cfg->C == 4)
          e = bd ? launch_sweep_ct<real, PSFM_PAD_BORDER, 0, 4, 0, true>(SP, grid, threads, smem, st)
                 : launch_sweep_ct<real, PSFM_PAD_ZEROS, 0, 4, 0, true>(SP, grid, threads, smem, st);
      }
    } else if (fisheye) {
      // fp32, two contexts, plain inputs: own instantiation; everything else (any C, pre-map, upsampled inverse
      // depths, fp64) runs the run-time-generic one
      bool done = false;
      if constexpr (sizeof(real) == 4) {
        if (var == 0 && cfg->C == 2) {
          e = cfg->padding == PSFM_PAD_BORDER ? launch_sweep_ct<real, PSFM_PAD_BORDER, 0, 2, 1>(SP, grid, threads, smem, st)
                                              : launch_sweep_ct<real, PSFM_PAD_ZEROS, 0, 2, 1>(SP, grid, threads, smem, st);
          done = true;
        }
      }
      if (!done)
        e = cfg->padding == PSFM_PAD_BORDER ? launch_sweep_ct<real, PSFM_PAD_BORDER, 2, 0, 1>(SP, grid, threads, smem, st)
                                            : launch_sweep_ct<real, PSFM_PAD_ZEROS, 2, 0, 1>(SP, grid, threads, smem, st);
    }
    else
      e = cfg->padding == PSFM_PAD_BORDER ? PSFM_SWEEP(PSFM_PAD_BORDER) : PSFM_SWEEP(PSFM_PAD_ZEROS);
#undef PSFM_SWEEP
    if (e != cudaSuccess) return e;
  }
  for (int s = 0; s < cfg->n; ++s) {
    const ScaleDesc& D = SP.P.sc[s];
    if (D.d_full == D.d_invd) continue;
    ProfScope ps(st, kTagFinish, 1);
    const int fy = D.H / D.Hd, fx = D.W / D.Wd;
    const bool pow2 = fy * D.Hd == D.H && fx * D.Wd == D.W && (fy & (fy - 1)) == 0 && (fx & (fx - 1)) == 0 && fx <= 256;
    if (pow2)   // floor(dst * (1 / 2^k)) == dst >> k exactly
      pool_pow2_kernel<real><<<dim3(cdiv(D.W, 256), D.Hd, cfg->B), 256, 0, st>>>(D.H, D.W, D.Hd, D.Wd, fy, fx,
                                                                               (const real*)D.d_full, (real*)D.d_invd);
    else
      pool_kernel<real><<<dim3(cdiv(D.Hd * D.Wd, 256), cfg->B), 256, 0, st>>>(D.H, D.W, D.Hd, D.Wd, D.rs, D.cs,
                                                                            (const real*)D.d_full, (real*)D.d_invd);
  }
  FinalParams F;
  F.B = cfg->B; F.C = cfg->C; F.n = cfg->n; F.smooth = cfg->smooth_w > 0.0;
  for (int s = 0; s < cfg->n; ++s) {
    F.nunits[s] = pl.nunits[s];
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
  if (cfg->real_dtype == PSFM_F64) return run_sweep<double, double>(cfg, io, ws_split, ws_stream, loss_out, st);
  if (cfg->img_dtype == PSFM_BF16) return run_sweep<float, __nv_bfloat16>(cfg, io, ws_split, ws_stream, loss_out, st);
  if (cfg->img_dtype == PSFM_U8) return run_sweep<float, uint8_t>(cfg, io, ws_split, ws_stream, loss_out, st);
  return run_sweep<float, float>(cfg, io, ws_split, ws_stream, loss_out, st);
}

}  // namespace psfm

