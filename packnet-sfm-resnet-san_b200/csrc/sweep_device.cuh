// sweep_device.cuh -- per-lane building blocks of the row-sweep kernels (loss_sweep.cu).
//
// A warp owns a strip of 32 adjacent columns and marches down the rows; lane l is column
// c0 - halo + l.  Horizontal neighbours are exchanged with warp shuffles, the vertical 3-row
// windows live in register rings.  Arithmetic follows SURVEY.md Appendix A (reference
// losses/multiview_photometric_loss.py:15-54,199-267, geometry/camera.py:111-190,
// torch grid_sampler_2d bilinear / align_corners=True); the algebra is rearranged (pose and
// intrinsics pre-multiplied, window statistics of the target image precomputed) but is the same
// function of the inputs, which the fp64 instantiation checks to 1e-10 against the oracle.
#pragma once
#include "psfm_device.cuh"
#include "pair_math.cuh"

namespace psfm {

constexpr unsigned kFull = 0xffffffffu;

// elementary functions of the fisheye camera model (geometry/camera.py:194-394)
__device__ __forceinline__ double rroot(double x) { return sqrt(x); }
__device__ __forceinline__ float rroot(float x) { return sqrtf(x); }
__device__ __forceinline__ double rtan(double x) { return tan(x); }
__device__ __forceinline__ float rtan(float x) { return tanf(x); }
__device__ __forceinline__ double ratan(double x) { return atan(x); }
__device__ __forceinline__ float ratan(float x) { return atanf(x); }
// reciprocal: exact for double; MUFU.RCP + one Newton step (< 1 ulp) for float
__device__ __forceinline__ double rrcp(double x) { return 1.0 / x; }
__device__ __forceinline__ float rrcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  const float e = fmaf(-x, r, 1.0f);
  return fmaf(r, e, r);
}

// reciprocal for the SSIM denominator: MUFU.RCP alone (1 ulp) is ample for a value in [0, 1]
__device__ __forceinline__ double rrcp_fast(double x) { return 1.0 / x; }
__device__ __forceinline__ float rrcp_fast(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <typename real> __device__ __forceinline__ real shfl_up1(real v) { return __shfl_up_sync(kFull, v, 1); }
template <typename real> __device__ __forceinline__ real shfl_dn1(real v) { return __shfl_down_sync(kFull, v, 1); }

template <typename real> __device__ __forceinline__ real warp_allsum(real v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------
// SSIM distance of one channel from the 3x3 window SUMS of x, x^2, x*y and the precomputed
// target statistics my = mean(y), qy = E[y^2] - my^2 + C2 (reference :15-54, :199-216).
// When GRAD: coef[0..2] = d dist / d (sum x, sum x^2 (already x2 for d x^2/dx), sum xy).
// ------------------------------------------------------------------------------------------
template <typename real, bool GRAD>
__device__ __forceinline__ real ssim_dist_sums(real Sx, real Sxx, real Sxy, real my, real qy, real C1, real C2,
                                               real* coef) {
  const real r9 = real(1) / real(9);
  const real mx = Sx * r9, exx = Sxx * r9, exy = Sxy * r9;
  const real mx2 = mx * mx, mxy = mx * my;
  const real a1 = real(2) * mxy + C1;
  const real v1 = real(2) * (exy - mxy) + C2;
  const real b1 = mx2 + (my * my + C1);
  const real v2 = (exx - mx2) + qy;
  const real inv = rrcp_fast(b1 * v2);
  const real ssim = (a1 * v1) * inv;
  const real dist = real(0.5) - real(0.5) * ssim;
  if (GRAD) {
    // d dist = -0.5 d ssim inside the clamp range, 0 outside.  Outside (which includes a non-finite
    // ssim from a degenerate window of a padding row) every factor is replaced by a finite one, so
    // the coefficients are always finite and can be scaled by a zero selection weight.
    const bool in = (dist >= real(0)) && (dist <= real(1));
    const real act = in ? real(-0.5) * r9 * inv : real(0);
    const real ss = in ? ssim : real(0);
    coef[0] = act * (real(2) * my * (v1 - a1) - ss * real(2) * mx * (v2 - b1));
    coef[1] = act * (real(-2) * ss * b1);
    coef[2] = act * (real(2) * a1);
  }
  return rmin(rmax(dist, real(0)), real(1));
}

// ------------------------------------------------------------------------------------------
// bilinear sample of a 3-plane image at (ix, iy) with value and coordinate gradients
// (grid_sample bilinear, align_corners=True), split in two so that the 12 loads can be in flight
// while other rows are processed.  Out-of-image taps are removed through their WEIGHTS (zeros
// padding) and every address is clamped into the plane, so all loads are unconditional.
// PAD == border clips the coordinate and zeroes its gradient where clipped.
//   wgt[8] = {ax0, ax1, ay0, ay1, sx0, sx1, sy0, sy1}: value weights and d/d(ix,iy) tap signs
//   tap[c * 4 + {0,1,2,3}] = {v00, v01, v10, v11} of channel c
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t, int PAD>
__device__ __forceinline__ void sample3_issue(const img_t* __restrict__ p0, const img_t* __restrict__ p1,
                                              const img_t* __restrict__ p2, int H, int W, real ix, real iy,
                                              real* wgt, real* tap) {
  real mx = real(1), my = real(1);
  if (PAD == PSFM_PAD_BORDER) {
    if (!(ix > real(0))) { ix = real(0); mx = real(0); }
    else if (ix >= real(W - 1)) { ix = real(W - 1); mx = real(0); }
    if (!(iy > real(0))) { iy = real(0); my = real(0); }
    else if (iy >= real(H - 1)) { iy = real(H - 1); my = real(0); }
  } else {
    // fully outside beyond [-1, size]: keep the value there so the int conversion is defined
    ix = rmin(rmax(ix, real(-2)), real(W + 1));
    iy = rmin(rmax(iy, real(-2)), real(H + 1));
  }
  const real fx = rfloor(ix), fy = rfloor(iy);
  const int x0 = (int)fx, y0 = (int)fy;
  const real wx1 = ix - fx, wx0 = (fx + real(1)) - ix;
  const real wy1 = iy - fy, wy0 = (fy + real(1)) - iy;
  const bool bx0 = (unsigned)x0 < (unsigned)W, bx1 = (unsigned)(x0 + 1) < (unsigned)W;
  const bool by0 = (unsigned)y0 < (unsigned)H, by1 = (unsigned)(y0 + 1) < (unsigned)H;
  wgt[0] = bx0 ? wx0 : real(0); wgt[1] = bx1 ? wx1 : real(0);
  wgt[2] = by0 ? wy0 : real(0); wgt[3] = by1 ? wy1 : real(0);
  wgt[4] = bx0 ? mx : real(0); wgt[5] = bx1 ? mx : real(0);
  wgt[6] = by0 ? my : real(0); wgt[7] = by1 ? my : real(0);
  const int xc0 = min(max(x0, 0), W - 1), xc1 = min(max(x0 + 1, 0), W - 1);
  const int yc0 = min(max(y0, 0), H - 1), yc1 = min(max(y0 + 1, 0), H - 1);
  const int r0 = yc0 * W, r1 = yc1 * W;
  const int o00 = r0 + xc0, o01 = r0 + xc1, o10 = r1 + xc0, o11 = r1 + xc1;
  const img_t* pl[3] = {p0, p1, p2};
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    tap[c * 4 + 0] = ldimg<real>(pl[c], o00); tap[c * 4 + 1] = ldimg<real>(pl[c], o01);
    tap[c * 4 + 2] = ldimg<real>(pl[c], o10); tap[c * 4 + 3] = ldimg<real>(pl[c], o11);
  }
}

template <typename real>
__device__ __forceinline__ void sample3_finish(const real* wgt, const real* tap, real* out, real* gx, real* gy) {
  const real ax0 = wgt[0], ax1 = wgt[1], ay0 = wgt[2], ay1 = wgt[3];
  const real sx0 = wgt[4], sx1 = wgt[5], sy0 = wgt[6], sy1 = wgt[7];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const real v00 = tap[c * 4], v01 = tap[c * 4 + 1], v10 = tap[c * 4 + 2], v11 = tap[c * 4 + 3];
    const real top = ax0 * v00 + ax1 * v01, bot = ax0 * v10 + ax1 * v11;
    out[c] = ay0 * top + ay1 * bot;
    gx[c] = ay0 * (sx1 * v01 - sx0 * v00) + ay1 * (sx1 * v11 - sx0 * v10);
    gy[c] = ax0 * (sy1 * v10 - sy0 * v00) + ax1 * (sy1 * v11 - sy0 * v01);
  }
}

// ------------------------------------------------------------------------------------------
// The same sample from a packed context image (pack_kernel of loss_sweep.cu): pixels are
// (r, g, b, 0) 4-vectors with a zero border of kPadLo / kPadHi pixels, row pitch Wp.  One tap is one
// 16-byte load; taps outside the image read the border, which is exactly what zeros padding
// contributes to the value and to d/d(ix, iy).  The bilinear form is the factored one
//   top = v00 + wx (v01 - v00), bot = v10 + wx (v11 - v10), out = top + wy (bot - top),
//   d/d iy = bot - top, d/d ix = (v01 - v00) + wy ((v11 - v10) - (v01 - v00)).
// PAD == border clips the coordinate; mx / my are 0 where clipped (zero gradient there).
// ------------------------------------------------------------------------------------------
template <typename real, typename vec4, int PAD>
__device__ __forceinline__ void sample4_issue(const vec4* __restrict__ base, int H, int W, int Wp, int pad_lo, real ix,
                                              real iy, real& wx, real& wy, real& mx, real& my, vec4* tap) {
  mx = real(1); my = real(1);
  if (PAD == PSFM_PAD_BORDER) {
    if (!(ix > real(0))) { ix = real(0); mx = real(0); }
    else if (ix >= real(W - 1)) { ix = real(W - 1); mx = real(0); }
    if (!(iy > real(0))) { iy = real(0); my = real(0); }
    else if (iy >= real(H - 1)) { iy = real(H - 1); my = real(0); }
  } else {
    // fully outside beyond [-1, size]: all four taps land in the zero border
    ix = rmin(rmax(ix, real(-2)), real(W + 1));
    iy = rmin(rmax(iy, real(-2)), real(H + 1));
  }
  const real fx = rfloor(ix), fy = rfloor(iy);
  wx = ix - fx; wy = iy - fy;
  const int o = ((int)fy + pad_lo) * Wp + ((int)fx + pad_lo);
  const vec4* p0 = base + o;
  const vec4* p1 = p0 + Wp;
  tap[0] = ld4(p0); tap[1] = ld4(p0 + 1); tap[2] = ld4(p1); tap[3] = ld4(p1 + 1);
}

template <typename real>
__device__ __forceinline__ void bilerp(real v00, real v01, real v10, real v11, real wx, real wy, real& out, real& gx,
                                       real& gy) {
  const real dxt = v01 - v00, dxb = v11 - v10;
  const real top = dxt * wx + v00, bot = dxb * wx + v10;
  gy = bot - top;
  out = gy * wy + top;
  gx = (dxb - dxt) * wy + dxt;
}

// ------------------------------------------------------------------------------------------
// SSIM distance of one channel, forward and gradient halves (same function as ssim_dist_sums;
// the gradient half takes the selection weight so that no separate scaling pass is needed).
//   fwd: returns clamp((1 - ssim) / 2, 0, 1) and keeps {mx, a1, v1, b1, v2, inv, ssim} in st[7].
//   grad: coef[0..2] = wgt * d dist / d (sum x, sum x^2 (already x2 for d x^2/dx), sum xy); wgt = 0 is
//         safe (every factor is finite, also for a degenerate window of a padding row).
// ------------------------------------------------------------------------------------------
template <typename real>
__device__ __forceinline__ real ssim_fwd(real Sx, real Sxx, real Sxy, real my, real qy, real C1, real C2, real* st) {
  const real r9 = real(1) / real(9);
  const real mx = Sx * r9;
  const real mxy = mx * my;
  const real a1 = real(2) * mxy + C1;
  const real v1 = (Sxy * (real(2) * r9) + (C1 + C2)) - a1;
  const real mx2 = mx * mx;
  const real b1 = mx2 + (my * my + C1);
  const real v2 = (Sxx * r9 + qy) - mx2;
  const real inv = rrcp_fast(b1 * v2);
  const real ssim = (a1 * v1) * inv;
  const real dist = real(0.5) - real(0.5) * ssim;
  st[0] = mx; st[1] = a1; st[2] = v1; st[3] = b1; st[4] = v2; st[5] = inv; st[6] = ssim;
  st[7] = dist;
  return rmin(rmax(dist, real(0)), real(1));
}

template <typename real>
__device__ __forceinline__ void ssim_grad(const real* st, real my, real wgt, real* coef) {
  const real mx = st[0], a1 = st[1], v1 = st[2], b1 = st[3], v2 = st[4], inv = st[5], dist = st[7];
  const bool in = (dist >= real(0)) && (dist <= real(1));
  // d dist = -0.5 d ssim inside the clamp range; act2 = 2 * wgt * (-0.5 / 9) / (b1 v2)
  const real act2 = in ? (wgt * (real(-1) / real(9))) * inv : real(0);
  const real ss = in ? st[6] : real(0);
  coef[0] = act2 * (my * (v1 - a1) - (ss * mx) * (v2 - b1));
  coef[1] = -act2 * (ss * b1);
  coef[2] = act2 * a1;
}

// ------------------------------------------------------------------------------------------
// Value-type generic forms (V = real for one channel, P2<real> for a channel pair; pair_math.cuh)
// ------------------------------------------------------------------------------------------
template <typename real> __device__ __forceinline__ P2<real> shfl_up1(P2<real> v) {
  return P2<real>(shfl_up1(v.lo()), shfl_up1(v.hi()));
}
template <typename real> __device__ __forceinline__ P2<real> shfl_dn1(P2<real> v) {
  return P2<real>(shfl_dn1(v.lo()), shfl_dn1(v.hi()));
}
__device__ __forceinline__ float vrcp_fast(float x) { return rrcp_fast(x); }
__device__ __forceinline__ double vrcp_fast(double x) { return rrcp_fast(x); }
template <typename real> __device__ __forceinline__ P2<real> vrcp_fast(P2<real> x) {
  return P2<real>(rrcp_fast(x.lo()), rrcp_fast(x.hi()));
}
// sum over the lanes of clamp(v, 0, 1)
__device__ __forceinline__ float vclamp01_sum(float v) { return rmin(rmax(v, 0.0f), 1.0f); }
__device__ __forceinline__ double vclamp01_sum(double v) { return rmin(rmax(v, 0.0), 1.0); }
template <typename real> __device__ __forceinline__ real vclamp01_sum(P2<real> v) {
  return rmin(rmax(v.lo(), real(0)), real(1)) + rmin(rmax(v.hi(), real(0)), real(1));
}
// lane-wise (0 <= d <= 1) ? a : 0
__device__ __forceinline__ float vsel_in01(float d, float a) { return (d >= 0.0f && d <= 1.0f) ? a : 0.0f; }
__device__ __forceinline__ double vsel_in01(double d, double a) { return (d >= 0.0 && d <= 1.0) ? a : 0.0; }
template <typename real> __device__ __forceinline__ P2<real> vsel_in01(P2<real> d, P2<real> a) {
  return P2<real>(vsel_in01(d.lo(), a.lo()), vsel_in01(d.hi(), a.hi()));
}
__device__ __forceinline__ float vneg(float a) { return -a; }
__device__ __forceinline__ double vneg(double a) { return -a; }
template <typename real> __device__ __forceinline__ P2<real> vneg(P2<real> a) { return P2<real>(-a.lo(), -a.hi()); }

template <typename V> struct SsimSt { V mx, a1, v1, b1, v2, inv, ssim, dist; };

// forward half: returns the UNclamped distance (1 - ssim) / 2 per lane and keeps the factors
template <typename real, typename V>
__device__ __forceinline__ V ssim_fwd_v(V Sx, V Sxx, V Sxy, V my, V qy, real C1, real C2, SsimSt<V>& st) {
  typedef Bc<V, real> B;
  const V r9 = B::of(real(1) / real(9));
  st.mx = Sx * r9;
  const V mxy = st.mx * my;
  st.a1 = vfma(mxy, B::of(real(2)), B::of(C1));
  st.v1 = vfma(Sxy, B::of(real(2) / real(9)), B::of(C1 + C2)) - st.a1;
  const V mx2 = st.mx * st.mx;
  st.b1 = mx2 + vfma(my, my, B::of(C1));
  st.v2 = vfma(Sxx, r9, qy) - mx2;
  st.inv = vrcp_fast(st.b1 * st.v2);
  st.ssim = (st.a1 * st.v1) * st.inv;
  st.dist = vfma(st.ssim, B::of(real(-0.5)), B::of(real(0.5)));
  return st.dist;
}

// gradient half: coef[0..2] = wgt * d clamp(dist) / d (sum x, sum x^2 (already x2 for d x^2/dx), sum xy)
template <typename real, typename V>
__device__ __forceinline__ void ssim_grad_v(const SsimSt<V>& st, V my, real wgt, V* coef) {
  typedef Bc<V, real> B;
  // inside the clamp range d dist = -0.5 d ssim, outside 0; every factor stays finite (see ssim_dist_sums)
  const V act2 = vsel_in01(st.dist, B::of(wgt * (real(-1) / real(9))) * st.inv);
  const V ssn = vsel_in01(st.dist, vneg(st.ssim));
  const V t1 = my * (st.v1 - st.a1);
  coef[0] = act2 * vfma(ssn * st.mx, st.v2 - st.b1, t1);
  coef[1] = act2 * (ssn * st.b1);
  coef[2] = act2 * st.a1;
}

template <typename V>
__device__ __forceinline__ void bilerp_v(V v00, V v01, V v10, V v11, V wx, V wy, V& out, V& gx, V& gy) {
  const V dxt = v01 - v00, dxb = v11 - v10;
  const V top = vfma(dxt, wx, v00), bot = vfma(dxb, wx, v10);
  gy = bot - top;
  out = vfma(gy, wy, top);
  gx = vfma(dxb - dxt, wy, dxt);
}

}  // namespace psfm
