// px2_math.cuh -- value type of the two-pixels-per-lane sweep (loss_sweep2.cu).
//
// A lane owns two horizontally adjacent pixels; a per-channel quantity of both is held as THREE packed pairs
//   a0 = (channel 0, channel 1) of pixel 0,  a1 = (channel 0, channel 1) of pixel 1,  b = (channel 2 of pixel 0, of pixel 1)
// so that every per-channel operation of the loss is three FADD2 / FMUL2 / FFMA2 for six values (pair_math.cuh: one
// issue slot per packed instruction), and a 3-tap horizontal window needs two shuffles per pair instead of two per
// value: the inner neighbours are in the lane's own registers.  A per-pixel scalar (depth, selection weight, mask) is
// a pair (pixel 0, pixel 1); bc() broadcasts it over the channels.
#pragma once
#include "sweep_device.cuh"

namespace psfm {

template <typename real> struct C3 {
  P2<real> a0, a1, b;
  __device__ __forceinline__ C3() {}
  __device__ __forceinline__ C3(P2<real> x, P2<real> y, P2<real> z) : a0(x), a1(y), b(z) {}
  __device__ __forceinline__ explicit C3(real s) : a0(s), a1(s), b(s) {}
};

template <typename real> __device__ __forceinline__ C3<real> operator+(const C3<real>& x, const C3<real>& y) {
  return C3<real>(x.a0 + y.a0, x.a1 + y.a1, x.b + y.b);
}
template <typename real> __device__ __forceinline__ C3<real> operator-(const C3<real>& x, const C3<real>& y) {
  return C3<real>(x.a0 - y.a0, x.a1 - y.a1, x.b - y.b);
}
template <typename real> __device__ __forceinline__ C3<real> operator*(const C3<real>& x, const C3<real>& y) {
  return C3<real>(x.a0 * y.a0, x.a1 * y.a1, x.b * y.b);
}
template <typename real> __device__ __forceinline__ C3<real> vfma(const C3<real>& x, const C3<real>& y, const C3<real>& z) {
  return C3<real>(vfma(x.a0, y.a0, z.a0), vfma(x.a1, y.a1, z.a1), vfma(x.b, y.b, z.b));
}
// per-pixel pair (pixel 0, pixel 1) broadcast over the channels
template <typename real> __device__ __forceinline__ C3<real> bc(const P2<real>& s) {
  return C3<real>(P2<real>(s.lo()), P2<real>(s.hi()), s);
}
// sum over the three channels, per pixel
template <typename real> __device__ __forceinline__ P2<real> chsum(const C3<real>& v) {
  return P2<real>(v.a0.lo() + v.a0.hi(), v.a1.lo() + v.a1.hi()) + v.b;
}
template <typename real> __device__ __forceinline__ P2<real> pabs(const P2<real>& v) { return P2<real>(rabs(v.lo()), rabs(v.hi())); }
template <typename real> __device__ __forceinline__ C3<real> cabs(const C3<real>& v) { return C3<real>(pabs(v.a0), pabs(v.a1), pabs(v.b)); }
template <typename real> __device__ __forceinline__ P2<real> pclamp01(const P2<real>& v) {
  return P2<real>(rmin(rmax(v.lo(), real(0)), real(1)), rmin(rmax(v.hi(), real(0)), real(1)));
}

// horizontal neighbours of a lane's pixel pair: la = left neighbour of pixel 0 (channels 0, 1), ra = right neighbour of
// pixel 1 (channels 0, 1), lrb = (left neighbour of pixel 0, right neighbour of pixel 1) of channel 2
template <typename real> struct Nb3 { P2<real> la, ra, lrb; };
template <typename real> __device__ __forceinline__ Nb3<real> neighbours(const C3<real>& v) {
  Nb3<real> n;
  n.la = P2<real>(shfl_up1(v.a1.lo()), shfl_up1(v.a1.hi()));
  n.ra = P2<real>(shfl_dn1(v.a0.lo()), shfl_dn1(v.a0.hi()));
  n.lrb = P2<real>(shfl_up1(v.b.hi()), shfl_dn1(v.b.lo()));
  return n;
}
// 3-tap horizontal sums of v and of v^2
template <typename real> __device__ __forceinline__ C3<real> hsum(const C3<real>& v, const Nb3<real>& n) {
  const P2<real> s = v.a0 + v.a1;
  const real sb = v.b.lo() + v.b.hi();
  return C3<real>(n.la + s, s + n.ra, n.lrb + P2<real>(sb));
}
template <typename real> __device__ __forceinline__ C3<real> hsumsq(const C3<real>& v, const Nb3<real>& n) {
  const P2<real> q = vfma(v.a1, v.a1, v.a0 * v.a0);
  const real qb = vfma(v.b.hi(), v.b.hi(), v.b.lo() * v.b.lo());
  return C3<real>(vfma(n.la, n.la, q), vfma(n.ra, n.ra, q), vfma(n.lrb, n.lrb, P2<real>(qb)));
}

template <typename real> struct Ssim3 { SsimSt<P2<real> > a0, a1, b; };

template <typename real>
__device__ __forceinline__ C3<real> ssim_fwd3(const C3<real>& Sx, const C3<real>& Sxx, const C3<real>& Sxy, const C3<real>& my,
                                              const C3<real>& qy, real C1, real C2, Ssim3<real>& st) {
  return C3<real>(ssim_fwd_v<real, P2<real> >(Sx.a0, Sxx.a0, Sxy.a0, my.a0, qy.a0, C1, C2, st.a0),
                  ssim_fwd_v<real, P2<real> >(Sx.a1, Sxx.a1, Sxy.a1, my.a1, qy.a1, C1, C2, st.a1),
                  ssim_fwd_v<real, P2<real> >(Sx.b, Sxx.b, Sxy.b, my.b, qy.b, C1, C2, st.b));
}

// gradient half with a lane-wise weight (sweep_device.cuh: ssim_grad_v takes one scalar)
template <typename real, typename V>
__device__ __forceinline__ void ssim_grad_w(const SsimSt<V>& st, V my, V wgt9 /* weight * (-1/9) */, V* coef) {
  const V act2 = vsel_in01(st.dist, wgt9 * st.inv);
  const V ssn = vsel_in01(st.dist, vneg(st.ssim));
  const V t1 = my * (st.v1 - st.a1);
  coef[0] = act2 * vfma(ssn * st.mx, st.v2 - st.b1, t1);
  coef[1] = act2 * (ssn * st.b1);
  coef[2] = act2 * st.a1;
}

}  // namespace psfm
