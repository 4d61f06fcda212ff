// pair_math.cuh -- two values of the arithmetic type processed by ONE instruction.
//
// sm_100a has packed fp32 arithmetic (FADD2 / FMUL2 / FFMA2 on an aligned register pair).  Measured on
// B200 (tools/ubench): a packed instruction occupies the FMA pipe for two issue cycles, so it adds no FP32
// throughput, but it takes ONE issue slot and has the latency of a scalar FFMA (4.5 cycles dependent).
// The sweep kernel is bound by issue slots and dependent-issue latency at two warps per scheduler, not
// by the FMA pipe, so its per-channel arithmetic runs on (channel 0, channel 1) pairs + a scalar channel 2.
// The double instantiation (parity testing) is two scalar operations per pair.
#pragma once
#include <cuda_runtime.h>

namespace psfm {

template <typename real> struct P2;

template <> struct P2<float> {
  float2 v;
  __device__ __forceinline__ P2() {}
  __device__ __forceinline__ P2(float a, float b) { v.x = a; v.y = b; }
  __device__ __forceinline__ explicit P2(float s) { v.x = s; v.y = s; }
  __device__ __forceinline__ float lo() const { return v.x; }
  __device__ __forceinline__ float hi() const { return v.y; }
};
template <> struct P2<double> {
  double a, b;
  __device__ __forceinline__ P2() {}
  __device__ __forceinline__ P2(double x, double y) : a(x), b(y) {}
  __device__ __forceinline__ explicit P2(double s) : a(s), b(s) {}
  __device__ __forceinline__ double lo() const { return a; }
  __device__ __forceinline__ double hi() const { return b; }
};

// ---- float: packed instructions ------------------------------------------------------------------
__device__ __forceinline__ P2<float> operator+(P2<float> x, P2<float> y) { P2<float> r; r.v = __fadd2_rn(x.v, y.v); return r; }
__device__ __forceinline__ P2<float> operator*(P2<float> x, P2<float> y) { P2<float> r; r.v = __fmul2_rn(x.v, y.v); return r; }
__device__ __forceinline__ P2<float> operator-(P2<float> x, P2<float> y) {
  P2<float> r;
  unsigned long long a = *reinterpret_cast<unsigned long long*>(&x.v), b = *reinterpret_cast<unsigned long long*>(&y.v), c;
  asm("sub.f32x2 %0, %1, %2;" : "=l"(c) : "l"(a), "l"(b));
  r.v = *reinterpret_cast<float2*>(&c);
  return r;
}
__device__ __forceinline__ P2<float> vfma(P2<float> x, P2<float> y, P2<float> z) { P2<float> r; r.v = __ffma2_rn(x.v, y.v, z.v); return r; }
// ---- double: two scalar operations ---------------------------------------------------------------
__device__ __forceinline__ P2<double> operator+(P2<double> x, P2<double> y) { return P2<double>(x.a + y.a, x.b + y.b); }
__device__ __forceinline__ P2<double> operator*(P2<double> x, P2<double> y) { return P2<double>(x.a * y.a, x.b * y.b); }
__device__ __forceinline__ P2<double> operator-(P2<double> x, P2<double> y) { return P2<double>(x.a - y.a, x.b - y.b); }
__device__ __forceinline__ P2<double> vfma(P2<double> x, P2<double> y, P2<double> z) {
  return P2<double>(fma(x.a, y.a, z.a), fma(x.b, y.b, z.b));
}
// ---- scalars through the same names, so that one template serves a pair and a single channel -------
__device__ __forceinline__ float vfma(float x, float y, float z) { return fmaf(x, y, z); }
__device__ __forceinline__ double vfma(double x, double y, double z) { return fma(x, y, z); }

// broadcast of a scalar to the value type V (identity for a scalar V)
template <typename V, typename real> struct Bc;
template <typename real> struct Bc<P2<real>, real> { static __device__ __forceinline__ P2<real> of(real s) { return P2<real>(s); } };
template <typename real> struct Bc<real, real> { static __device__ __forceinline__ real of(real s) { return s; } };

// lane-wise helpers on pairs
template <typename real, typename F> __device__ __forceinline__ P2<real> pmap(P2<real> x, F f) { return P2<real>(f(x.lo()), f(x.hi())); }

}  // namespace psfm
