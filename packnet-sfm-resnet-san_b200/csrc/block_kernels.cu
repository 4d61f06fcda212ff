// block_kernels.cu -- the reference's building-block functions of the photometric-loss path as
// stand-alone operators (each with its backward), for callers that use them directly instead of
// MultiViewPhotometricLoss.forward:
//   reconstruct / project      Camera.reconstruct, Camera.project        (geometry/camera.py:111-190)
//   transform_points           Pose.transform_points                      (geometry/pose.py:80-86)
//   ssim / photometric map     SSIM, calc_photometric_loss                (losses/multiview_photometric_loss.py:15-54,199-267)
//   normalize / smoothness     inv_depths_normalize, calc_smoothness      (utils/depth.py:146-198, utils/image.py:85-113)
//   reductions                 reduce_photometric_loss, |x|.mean()        (losses/...:269-297,301-327)
// All tensors are contiguous `real` (fp32 or fp64).  Reductions are deterministic (block partials
// in fixed slots + a fixed-order finish).
#include <cuda_runtime.h>
#include "psfm_device.cuh"
#include "profile.h"

namespace psfm {

constexpr int kBT = 256;

static inline unsigned nblk_for(size_t n) { return (unsigned)((n + kBT - 1) / kBT); }

// lifting matrix of the reference (`Kinv`, camera.py:73-81) of batch item b
template <typename real> __device__ __forceinline__ void load_kinv(const real* K, int b, real* ki) {
#pragma unroll
  for (int i = 0; i < 9; ++i) ki[i] = K[b * 9 + i];
  const real fx = ki[0], fy = ki[4], cx = ki[2], cy = ki[5];
  ki[0] = real(1) / fx;
  ki[4] = real(1) / fy;
  ki[2] = real(-1) * cx / fx;
  ki[5] = real(-1) * cy / fy;
}

// [B, nblk, 12] partials -> d mat [B,4,4] (rows 0-2), fixed order
template <typename real>
__global__ void __launch_bounds__(32) mat_finish_kernel(const real* part, int nblk, real* d_mat) {
  const int b = blockIdx.x, t = threadIdx.x;
  if (t < 12) {
    double a = 0.0;
    for (int i = 0; i < nblk; ++i) a += (double)part[((size_t)b * nblk + i) * 12 + t];
    d_mat[(size_t)b * 16 + t] = (real)a;
  } else if (t < 16) {
    d_mat[(size_t)b * 16 + t] = real(0);
  }
}

// ------------------------------------------------------------------------------------------
// Camera.reconstruct: X = [R|t] (Kinv (u, v, 1) * depth), [R|t] = Twc rows or identity when null
// ------------------------------------------------------------------------------------------
template <typename real>
__global__ void __launch_bounds__(kBT) reconstruct_kernel(int H, int W, const real* depth, const real* K,
                                                          const real* Twc, real* out) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  real ki[9];
  load_kinv<real>(K, b, ki);
  const real u = (real)(int)(i % W), v = (real)(int)(i / W), d = depth[b * plane + i];
  real X[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) X[k] = (ki[k * 3] * u + ki[k * 3 + 1] * v + ki[k * 3 + 2]) * d;
  if (Twc) {
    const real* T = Twc + (size_t)b * 16;
    const real x0 = X[0], x1 = X[1], x2 = X[2];
#pragma unroll
    for (int k = 0; k < 3; ++k) X[k] = T[k * 4] * x0 + T[k * 4 + 1] * x1 + T[k * 4 + 2] * x2 + T[k * 4 + 3];
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) out[((size_t)b * 3 + k) * plane + i] = X[k];
}

template <typename real>
__global__ void __launch_bounds__(kBT) reconstruct_bwd_kernel(int H, int W, const real* K, const real* Twc,
                                                              const real* gout, real* d_depth) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  real ki[9];
  load_kinv<real>(K, b, ki);
  const real u = (real)(int)(i % W), v = (real)(int)(i / W);
  real g[3], ray[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    g[k] = gout[((size_t)b * 3 + k) * plane + i];
    ray[k] = ki[k * 3] * u + ki[k * 3 + 1] * v + ki[k * 3 + 2];
  }
  if (Twc) {   // d/d Xc = R^T g
    const real* T = Twc + (size_t)b * 16;
    const real g0 = g[0], g1 = g[1], g2 = g[2];
#pragma unroll
    for (int k = 0; k < 3; ++k) g[k] = T[k] * g0 + T[4 + k] * g1 + T[8 + k] * g2;
  }
  d_depth[b * plane + i] = ray[0] * g[0] + ray[1] * g[1] + ray[2] * g[2];
}

// ------------------------------------------------------------------------------------------
// Camera.project: p = K ([R|t] X); Z = max(p2, 1e-5); out[b, v, u] = (2 (p0/Z)/(W-1) - 1, 2 (p1/Z)/(H-1) - 1)
// ------------------------------------------------------------------------------------------
template <typename real>
__device__ __forceinline__ void project_point(const real* K, const real* Tcw, int b, const real* X, real* Xc, real* p) {
  if (Tcw) {
    const real* T = Tcw + (size_t)b * 16;
#pragma unroll
    for (int k = 0; k < 3; ++k) Xc[k] = T[k * 4] * X[0] + T[k * 4 + 1] * X[1] + T[k * 4 + 2] * X[2] + T[k * 4 + 3];
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) Xc[k] = X[k];
  }
  const real* Kb = K + b * 9;
#pragma unroll
  for (int k = 0; k < 3; ++k) p[k] = Kb[k * 3] * Xc[0] + Kb[k * 3 + 1] * Xc[1] + Kb[k * 3 + 2] * Xc[2];
}

template <typename real>
__global__ void __launch_bounds__(kBT) project_kernel(int H, int W, const real* X, const real* K, const real* Tcw,
                                                      real* out) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  real x[3], Xc[3], p[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) x[k] = X[((size_t)b * 3 + k) * plane + i];
  project_point<real>(K, Tcw, b, x, Xc, p);
  const real Z = rmax(p[2], real(1e-5));
  out[(b * plane + i) * 2] = real(2) * (p[0] / Z) / real(W - 1) - real(1);
  out[(b * plane + i) * 2 + 1] = real(2) * (p[1] / Z) / real(H - 1) - real(1);
}

template <typename real>
__global__ void __launch_bounds__(kBT) project_bwd_kernel(int H, int W, const real* X, const real* K, const real* Tcw,
                                                          const real* gout, real* dX, real* part) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  __shared__ real scratch[12 * (kBT / 32)];
  real acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = real(0);
  if (i < plane) {
    real x[3], Xc[3], p[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) x[k] = X[((size_t)b * 3 + k) * plane + i];
    project_point<real>(K, Tcw, b, x, Xc, p);
    const real Z = rmax(p[2], real(1e-5));
    const real gxn = gout[(b * plane + i) * 2], gyn = gout[(b * plane + i) * 2 + 1];
    real dp[3];
    dp[0] = gxn * real(2) / real(W - 1) / Z;
    dp[1] = gyn * real(2) / real(H - 1) / Z;
    dp[2] = (p[2] >= real(1e-5)) ? -(dp[0] * p[0] + dp[1] * p[1]) / Z : real(0);
    const real* Kb = K + b * 9;
    real dXc[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) dXc[k] = Kb[k] * dp[0] + Kb[3 + k] * dp[1] + Kb[6 + k] * dp[2];
    if (Tcw) {
      const real* T = Tcw + (size_t)b * 16;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        dX[((size_t)b * 3 + k) * plane + i] = T[k] * dXc[0] + T[4 + k] * dXc[1] + T[8 + k] * dXc[2];
        acc[k * 4] = dXc[k] * x[0]; acc[k * 4 + 1] = dXc[k] * x[1]; acc[k * 4 + 2] = dXc[k] * x[2]; acc[k * 4 + 3] = dXc[k];
      }
    } else {
#pragma unroll
      for (int k = 0; k < 3; ++k) dX[((size_t)b * 3 + k) * plane + i] = dXc[k];
    }
  }
  if (part) {
    block_reduce<real, 12>(acc, scratch);
    if (threadIdx.x < 12) part[((size_t)b * gridDim.x + blockIdx.x) * 12 + threadIdx.x] = acc[threadIdx.x];
  }
}

// ------------------------------------------------------------------------------------------
// FisheyeCamera (VADAS), frame 'c' (reference geometry/camera.py:194-394).  intr: [B,11] rows
// (k0..k6, s, div, ux, uy).  reconstruct: theta = r_d is the reference's placeholder inverse of the
// distortion polynomial (:278-283); project: r_d = sum k_i atan(r)^i (:352-383).
// ------------------------------------------------------------------------------------------
template <typename real> __device__ __forceinline__ real fe_eps() { return (real)2.220446049250313e-16; }   // sys.float_info.epsilon

template <typename real>
__device__ __forceinline__ void fisheye_ray(const real* fi, real u, real v, real& rx, real& ry) {
  const real xd = (u - fi[9]) / fi[7], yd = (v - fi[10]) / fi[8];
  const real rd = sqrt(xd * xd + yd * yd);
  const real sc = tan(rd) / (rd < fe_eps<real>() ? fe_eps<real>() : rd);
  rx = sc * xd; ry = sc * yd;
}

template <typename real>
__global__ void __launch_bounds__(kBT) fisheye_reconstruct_kernel(int H, int W, const real* depth, const real* intr, real* out) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  real rx, ry;
  fisheye_ray<real>(intr + (size_t)b * 11, (real)(int)(i % W), (real)(int)(i / W), rx, ry);
  const real d = depth[b * plane + i];
  out[((size_t)b * 3) * plane + i] = rx * d;
  out[((size_t)b * 3 + 1) * plane + i] = ry * d;
  out[((size_t)b * 3 + 2) * plane + i] = d;
}

template <typename real>
__global__ void __launch_bounds__(kBT) fisheye_reconstruct_bwd_kernel(int H, int W, const real* intr, const real* gout,
                                                                      real* d_depth) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  real rx, ry;
  fisheye_ray<real>(intr + (size_t)b * 11, (real)(int)(i % W), (real)(int)(i / W), rx, ry);
  d_depth[b * plane + i] = rx * gout[((size_t)b * 3) * plane + i] + ry * gout[((size_t)b * 3 + 1) * plane + i] +
                           gout[((size_t)b * 3 + 2) * plane + i];
}

// pixel coordinates (px, py) of the camera-frame point X; when J: J[0..2] = d px / d X, J[3..5] = d py / d X.
// torch's gradient of sqrt at r = 0 is undefined (inf * 0); here it is 0.
template <typename real>
__device__ __forceinline__ void fisheye_pixel(const real* fi, const real* X, real& px, real& py, real* J) {
  const real eps = fe_eps<real>();
  const real Z = rmax(X[2], eps);
  const real xn = X[0] / Z, yn = X[1] / Z;
  const real r2 = xn * xn + yn * yn, r = sqrt(r2);
  const real th = atan(r);
  real poly = fi[6], dpoly = real(6) * fi[6];
#pragma unroll
  for (int i = 5; i >= 1; --i) { poly = poly * th + fi[i]; dpoly = dpoly * th + (real)i * fi[i]; }
  poly = poly * th + fi[0];
  const bool tiny = r < eps;
  const real g = poly / (tiny ? eps : r);
  px = fi[7] * (g * xn) + fi[9];
  py = fi[8] * (g * yn) + fi[10];
  if (J) {
    const real h = tiny ? real(0) : (dpoly / (real(1) + r2) - g) / r2;      // (d g / d r) / r
    const real a00 = fi[7] * (g + h * xn * xn), a01 = fi[7] * h * xn * yn, a10 = fi[8] * h * xn * yn,
               a11 = fi[8] * (g + h * yn * yn);
    const real zk = (X[2] >= eps) ? real(-1) / Z : real(0);
    J[0] = a00 / Z; J[1] = a01 / Z; J[2] = (a00 * xn + a01 * yn) * zk;
    J[3] = a10 / Z; J[4] = a11 / Z; J[5] = (a10 * xn + a11 * yn) * zk;
  }
}

// Hn, Wn: the size the coordinates are normalised with (the map itself, or FisheyeCamera.image_size for [B,3,N] clouds)
template <typename real>
__global__ void __launch_bounds__(kBT) fisheye_project_kernel(size_t N, int Hn, int Wn, const real* X, const real* intr, real* out) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= N) return;
  real x[3], px, py;
#pragma unroll
  for (int k = 0; k < 3; ++k) x[k] = X[((size_t)b * 3 + k) * N + i];
  fisheye_pixel<real>(intr + (size_t)b * 11, x, px, py, nullptr);
  out[(b * N + i) * 2] = real(2) * px / real(Wn - 1) - real(1);
  out[(b * N + i) * 2 + 1] = real(2) * py / real(Hn - 1) - real(1);
}

template <typename real>
__global__ void __launch_bounds__(kBT) fisheye_project_bwd_kernel(size_t N, int Hn, int Wn, const real* X, const real* intr,
                                                                  const real* gout, real* dX) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= N) return;
  real x[3], px, py, J[6];
#pragma unroll
  for (int k = 0; k < 3; ++k) x[k] = X[((size_t)b * 3 + k) * N + i];
  fisheye_pixel<real>(intr + (size_t)b * 11, x, px, py, J);
  const real gx = gout[(b * N + i) * 2] * real(2) / real(Wn - 1), gy = gout[(b * N + i) * 2 + 1] * real(2) / real(Hn - 1);
#pragma unroll
  for (int k = 0; k < 3; ++k) dX[((size_t)b * 3 + k) * N + i] = gx * J[k] + gy * J[3 + k];
}

// ------------------------------------------------------------------------------------------
// grid_sample(image [B,3,Hi,Wi], grid [B,Ho,Wo,2], bilinear, align_corners=True, zeros | border)
// (the reference's call: geometry/camera_utils.py:58-59).  Backward: d grid.
// ------------------------------------------------------------------------------------------
template <typename real>
__global__ void __launch_bounds__(kBT) grid_sample_kernel(int Hi, int Wi, size_t No, int pad, const real* img, const real* grid,
                                                          real* out) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= No) return;
  const real ix = (grid[(b * No + i) * 2] + real(1)) * real(0.5) * real(Wi - 1);
  const real iy = (grid[(b * No + i) * 2 + 1] + real(1)) * real(0.5) * real(Hi - 1);
  real v[3];
  bilinear3<real, real, false>(img + (size_t)b * 3 * Hi * Wi, (size_t)Hi * Wi, Hi, Wi, pad, ix, iy, v, nullptr, nullptr);
#pragma unroll
  for (int c = 0; c < 3; ++c) out[((size_t)b * 3 + c) * No + i] = v[c];
}

template <typename real>
__global__ void __launch_bounds__(kBT) grid_sample_bwd_kernel(int Hi, int Wi, size_t No, int pad, const real* img,
                                                              const real* grid, const real* gout, real* d_grid) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= No) return;
  const real ix = (grid[(b * No + i) * 2] + real(1)) * real(0.5) * real(Wi - 1);
  const real iy = (grid[(b * No + i) * 2 + 1] + real(1)) * real(0.5) * real(Hi - 1);
  real v[3], gx[3], gy[3];
  bilinear3<real, real, true>(img + (size_t)b * 3 * Hi * Wi, (size_t)Hi * Wi, Hi, Wi, pad, ix, iy, v, gx, gy);
  real ax = real(0), ay = real(0);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const real g = gout[((size_t)b * 3 + c) * No + i];
    ax += g * gx[c]; ay += g * gy[c];
  }
  d_grid[(b * No + i) * 2] = ax * real(0.5) * real(Wi - 1);
  d_grid[(b * No + i) * 2 + 1] = ay * real(0.5) * real(Hi - 1);
}

// ------------------------------------------------------------------------------------------
// SupervisedLoss.calculate_loss, one scale (reference losses/supervised_loss.py:292-417) with the loss functions
// nn.L1Loss / nn.MSELoss / SilogLoss (:67-82, :84-92) and SSISilogLoss (losses/ssi_silog_loss.py:53-429).
//   sparse methods: valid = (gt > 0) [& mask]; pred and gt are masked_fill'ed with 1e-6 where not valid, and the
//   loss function then averages over ALL N elements (:323-345); dense methods use every element as it is.
//   l1: mean |p - g|;  mse: mean (p - g)^2;  silog: sqrt(mean(l^2) - ratio2 * mean(l)^2), l = log p - log g.
//   ssi-silog: means over the M valid elements only; ssi = var(d) + alpha mean(d)^2 on inverse depths (:176-197),
//   silog = sqrt(|mean(l^2) - ratio2 mean(l)^2| + 1e-8) with l = log P - log G on depths P = 1 / max(p, 1e-6) clamped to
//   [cmin, cmax] (:218-262); loss = w_ssi ssi + w_silog silog (:413-415); 0 when M < 100 (:398-400), 1 (constant) when
//   a NaN shows up (:403-408).  The multi-scale gradient term (gradient_weight, default 0) is not provided.
// hp (host, double[6]): {ratio2, alpha, w_ssi, w_silog, cmin, cmax}.
// state (double[16], caller-owned, forward -> backward):
//   {sum |d|, sum d^2, sum l, sum l^2, valid count M, sum d, N, loss, ssi, silog, mean d, var d, mean l^2, ratio2 mean(l)^2,
//    sign of the silog radicand, 1 if the loss is a constant (no gradient)}.
// ------------------------------------------------------------------------------------------
constexpr int kSupVals = 6;
struct SupHp { double ratio2, alpha, w_ssi, w_silog, cmin, cmax; };

// log-difference of the method; dl = d l / d p
template <typename real>
__device__ __forceinline__ real sup_logdiff(int method, real p, real g, real cmin, real cmax, real& dl) {
  if (method == PSFM_SUP_SSI_SILOG) {
    const real P0 = real(1) / rmax(p, real(1e-6)), G0 = real(1) / rmax(g, real(1e-6));      // inv2depth
    const real P = rmin(rmax(P0, cmin), cmax), G = rmin(rmax(G0, cmin), cmax);
    dl = (p >= real(1e-6) && P0 >= cmin && P0 <= cmax) ? real(-1) / p : real(0);
    return log(P) - log(G);
  }
  dl = real(1) / p;
  return log(p) - log(g);
}

// one element's contribution to the six sums
template <typename real>
__device__ __forceinline__ void sup_accumulate(int method, int sparse, real p, real g, bool masked_out, const SupHp& hp, real* v) {
  const bool valid = !(sparse || method == PSFM_SUP_SSI_SILOG) || (g > real(0) && !masked_out);
  if (!valid) { p = real(1e-6); g = real(1e-6); }
  const real d = p - g;
  v[0] += rabs(d);
  v[1] += d * d;
  if (method >= PSFM_SUP_SILOG && (valid || method == PSFM_SUP_SILOG)) {
    real dl;
    const real l = sup_logdiff<real>(method, p, g, (real)hp.cmin, (real)hp.cmax, dl);
    v[2] += l; v[3] += l * l;
  }
  v[4] += valid ? real(1) : real(0);
  v[5] += d;
}

// 16-byte vector loads (4 floats / 2 doubles per thread and iteration), grid-stride; the grid is a fixed function of N
// (sup_grid), so the order of every sum is fixed.  The up-to-3 tail elements go to thread 0 of block 0.
template <typename real>
__global__ void __launch_bounds__(kBT) sup_sums_kernel(size_t N, int method, int sparse, const real* pred, const real* gt,
                                                       const uint8_t* mask, SupHp hp, double* part) {
  constexpr int V = 16 / sizeof(real);
  __shared__ real scratch[kSupVals * (kBT / 32)];
  real v[kSupVals];
#pragma unroll
  for (int k = 0; k < kSupVals; ++k) v[k] = real(0);
  const size_t nvec = N / V;
#pragma unroll 2
  for (size_t q = (size_t)blockIdx.x * kBT + threadIdx.x; q < nvec; q += (size_t)gridDim.x * kBT) {
    alignas(16) real pv[V], gv[V];
    *reinterpret_cast<int4*>(pv) = __ldg(reinterpret_cast<const int4*>(pred) + q);
    *reinterpret_cast<int4*>(gv) = __ldg(reinterpret_cast<const int4*>(gt) + q);
#pragma unroll
    for (int e = 0; e < V; ++e) sup_accumulate<real>(method, sparse, pv[e], gv[e], mask && mask[q * V + e] == 0, hp, v);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (size_t i = nvec * V; i < N; ++i) sup_accumulate<real>(method, sparse, pred[i], gt[i], mask && mask[i] == 0, hp, v);
  block_reduce<real, kSupVals>(v, scratch);
  if (threadIdx.x < kSupVals) part[(size_t)blockIdx.x * kSupVals + threadIdx.x] = (double)v[threadIdx.x];
}

template <typename real>
__global__ void __launch_bounds__(256) sup_finish_kernel(const double* part, int nblk, size_t N, int method, SupHp hp,
                                                         double* state, real* loss_out) {
  __shared__ double sh[kSupVals][256];
  double a[kSupVals];
#pragma unroll
  for (int k = 0; k < kSupVals; ++k) a[k] = 0.0;
  for (int i = threadIdx.x; i < nblk; i += 256)
#pragma unroll
    for (int k = 0; k < kSupVals; ++k) a[k] += part[(size_t)i * kSupVals + k];
#pragma unroll
  for (int k = 0; k < kSupVals; ++k) sh[k][threadIdx.x] = a[k];
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o)
#pragma unroll
      for (int k = 0; k < kSupVals; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double n = (double)N;
    double loss, st[8] = {0, 0, 0, 0, 0, 0, 1.0, 0};   // ssi, silog, mean d, var d, silog1, silog2, sign, constant
    if (method == PSFM_SUP_L1) loss = sh[0][0] / n;
    else if (method == PSFM_SUP_MSE) loss = sh[1][0] / n;
    else if (method == PSFM_SUP_SILOG) { const double m = sh[2][0] / n; loss = sqrt(sh[3][0] / n - hp.ratio2 * m * m); }
    else {
      const double M = sh[4][0];
      const double md = sh[5][0] / M, var = sh[1][0] / M - md * md;
      const double ml = sh[2][0] / M, s1 = sh[3][0] / M, s2 = hp.ratio2 * ml * ml;
      double rad = s1 - s2;
      st[6] = rad < 0.0 ? -1.0 : 1.0;
      rad = rad < 0.0 ? -rad : rad;
      st[0] = var + hp.alpha * md * md; st[1] = sqrt(rad + 1e-8);
      st[2] = md; st[3] = var; st[4] = s1; st[5] = s2;
      loss = hp.w_ssi * st[0] + hp.w_silog * st[1];
      if (M < 100.0) { loss = 0.0; st[7] = 1.0; }
      else if (loss != loss) { loss = 1.0; st[7] = 1.0; }
    }
#pragma unroll
    for (int k = 0; k < kSupVals; ++k) state[k] = sh[k][0];
    state[6] = n; state[7] = loss;
#pragma unroll
    for (int k = 0; k < 8; ++k) state[8 + k] = st[k];
    loss_out[0] = (real)loss;
  }
}

template <typename real>
__device__ __forceinline__ real sup_grad_elem(int method, int sparse, real p, real g, bool masked_out, const SupHp& hp,
                                              const double* state, real up, bool live) {
  const bool valid = !(sparse || method == PSFM_SUP_SSI_SILOG) || (g > real(0) && !masked_out);
  if (!valid || !live) return real(0);           // masked_fill / masked selection: no gradient reaches an invalid element
  const real n = (real)state[6];
  const real d = p - g;
  if (method == PSFM_SUP_L1) return up * rsign(d) / n;
  if (method == PSFM_SUP_MSE) return up * real(2) * d / n;
  real dl;
  const real l = sup_logdiff<real>(method, p, g, (real)hp.cmin, (real)hp.cmax, dl);
  if (method == PSFM_SUP_SILOG) {
    const real m = (real)(state[2] / state[6]), loss = (real)state[7];
    return up * (l - (real)hp.ratio2 * m) / (n * loss) * dl;
  }
  const real M = (real)state[4], md = (real)state[10], ml = (real)(state[2] / state[4]);
  const real g_ssi = real(2) * (d - (real)(1.0 - hp.alpha) * md) / M;
  const real g_sil = (real)state[14] * (l - (real)hp.ratio2 * ml) / (M * (real)state[9]) * dl;
  return up * ((real)hp.w_ssi * g_ssi + (real)hp.w_silog * g_sil);
}

template <typename real>
__global__ void __launch_bounds__(kBT) sup_grad_kernel(size_t N, int method, int sparse, const real* pred, const real* gt,
                                                       const uint8_t* mask, SupHp hp, const double* state,
                                                       const real* grad_loss, real* d_pred) {
  constexpr int V = 16 / sizeof(real);
  const bool live = state[15] == 0.0;
  const real up = grad_loss[0];
  const size_t nvec = N / V;
#pragma unroll 2
  for (size_t q = (size_t)blockIdx.x * kBT + threadIdx.x; q < nvec; q += (size_t)gridDim.x * kBT) {
    alignas(16) real pv[V], gv[V], ov[V];
    *reinterpret_cast<int4*>(pv) = __ldg(reinterpret_cast<const int4*>(pred) + q);
    *reinterpret_cast<int4*>(gv) = __ldg(reinterpret_cast<const int4*>(gt) + q);
#pragma unroll
    for (int e = 0; e < V; ++e)
      ov[e] = sup_grad_elem<real>(method, sparse, pv[e], gv[e], mask && mask[q * V + e] == 0, hp, state, up, live);
    reinterpret_cast<int4*>(d_pred)[q] = *reinterpret_cast<const int4*>(ov);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (size_t i = nvec * V; i < N; ++i)
      d_pred[i] = sup_grad_elem<real>(method, sparse, pred[i], gt[i], mask && mask[i] == 0, hp, state, up, live);
}

// ------------------------------------------------------------------------------------------
// DualHeadDepthLoss.forward (reference losses/dual_head_depth_loss.py:23-201; decompose_depth / dual_head_to_depth,
// networks/layers/resnet/layers.py:76-149): valid = min_depth < gt < max_depth; over the M valid pixels
//   integer = mean |ip - floor(gt) / max_depth|, fractional = mean |fp - (gt - floor gt)|,
//   consistency = mean |ip max_depth + fp - gt|, loss = w_i integer + w_f fractional + w_c consistency (0 when M = 0).
// hp (host double[5]): {max_depth, min_depth, w_i, w_f, w_c}.
// state (device double[8]): {sum_i, sum_f, sum_c, M, #(|integer error| < 1 m), sum frac_err^2, 0, 1 if M = 0}.
// ------------------------------------------------------------------------------------------
constexpr int kDualVals = 6;
struct DualHp { double max_depth, min_depth, wi, wf, wc; };

template <typename real>
__device__ __forceinline__ bool dual_terms(real ip, real fp, real gt, const DualHp& hp, real& ei, real& ef, real& ec) {
  const real mx = (real)hp.max_depth;
  const bool valid = gt > (real)hp.min_depth && gt < mx;
  const real fl = floor(gt);
  ei = ip - fl / mx;
  ef = fp - (gt - fl);
  ec = (ip * mx + fp) - gt;
  return valid;
}

template <typename real>
__global__ void __launch_bounds__(kBT) dual_sums_kernel(size_t N, const real* ip, const real* fp, const real* gt, DualHp hp,
                                                        double* part) {
  __shared__ real scratch[kDualVals * (kBT / 32)];
  real v[kDualVals];
#pragma unroll
  for (int k = 0; k < kDualVals; ++k) v[k] = real(0);
#pragma unroll 4
  for (size_t i = (size_t)blockIdx.x * kBT + threadIdx.x; i < N; i += (size_t)gridDim.x * kBT) {
    real ei, ef, ec;
    if (dual_terms<real>(__ldg(ip + i), __ldg(fp + i), __ldg(gt + i), hp, ei, ef, ec)) {
      v[0] += rabs(ei); v[1] += rabs(ef); v[2] += rabs(ec); v[3] += real(1);
      v[4] += (rabs(ei * (real)hp.max_depth) < real(1)) ? real(1) : real(0);
      v[5] += ef * ef;
    }
  }
  block_reduce<real, kDualVals>(v, scratch);
  if (threadIdx.x < kDualVals) part[(size_t)blockIdx.x * kDualVals + threadIdx.x] = (double)v[threadIdx.x];
}

template <typename real>
__global__ void __launch_bounds__(256) dual_finish_kernel(const double* part, int nblk, DualHp hp, double* state, real* loss_out) {
  __shared__ double sh[kDualVals][256];
  double a[kDualVals];
#pragma unroll
  for (int k = 0; k < kDualVals; ++k) a[k] = 0.0;
  for (int i = threadIdx.x; i < nblk; i += 256)
#pragma unroll
    for (int k = 0; k < kDualVals; ++k) a[k] += part[(size_t)i * kDualVals + k];
#pragma unroll
  for (int k = 0; k < kDualVals; ++k) sh[k][threadIdx.x] = a[k];
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o)
#pragma unroll
      for (int k = 0; k < kDualVals; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double M = sh[3][0];
    const bool none = M == 0.0;
    const double li = none ? 0.0 : sh[0][0] / M, lf = none ? 0.0 : sh[1][0] / M, lc = none ? 0.0 : sh[2][0] / M;
#pragma unroll
    for (int k = 0; k < kDualVals; ++k) state[k] = sh[k][0];
    state[6] = 0.0; state[7] = none ? 1.0 : 0.0;
    loss_out[0] = (real)(hp.wi * li + hp.wf * lf + hp.wc * lc);
    loss_out[1] = (real)li; loss_out[2] = (real)lf; loss_out[3] = (real)lc;
  }
}

template <typename real>
__global__ void __launch_bounds__(kBT) dual_grad_kernel(size_t N, const real* ip, const real* fp, const real* gt, DualHp hp,
                                                        const double* state, const real* grad_loss, real* d_ip, real* d_fp) {
  const bool live = state[7] == 0.0;
  const real up = live ? grad_loss[0] / (real)state[3] : real(0);
#pragma unroll 4
  for (size_t i = (size_t)blockIdx.x * kBT + threadIdx.x; i < N; i += (size_t)gridDim.x * kBT) {
    real ei, ef, ec, gi = real(0), gf = real(0);
    if (dual_terms<real>(__ldg(ip + i), __ldg(fp + i), __ldg(gt + i), hp, ei, ef, ec) && live) {
      const real sc = rsign(ec) * (real)hp.wc;
      gi = up * ((real)hp.wi * rsign(ei) + sc * (real)hp.max_depth);
      gf = up * ((real)hp.wf * rsign(ef) + sc);
    }
    d_ip[i] = gi; d_fp[i] = gf;
  }
}

// ------------------------------------------------------------------------------------------
// Pose.transform_points: out = R X + t on [B,3,N]
// ------------------------------------------------------------------------------------------
template <typename real>
__global__ void __launch_bounds__(kBT) transform_kernel(size_t N, const real* mat, const real* X, real* out) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= N) return;
  const real* T = mat + (size_t)b * 16;
  const real x0 = X[((size_t)b * 3) * N + i], x1 = X[((size_t)b * 3 + 1) * N + i], x2 = X[((size_t)b * 3 + 2) * N + i];
#pragma unroll
  for (int k = 0; k < 3; ++k) out[((size_t)b * 3 + k) * N + i] = T[k * 4] * x0 + T[k * 4 + 1] * x1 + T[k * 4 + 2] * x2 + T[k * 4 + 3];
}

template <typename real>
__global__ void __launch_bounds__(kBT) transform_bwd_kernel(size_t N, const real* mat, const real* X, const real* gout,
                                                            real* dX, real* part) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  __shared__ real scratch[12 * (kBT / 32)];
  real acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = real(0);
  if (i < N) {
    const real* T = mat + (size_t)b * 16;
    real g[3], x[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { g[k] = gout[((size_t)b * 3 + k) * N + i]; x[k] = X[((size_t)b * 3 + k) * N + i]; }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      dX[((size_t)b * 3 + k) * N + i] = T[k] * g[0] + T[4 + k] * g[1] + T[8 + k] * g[2];
      acc[k * 4] = g[k] * x[0]; acc[k * 4 + 1] = g[k] * x[1]; acc[k * 4 + 2] = g[k] * x[2]; acc[k * 4 + 3] = g[k];
    }
  }
  block_reduce<real, 12>(acc, scratch);
  if (threadIdx.x < 12) part[((size_t)b * gridDim.x + blockIdx.x) * 12 + threadIdx.x] = acc[threadIdx.x];
}

// ------------------------------------------------------------------------------------------
// SSIM (3x3, reflection padding) and the photometric map.
//   mode 0: SSIM index            [B,3,H,W]   (free function SSIM, losses/...:15-54)
//   mode 1: clamp((1-SSIM)/2,0,1) [B,3,H,W]   (method SSIM, :199-216)
//   mode 2: ssim_w * mean_c(mode 1) + (1 - ssim_w) * mean_c |x - y|   [B,1,H,W]   (:236-245)
//   mode 3: |x - y|               [B,3,H,W]   (ssim_loss_weight == 0, :247)
// ------------------------------------------------------------------------------------------
template <typename real> struct Win { real sx, sy, sxx, syy, sxy; };

template <typename real>
__device__ __forceinline__ Win<real> window_sums(const real* x, const real* y, int H, int W, int v, int u) {
  Win<real> w = {real(0), real(0), real(0), real(0), real(0)};
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      const size_t o = (size_t)reflect1(v + dy, H) * W + reflect1(u + dx, W);
      const real a = x[o], b = y[o];
      w.sx += a; w.sy += b; w.sxx += a * a; w.syy += b * b; w.sxy += a * b;
    }
  return w;
}

// value f (SSIM index, or its clamped distance when `dist`) and df / d(sum x, sum x^2, sum xy)
template <typename real>
__device__ __forceinline__ real ssim_value_grad(const Win<real>& w, real C1, real C2, bool dist, real* c) {
  const real r9 = real(1) / real(9);
  const real mx = w.sx * r9, my = w.sy * r9, exx = w.sxx * r9, eyy = w.syy * r9, exy = w.sxy * r9;
  const real sx = exx - mx * mx, sy = eyy - my * my, sxy = exy - mx * my;
  const real a1 = real(2) * mx * my + C1, v1 = real(2) * sxy + C2;
  const real b1 = mx * mx + my * my + C1, v2 = sx + sy + C2;
  const real inv = real(1) / (b1 * v2);
  const real ssim = (a1 * v1) * inv;
  real k = r9;
  real val = ssim;
  if (dist) {
    const real d = (real(1) - ssim) / real(2);
    k = (d >= real(0) && d <= real(1)) ? real(-0.5) * r9 : real(0);
    val = rmin(rmax(d, real(0)), real(1));
  }
  c[0] = k * (real(2) * my * (v1 - a1) * inv - ssim * real(2) * mx * (v2 - b1) * inv);
  c[1] = k * (-ssim * b1 * inv);
  c[2] = k * (real(2) * a1 * inv);
  return val;
}

template <typename real>
__global__ void __launch_bounds__(kBT) photo_map_kernel(int H, int W, int mode, real ssim_w, real C1, real C2,
                                                        const real* x, const real* y, real* out) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  const int v = (int)(i / W), u = (int)(i % W);
  real ds = real(0), ls = real(0);
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    const real* xc = x + ((size_t)b * 3 + ch) * plane;
    const real* yc = y + ((size_t)b * 3 + ch) * plane;
    const real l1 = rabs(xc[i] - yc[i]);
    real val = real(0), c[3];
    if (mode != 3) val = ssim_value_grad<real>(window_sums<real>(xc, yc, H, W, v, u), C1, C2, mode != 0, c);
    if (mode == 3) out[((size_t)b * 3 + ch) * plane + i] = l1;
    else if (mode != 2) out[((size_t)b * 3 + ch) * plane + i] = val;
    ds += val;
    ls += l1;
  }
  if (mode == 2) out[b * plane + i] = ssim_w * (ds / real(3)) + (real(1) - ssim_w) * (ls / real(3));
}

// number of window offsets d in {-1,0,1} with reflect(p + d) == q
__device__ __forceinline__ int reflect_mult(int p, int q, int n) {
  int m = 0;
#pragma unroll
  for (int d = -1; d <= 1; ++d) m += (reflect1(p + d, n) == q) ? 1 : 0;
  return m;
}

// d loss / d x (and / d y) of the maps above given gout (3 channels for modes 0, 1, 3; 1 channel for mode 2)
template <typename real>
__global__ void __launch_bounds__(kBT) photo_map_bwd_kernel(int H, int W, int mode, real ssim_w, real C1, real C2,
                                                            const real* x, const real* y, const real* gout, real* dx,
                                                            real* dy) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  const int v = (int)(i / W), u = (int)(i % W);
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    const real* xc = x + ((size_t)b * 3 + ch) * plane;
    const real* yc = y + ((size_t)b * 3 + ch) * plane;
    const real* g = (mode == 2) ? gout + b * plane : gout + ((size_t)b * 3 + ch) * plane;
    const real xq = xc[i], yq = yc[i];
    real ax = real(0), ay = real(0);
    if (mode != 3) {
      const real scale = (mode == 2) ? ssim_w / real(3) : real(1);
      for (int pv = v - 1; pv <= v + 1; ++pv) {
        if (pv < 0 || pv >= H) continue;
        const int mv = reflect_mult(pv, v, H);
        for (int pu = u - 1; pu <= u + 1; ++pu) {
          if (pu < 0 || pu >= W) continue;
          const int m = mv * reflect_mult(pu, u, W);
          if (m == 0) continue;
          const real w = g[(size_t)pv * W + pu] * scale * (real)m;
          const Win<real> win = window_sums<real>(xc, yc, H, W, pv, pu);
          real c[3];
          ssim_value_grad<real>(win, C1, C2, mode != 0, c);
          ax += w * (c[0] + real(2) * xq * c[1] + yq * c[2]);
          if (dy) {
            const Win<real> sw = {win.sy, win.sx, win.syy, win.sxx, win.sxy};
            ssim_value_grad<real>(sw, C1, C2, mode != 0, c);
            ay += w * (c[0] + real(2) * yq * c[1] + xq * c[2]);
          }
        }
      }
    }
    if (mode >= 2) {
      const real l = (mode == 2) ? g[i] * (real(1) - ssim_w) / real(3) : g[i];
      const real s = rsign(xq - yq);
      ax += l * s;
      ay -= l * s;
    }
    dx[((size_t)b * 3 + ch) * plane + i] = ax;
    if (dy) dy[((size_t)b * 3 + ch) * plane + i] = ay;
  }
}

// ------------------------------------------------------------------------------------------
// deterministic per-batch-item sums: part[b][blk] = sum over the block's elements of x (* y),
// then a fixed-order finish.  op 0: x, op 1: x * y, op 2: |x|.
// ------------------------------------------------------------------------------------------
template <typename real>
__global__ void __launch_bounds__(kBT) rowsum_kernel(size_t N, int op, const real* x, const real* y, double* part) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  __shared__ real scratch[kBT / 32];
  real v[1] = {real(0)};
  if (i < N) {
    const real a = x[b * N + i];
    v[0] = op == 0 ? a : (op == 1 ? a * y[b * N + i] : rabs(a));
  }
  block_reduce<real, 1>(v, scratch);
  if (threadIdx.x == 0) part[(size_t)b * gridDim.x + blockIdx.x] = (double)v[0];
}

// out[b] = scale * sum_blk part[b][blk]   (one block per b; when total != 0 a single block sums all b)
template <typename real>
__global__ void __launch_bounds__(256) rowsum_finish_kernel(const double* part, int nblk, double scale, real* out) {
  __shared__ double sh[256];
  const int b = blockIdx.x;
  double a = 0.0;
  for (int i = threadIdx.x; i < nblk; i += 256) a += part[(size_t)b * nblk + i];
  sh[threadIdx.x] = a;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[b] = (real)(sh[0] * scale);
}

// inv_depths_normalize (utils/depth.py:146-162): out = d / max(mean_b, 1e-6)
template <typename real>
__global__ void __launch_bounds__(kBT) normalize_kernel(size_t N, const real* d, const real* mean, real* out) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i < N) out[b * N + i] = d[b * N + i] / rmax(mean[b], real(1e-6));
}

// d d = g / m - [mean >= 1e-6] * dot_b / (N m^2),  dot_b = sum_i g_i d_i
template <typename real>
__global__ void __launch_bounds__(kBT) normalize_bwd_kernel(size_t N, const real* g, const real* mean, const real* dot,
                                                            real* dd) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= N) return;
  const real m = rmax(mean[b], real(1e-6));
  const real c = (mean[b] >= real(1e-6)) ? dot[b] / ((real)N * m * m) : real(0);
  dd[b * N + i] = g[b * N + i] / m - c;
}

// calc_smoothness maps (utils/depth.py:165-198): sx [B,1,H,W-1], sy [B,1,H-1,W] from the normalised
// inverse depth dn [B,1,H,W] and the image [B,3,H,W]
template <typename real>
__device__ __forceinline__ real edge_weight(const real* img, size_t plane, size_t o, size_t o2) {
  real ig = real(0);
#pragma unroll
  for (int c = 0; c < 3; ++c) ig += rabs(img[c * plane + o] - img[c * plane + o2]);
  return rexp(-(ig / real(3)));
}

template <typename real>
__global__ void __launch_bounds__(kBT) smooth_maps_kernel(int H, int W, const real* dn, const real* img, real* sx,
                                                          real* sy) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  const int v = (int)(i / W), u = (int)(i % W);
  const real* d = dn + b * plane;
  const real* im = img + (size_t)b * 3 * plane;
  if (u + 1 < W) sx[(size_t)b * H * (W - 1) + (size_t)v * (W - 1) + u] = (d[i] - d[i + 1]) * edge_weight<real>(im, plane, i, i + 1);
  if (v + 1 < H) sy[(size_t)b * (H - 1) * W + i] = (d[i] - d[i + W]) * edge_weight<real>(im, plane, i, i + W);
}

template <typename real>
__global__ void __launch_bounds__(kBT) smooth_maps_bwd_kernel(int H, int W, const real* img, const real* gsx,
                                                              const real* gsy, real* ddn) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W, i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  const int v = (int)(i / W), u = (int)(i % W);
  const real* im = img + (size_t)b * 3 * plane;
  const real* gx = gsx + (size_t)b * H * (W - 1) + (size_t)v * (W - 1);
  const real* gy = gsy + (size_t)b * (H - 1) * W;
  real a = real(0);
  if (u + 1 < W) a += gx[u] * edge_weight<real>(im, plane, i, i + 1);
  if (u >= 1) a -= gx[u - 1] * edge_weight<real>(im, plane, i - 1, i);
  if (v + 1 < H) a += gy[i] * edge_weight<real>(im, plane, i, i + W);
  if (v >= 1) a -= gy[i - W] * edge_weight<real>(im, plane, i - W, i);
  ddn[b * plane + i] = a;
}

// ------------------------------------------------------------------------------------------
// reduce_photometric_loss for one scale (losses/...:269-297): candidates k = 0..K-1, each [B,Cg,H,W].
//   op min : mean over (b, pixel) of the min over all K * Cg channels (first index wins ties)
//   op mean: sum_k mean(cand_k) / K
// ------------------------------------------------------------------------------------------
constexpr int kMaxCand = 16;
struct CandPtrs { const void* p[kMaxCand]; void* g[kMaxCand]; };

template <typename real>
__global__ void __launch_bounds__(kBT) cand_reduce_kernel(int K, int Cg, size_t plane, int op, CandPtrs cp,
                                                          int32_t* argmin, double* part) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  __shared__ real scratch[kBT / 32];
  real v[1] = {real(0)};
  if (i < plane) {
    if (op == PSFM_REDUCE_MIN) {
      real best = real(0);
      int sel = -1;
      for (int k = 0; k < K; ++k)
        for (int c = 0; c < Cg; ++c) {
          const real a = ((const real*)cp.p[k])[((size_t)b * Cg + c) * plane + i];
          if (sel < 0 || a < best) { best = a; sel = k * Cg + c; }
        }
      argmin[b * plane + i] = sel;
      v[0] = best;
    } else {
      for (int k = 0; k < K; ++k)
        for (int c = 0; c < Cg; ++c) v[0] += ((const real*)cp.p[k])[((size_t)b * Cg + c) * plane + i];
    }
  }
  block_reduce<real, 1>(v, scratch);
  if (threadIdx.x == 0) part[(size_t)b * gridDim.x + blockIdx.x] = (double)v[0];
}

template <typename real>
__global__ void __launch_bounds__(kBT) cand_reduce_bwd_kernel(int K, int Cg, size_t plane, int op, real scale,
                                                              const real* gout, const int32_t* argmin, CandPtrs cp) {
  const int b = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i >= plane) return;
  const real g = gout[0] * scale;
  const int sel = (op == PSFM_REDUCE_MIN) ? argmin[b * plane + i] : -1;
  for (int k = 0; k < K; ++k)
    for (int c = 0; c < Cg; ++c)
      ((real*)cp.g[k])[((size_t)b * Cg + c) * plane + i] = (op != PSFM_REDUCE_MIN || sel == k * Cg + c) ? g : real(0);
}

// d mean|x| / dx = sign(x) * g / n
template <typename real>
__global__ void __launch_bounds__(kBT) abs_mean_bwd_kernel(size_t n, real scale, const real* x, const real* gout, real* dx) {
  const size_t i = (size_t)blockIdx.x * kBT + threadIdx.x;
  if (i < n) dx[i] = rsign(x[i]) * gout[0] * scale;
}


// ------------------------------------------------------------------------------------------
// Pose.from_vec / pose_vec2mat / euler2mat (geometry/pose.py:39-46, geometry/pose_utils.py:8-51):
// vec = [tx, ty, tz, rx, ry, rz] -> [R | t; 0 0 0 1] with R = Rx(rx) Ry(ry) Rz(rz).  One thread per pose.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void sincos_full(float a, float& s, float& c) { s = sinf(a); c = cosf(a); }
__device__ __forceinline__ void sincos_full(double a, double& s, double& c) { s = sin(a); c = cos(a); }

template <typename real>
__device__ __forceinline__ void euler_rot(const real* v, real* R, real* sc) {
  real sx, cx, sy, cy, sz, cz;
  sincos_full(v[3], sx, cx); sincos_full(v[4], sy, cy); sincos_full(v[5], sz, cz);
  R[0] = cy * cz;                R[1] = -cy * sz;               R[2] = sy;
  R[3] = sx * sy * cz + cx * sz; R[4] = cx * cz - sx * sy * sz; R[5] = -sx * cy;
  R[6] = sx * sz - cx * sy * cz; R[7] = cx * sy * sz + sx * cz; R[8] = cx * cy;
  sc[0] = sx; sc[1] = cx; sc[2] = sy; sc[3] = cy; sc[4] = sz; sc[5] = cz;
}

template <typename real>
__global__ void pose_vec_kernel(int B, const real* __restrict__ vec, real* __restrict__ mat) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  real R[9], sc[6];
  euler_rot<real>(vec + (size_t)b * 6, R, sc);
  real* m = mat + (size_t)b * 16;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    m[i * 4] = R[i * 3]; m[i * 4 + 1] = R[i * 3 + 1]; m[i * 4 + 2] = R[i * 3 + 2];
    m[i * 4 + 3] = vec[(size_t)b * 6 + i];
  }
  m[12] = real(0); m[13] = real(0); m[14] = real(0); m[15] = real(1);
}

template <typename real>
__global__ void pose_vec_bwd_kernel(int B, const real* __restrict__ vec, const real* __restrict__ g,
                                    real* __restrict__ d_vec) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  real R[9], sc[6], G[9];
  euler_rot<real>(vec + (size_t)b * 6, R, sc);
  const real sx = sc[0], cx = sc[1], sy = sc[2], cy = sc[3], sz = sc[4], cz = sc[5];
  const real* gm = g + (size_t)b * 16;
#pragma unroll
  for (int i = 0; i < 3; ++i) { G[i * 3] = gm[i * 4]; G[i * 3 + 1] = gm[i * 4 + 1]; G[i * 3 + 2] = gm[i * 4 + 2]; }
  real* o = d_vec + (size_t)b * 6;
  o[0] = gm[3]; o[1] = gm[7]; o[2] = gm[11];
  // d R / d rx: row 1 -> -row 2, row 2 -> row 1
  o[3] = (G[6] * R[3] + G[7] * R[4] + G[8] * R[5]) - (G[3] * R[6] + G[4] * R[7] + G[5] * R[8]);
  // d R / d ry
  o[4] = G[0] * (-sy * cz) + G[1] * (sy * sz) + G[2] * cy + G[3] * (sx * cy * cz) + G[4] * (-sx * cy * sz) +
         G[5] * (sx * sy) + G[6] * (-cx * cy * cz) + G[7] * (cx * cy * sz) + G[8] * (-cx * sy);
  // d R / d rz: column 0 -> column 1, column 1 -> -column 0
  o[5] = (G[0] * R[1] + G[3] * R[4] + G[6] * R[7]) - (G[1] * R[0] + G[4] * R[3] + G[7] * R[6]);
}

}  // namespace psfm

// ------------------------------------------------------------------------------------------
// C ABI (declared in include/psfm.h)
// ------------------------------------------------------------------------------------------
int psfm_fail(int code, const char* fmt, ...);      // psfm_abi.cu: sets the thread-local error string

using namespace psfm;

#define PSFM_BLOCK_CHECK(cond, msg) do { if (!(cond)) return psfm_fail(PSFM_ERR_INVALID, msg); } while (0)
#define PSFM_DT_CHECK(dt) PSFM_BLOCK_CHECK((dt) == PSFM_F32 || (dt) == PSFM_F64, "real_dtype must be f32|f64")
// run `call` with real = float or double according to dt
#define PSFM_DISPATCH(dt, ...) do { if ((dt) == PSFM_F64) { typedef double real; __VA_ARGS__; } else { typedef float real; __VA_ARGS__; } } while (0)

static int launch_result(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) return PSFM_OK;
  return psfm_fail(PSFM_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

static inline size_t al256(size_t x) { return (x + 255) / 256 * 256; }

extern "C" {

size_t psfm_block_workspace_bytes(int B, size_t N) {
  if (B < 1 || N < 1) return 0;
  return al256((size_t)B * nblk_for(N) * 12 * sizeof(double)) + al256((size_t)B * sizeof(double)) + 256;
}

int psfm_reconstruct(int B, int H, int W, int dt, const void* depth, const void* K, const void* Twc, void* out,
                     void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 1 && W >= 1 && depth && K && out, "psfm_reconstruct: bad argument");
  const dim3 grid(nblk_for((size_t)H * W), B);
  count_launches(1);
  PSFM_DISPATCH(dt, (reconstruct_kernel<real><<<grid, kBT, 0, (cudaStream_t)stream>>>(
                        H, W, (const real*)depth, (const real*)K, (const real*)Twc, (real*)out)));
  return launch_result("psfm_reconstruct");
}

int psfm_reconstruct_backward(int B, int H, int W, int dt, const void* K, const void* Twc, const void* grad_out,
                              void* d_depth, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 1 && W >= 1 && K && grad_out && d_depth, "psfm_reconstruct_backward: bad argument");
  const dim3 grid(nblk_for((size_t)H * W), B);
  count_launches(1);
  PSFM_DISPATCH(dt, (reconstruct_bwd_kernel<real><<<grid, kBT, 0, (cudaStream_t)stream>>>(
                        H, W, (const real*)K, (const real*)Twc, (const real*)grad_out, (real*)d_depth)));
  return launch_result("psfm_reconstruct_backward");
}

int psfm_project(int B, int H, int W, int dt, const void* X, const void* K, const void* Tcw, void* out, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 2 && W >= 2 && X && K && out, "psfm_project: bad argument");
  const dim3 grid(nblk_for((size_t)H * W), B);
  count_launches(1);
  PSFM_DISPATCH(dt, (project_kernel<real><<<grid, kBT, 0, (cudaStream_t)stream>>>(
                        H, W, (const real*)X, (const real*)K, (const real*)Tcw, (real*)out)));
  return launch_result("psfm_project");
}

int psfm_project_backward(int B, int H, int W, int dt, const void* X, const void* K, const void* Tcw,
                          const void* grad_out, void* dX, void* d_Tcw, void* ws, size_t ws_bytes, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 2 && W >= 2 && X && K && grad_out && dX, "psfm_project_backward: bad argument");
  const bool pose = Tcw && d_Tcw;
  if (pose && (!ws || ws_bytes < psfm_block_workspace_bytes(B, (size_t)H * W)))
    return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_project_backward: workspace too small");
  const unsigned nb = nblk_for((size_t)H * W);
  count_launches(pose ? 2 : 1);
  PSFM_DISPATCH(dt, {
    project_bwd_kernel<real><<<dim3(nb, B), kBT, 0, (cudaStream_t)stream>>>(
        H, W, (const real*)X, (const real*)K, (const real*)Tcw, (const real*)grad_out, (real*)dX, pose ? (real*)ws : nullptr);
    if (pose) mat_finish_kernel<real><<<B, 32, 0, (cudaStream_t)stream>>>((const real*)ws, (int)nb, (real*)d_Tcw);
  });
  return launch_result("psfm_project_backward");
}

int psfm_fisheye_reconstruct(int B, int H, int W, int dt, const void* depth, const void* intr, void* out, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 1 && W >= 1 && depth && intr && out, "psfm_fisheye_reconstruct: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (fisheye_reconstruct_kernel<real><<<dim3(nblk_for((size_t)H * W), B), kBT, 0, (cudaStream_t)stream>>>(
                        H, W, (const real*)depth, (const real*)intr, (real*)out)));
  return launch_result("psfm_fisheye_reconstruct");
}

int psfm_fisheye_reconstruct_backward(int B, int H, int W, int dt, const void* intr, const void* grad_out, void* d_depth,
                                      void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 1 && W >= 1 && intr && grad_out && d_depth, "psfm_fisheye_reconstruct_backward: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (fisheye_reconstruct_bwd_kernel<real><<<dim3(nblk_for((size_t)H * W), B), kBT, 0, (cudaStream_t)stream>>>(
                        H, W, (const real*)intr, (const real*)grad_out, (real*)d_depth)));
  return launch_result("psfm_fisheye_reconstruct_backward");
}

int psfm_fisheye_project(int B, size_t N, int Hn, int Wn, int dt, const void* X, const void* intr, void* out, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && N >= 1 && Hn >= 2 && Wn >= 2 && X && intr && out, "psfm_fisheye_project: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (fisheye_project_kernel<real><<<dim3(nblk_for(N), B), kBT, 0, (cudaStream_t)stream>>>(
                        N, Hn, Wn, (const real*)X, (const real*)intr, (real*)out)));
  return launch_result("psfm_fisheye_project");
}

int psfm_fisheye_project_backward(int B, size_t N, int Hn, int Wn, int dt, const void* X, const void* intr,
                                  const void* grad_out, void* dX, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && N >= 1 && Hn >= 2 && Wn >= 2 && X && intr && grad_out && dX,
                   "psfm_fisheye_project_backward: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (fisheye_project_bwd_kernel<real><<<dim3(nblk_for(N), B), kBT, 0, (cudaStream_t)stream>>>(
                        N, Hn, Wn, (const real*)X, (const real*)intr, (const real*)grad_out, (real*)dX)));
  return launch_result("psfm_fisheye_project_backward");
}

int psfm_grid_sample(int B, int Hi, int Wi, int Ho, int Wo, int dt, int padding, const void* image, const void* grid,
                     void* out, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && Hi >= 1 && Wi >= 1 && Ho >= 1 && Wo >= 1 && image && grid && out &&
                       (padding == PSFM_PAD_ZEROS || padding == PSFM_PAD_BORDER), "psfm_grid_sample: bad argument");
  count_launches(1);
  const size_t No = (size_t)Ho * Wo;
  PSFM_DISPATCH(dt, (grid_sample_kernel<real><<<dim3(nblk_for(No), B), kBT, 0, (cudaStream_t)stream>>>(
                        Hi, Wi, No, padding, (const real*)image, (const real*)grid, (real*)out)));
  return launch_result("psfm_grid_sample");
}

int psfm_grid_sample_backward(int B, int Hi, int Wi, int Ho, int Wo, int dt, int padding, const void* image,
                              const void* grid, const void* grad_out, void* d_grid, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && Hi >= 1 && Wi >= 1 && Ho >= 1 && Wo >= 1 && image && grid && grad_out && d_grid &&
                       (padding == PSFM_PAD_ZEROS || padding == PSFM_PAD_BORDER), "psfm_grid_sample_backward: bad argument");
  count_launches(1);
  const size_t No = (size_t)Ho * Wo;
  PSFM_DISPATCH(dt, (grid_sample_bwd_kernel<real><<<dim3(nblk_for(No), B), kBT, 0, (cudaStream_t)stream>>>(
                        Hi, Wi, No, padding, (const real*)image, (const real*)grid, (const real*)grad_out, (real*)d_grid)));
  return launch_result("psfm_grid_sample_backward");
}

// blocks of the supervised-loss kernels: one 16-byte vector per thread up to 8 CTAs per SM of a 148-SM part, grid-stride beyond
static inline unsigned sup_grid(size_t N) { const unsigned nb = nblk_for((N + 3) / 4); return nb < 148u * 8u ? nb : 148u * 8u; }

int psfm_supervised_loss(int method, int sparse, size_t N, int dt, const void* pred, const void* gt, const void* mask,
                         const double* hp, void* loss_out, void* state, void* ws, size_t ws_bytes, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(N >= 1 && pred && gt && hp && loss_out && state && method >= PSFM_SUP_L1 && method <= PSFM_SUP_SSI_SILOG,
                   "psfm_supervised_loss: bad argument");
  if (!ws || ws_bytes < psfm_block_workspace_bytes(1, N)) return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_supervised_loss: workspace too small");
  if ((((uintptr_t)pred) | ((uintptr_t)gt)) & 15) return psfm_fail(PSFM_ERR_INVALID, "psfm_supervised_loss: pred / gt must be 16-byte aligned");
  SupHp h = {hp[0], hp[1], hp[2], hp[3], hp[4], hp[5]};
  if (h.cmax <= h.cmin) h.cmax = h.cmin + 1.0;                 // losses/ssi_silog_loss.py:227-228
  const unsigned nb = sup_grid(N);
  cudaStream_t st = (cudaStream_t)stream;
  count_launches(2);
  PSFM_DISPATCH(dt, {
    sup_sums_kernel<real><<<nb, kBT, 0, st>>>(N, method, sparse, (const real*)pred, (const real*)gt, (const uint8_t*)mask, h, (double*)ws);
    sup_finish_kernel<real><<<1, 256, 0, st>>>((const double*)ws, (int)nb, N, method, h, (double*)state, (real*)loss_out);
  });
  return launch_result("psfm_supervised_loss");
}

int psfm_supervised_loss_backward(int method, int sparse, size_t N, int dt, const void* pred, const void* gt, const void* mask,
                                  const double* hp, const void* state, const void* grad_loss, void* d_pred, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(N >= 1 && pred && gt && hp && state && grad_loss && d_pred && method >= PSFM_SUP_L1 && method <= PSFM_SUP_SSI_SILOG,
                   "psfm_supervised_loss_backward: bad argument");
  if ((((uintptr_t)pred) | ((uintptr_t)gt)) & 15) return psfm_fail(PSFM_ERR_INVALID, "psfm_supervised_loss: pred / gt must be 16-byte aligned");
  SupHp h = {hp[0], hp[1], hp[2], hp[3], hp[4], hp[5]};
  if (h.cmax <= h.cmin) h.cmax = h.cmin + 1.0;
  if (((uintptr_t)d_pred) & 15) return psfm_fail(PSFM_ERR_INVALID, "psfm_supervised_loss_backward: d_pred must be 16-byte aligned");
  count_launches(1);
  PSFM_DISPATCH(dt, (sup_grad_kernel<real><<<sup_grid(N), kBT, 0, (cudaStream_t)stream>>>(
                        N, method, sparse, (const real*)pred, (const real*)gt, (const uint8_t*)mask, h,
                        (const double*)state, (const real*)grad_loss, (real*)d_pred)));
  return launch_result("psfm_supervised_loss_backward");
}

int psfm_dual_head_loss(size_t N, int dt, const void* integer_pred, const void* fractional_pred, const void* depth_gt,
                        const double* hp, void* loss_out, void* state, void* ws, size_t ws_bytes, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(N >= 1 && integer_pred && fractional_pred && depth_gt && hp && loss_out && state, "psfm_dual_head_loss: bad argument");
  if (!(hp[0] > hp[1]) || !(hp[0] > 0.0) || hp[1] < 0.0)
    return psfm_fail(PSFM_ERR_INVALID, "psfm_dual_head_loss: need max_depth > min_depth >= 0 (got %g, %g)", hp[0], hp[1]);
  if (!ws || ws_bytes < psfm_block_workspace_bytes(1, N)) return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_dual_head_loss: workspace too small");
  const DualHp h = {hp[0], hp[1], hp[2], hp[3], hp[4]};
  const unsigned nb = nblk_for(N) < 148u * 8u ? nblk_for(N) : 148u * 8u;
  cudaStream_t st = (cudaStream_t)stream;
  count_launches(2);
  PSFM_DISPATCH(dt, {
    dual_sums_kernel<real><<<nb, kBT, 0, st>>>(N, (const real*)integer_pred, (const real*)fractional_pred, (const real*)depth_gt, h, (double*)ws);
    dual_finish_kernel<real><<<1, 256, 0, st>>>((const double*)ws, (int)nb, h, (double*)state, (real*)loss_out);
  });
  return launch_result("psfm_dual_head_loss");
}

int psfm_dual_head_loss_backward(size_t N, int dt, const void* integer_pred, const void* fractional_pred, const void* depth_gt,
                                 const double* hp, const void* state, const void* grad_loss, void* d_integer, void* d_fractional,
                                 void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(N >= 1 && integer_pred && fractional_pred && depth_gt && hp && state && grad_loss && d_integer && d_fractional,
                   "psfm_dual_head_loss_backward: bad argument");
  const DualHp h = {hp[0], hp[1], hp[2], hp[3], hp[4]};
  const unsigned nb = nblk_for(N) < 148u * 8u ? nblk_for(N) : 148u * 8u;
  count_launches(1);
  PSFM_DISPATCH(dt, (dual_grad_kernel<real><<<nb, kBT, 0, (cudaStream_t)stream>>>(
                        N, (const real*)integer_pred, (const real*)fractional_pred, (const real*)depth_gt, h,
                        (const double*)state, (const real*)grad_loss, (real*)d_integer, (real*)d_fractional)));
  return launch_result("psfm_dual_head_loss_backward");
}

int psfm_pose_from_vec(int B, int dt, const void* vec, void* mat, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && vec && mat, "psfm_pose_from_vec: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (pose_vec_kernel<real><<<(B + 63) / 64, 64, 0, (cudaStream_t)stream>>>(B, (const real*)vec, (real*)mat)));
  return launch_result("psfm_pose_from_vec");
}

int psfm_pose_from_vec_backward(int B, int dt, const void* vec, const void* grad_mat, void* d_vec, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && vec && grad_mat && d_vec, "psfm_pose_from_vec_backward: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (pose_vec_bwd_kernel<real><<<(B + 63) / 64, 64, 0, (cudaStream_t)stream>>>(
                        B, (const real*)vec, (const real*)grad_mat, (real*)d_vec)));
  return launch_result("psfm_pose_from_vec_backward");
}

int psfm_transform_points(int B, size_t N, int dt, const void* mat, const void* X, void* out, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && N >= 1 && mat && X && out, "psfm_transform_points: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (transform_kernel<real><<<dim3(nblk_for(N), B), kBT, 0, (cudaStream_t)stream>>>(
                        N, (const real*)mat, (const real*)X, (real*)out)));
  return launch_result("psfm_transform_points");
}

int psfm_transform_points_backward(int B, size_t N, int dt, const void* mat, const void* X, const void* grad_out,
                                   void* dX, void* d_mat, void* ws, size_t ws_bytes, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && N >= 1 && mat && X && grad_out && dX && d_mat, "psfm_transform_points_backward: bad argument");
  if (!ws || ws_bytes < psfm_block_workspace_bytes(B, N))
    return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_transform_points_backward: workspace too small");
  const unsigned nb = nblk_for(N);
  count_launches(2);
  PSFM_DISPATCH(dt, {
    transform_bwd_kernel<real><<<dim3(nb, B), kBT, 0, (cudaStream_t)stream>>>(
        N, (const real*)mat, (const real*)X, (const real*)grad_out, (real*)dX, (real*)ws);
    mat_finish_kernel<real><<<B, 32, 0, (cudaStream_t)stream>>>((const real*)ws, (int)nb, (real*)d_mat);
  });
  return launch_result("psfm_transform_points_backward");
}

int psfm_photometric_map(int B, int H, int W, int dt, int mode, double ssim_w, double C1, double C2, const void* x,
                         const void* y, void* out, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 2 && W >= 2 && x && y && out && mode >= 0 && mode <= 3, "psfm_photometric_map: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (photo_map_kernel<real><<<dim3(nblk_for((size_t)H * W), B), kBT, 0, (cudaStream_t)stream>>>(
                        H, W, mode, (real)ssim_w, (real)C1, (real)C2, (const real*)x, (const real*)y, (real*)out)));
  return launch_result("psfm_photometric_map");
}

int psfm_photometric_map_backward(int B, int H, int W, int dt, int mode, double ssim_w, double C1, double C2,
                                  const void* x, const void* y, const void* grad_out, void* dx, void* dy, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 2 && W >= 2 && x && y && grad_out && dx && mode >= 0 && mode <= 3,
                   "psfm_photometric_map_backward: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (photo_map_bwd_kernel<real><<<dim3(nblk_for((size_t)H * W), B), kBT, 0, (cudaStream_t)stream>>>(
                        H, W, mode, (real)ssim_w, (real)C1, (real)C2, (const real*)x, (const real*)y,
                        (const real*)grad_out, (real*)dx, (real*)dy)));
  return launch_result("psfm_photometric_map_backward");
}

int psfm_normalize(int B, size_t N, int dt, const void* d, void* out, void* mean_out, void* ws, size_t ws_bytes,
                   void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && N >= 1 && d && out && mean_out, "psfm_normalize: bad argument");
  if (!ws || ws_bytes < psfm_block_workspace_bytes(B, N)) return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_normalize: workspace too small");
  const unsigned nb = nblk_for(N);
  cudaStream_t st = (cudaStream_t)stream;
  count_launches(3);
  PSFM_DISPATCH(dt, {
    rowsum_kernel<real><<<dim3(nb, B), kBT, 0, st>>>(N, 0, (const real*)d, nullptr, (double*)ws);
    rowsum_finish_kernel<real><<<B, 256, 0, st>>>((const double*)ws, (int)nb, 1.0 / (double)N, (real*)mean_out);
    normalize_kernel<real><<<dim3(nb, B), kBT, 0, st>>>(N, (const real*)d, (const real*)mean_out, (real*)out);
  });
  return launch_result("psfm_normalize");
}

int psfm_normalize_backward(int B, size_t N, int dt, const void* d, const void* mean, const void* grad_out, void* dd,
                            void* ws, size_t ws_bytes, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && N >= 1 && d && mean && grad_out && dd, "psfm_normalize_backward: bad argument");
  if (!ws || ws_bytes < psfm_block_workspace_bytes(B, N))
    return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_normalize_backward: workspace too small");
  const unsigned nb = nblk_for(N);
  cudaStream_t st = (cudaStream_t)stream;
  char* dot = (char*)ws + al256((size_t)B * nb * 12 * sizeof(double));
  count_launches(3);
  PSFM_DISPATCH(dt, {
    rowsum_kernel<real><<<dim3(nb, B), kBT, 0, st>>>(N, 1, (const real*)grad_out, (const real*)d, (double*)ws);
    rowsum_finish_kernel<real><<<B, 256, 0, st>>>((const double*)ws, (int)nb, 1.0, (real*)dot);
    normalize_bwd_kernel<real><<<dim3(nb, B), kBT, 0, st>>>(N, (const real*)grad_out, (const real*)mean, (const real*)dot, (real*)dd);
  });
  return launch_result("psfm_normalize_backward");
}

int psfm_smoothness_maps(int B, int H, int W, int dt, const void* dn, const void* image, void* sx, void* sy, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 2 && W >= 2 && dn && image && sx && sy, "psfm_smoothness_maps: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (smooth_maps_kernel<real><<<dim3(nblk_for((size_t)H * W), B), kBT, 0, (cudaStream_t)stream>>>(
                        H, W, (const real*)dn, (const real*)image, (real*)sx, (real*)sy)));
  return launch_result("psfm_smoothness_maps");
}

int psfm_smoothness_maps_backward(int B, int H, int W, int dt, const void* image, const void* gsx, const void* gsy,
                                  void* ddn, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(B >= 1 && H >= 2 && W >= 2 && image && gsx && gsy && ddn, "psfm_smoothness_maps_backward: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (smooth_maps_bwd_kernel<real><<<dim3(nblk_for((size_t)H * W), B), kBT, 0, (cudaStream_t)stream>>>(
                        H, W, (const real*)image, (const real*)gsx, (const real*)gsy, (real*)ddn)));
  return launch_result("psfm_smoothness_maps_backward");
}

int psfm_reduce_candidates(int K, int B, int Cg, int H, int W, int dt, int op, const void* const* cands, void* out,
                           int32_t* argmin, void* ws, size_t ws_bytes, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(K >= 1 && K <= kMaxCand && B >= 1 && Cg >= 1 && H >= 1 && W >= 1 && cands && out,
                   "psfm_reduce_candidates: bad argument (1..16 candidates)");
  if (op != PSFM_REDUCE_MIN && op != PSFM_REDUCE_MEAN) return psfm_fail(PSFM_ERR_INVALID, "Unknown photometric_reduce_op: %d", op);
  PSFM_BLOCK_CHECK(op != PSFM_REDUCE_MIN || argmin, "psfm_reduce_candidates: argmin is NULL");
  const size_t plane = (size_t)H * W;
  if (!ws || ws_bytes < psfm_block_workspace_bytes(B, plane)) return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_reduce_candidates: workspace too small");
  CandPtrs cp;
  for (int k = 0; k < kMaxCand; ++k) { cp.p[k] = k < K ? cands[k] : nullptr; cp.g[k] = nullptr; }
  const unsigned nb = nblk_for(plane);
  const double scale = op == PSFM_REDUCE_MIN ? 1.0 / ((double)B * plane) : 1.0 / ((double)K * B * Cg * plane);
  cudaStream_t st = (cudaStream_t)stream;
  count_launches(2);
  PSFM_DISPATCH(dt, {
    cand_reduce_kernel<real><<<dim3(nb, B), kBT, 0, st>>>(K, Cg, plane, op, cp, argmin, (double*)ws);
    rowsum_finish_kernel<real><<<1, 256, 0, st>>>((const double*)ws, (int)(nb * B), scale, (real*)out);
  });
  return launch_result("psfm_reduce_candidates");
}

int psfm_reduce_candidates_backward(int K, int B, int Cg, int H, int W, int dt, int op, const int32_t* argmin,
                                    const void* grad_out, void* const* d_cands, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(K >= 1 && K <= kMaxCand && B >= 1 && Cg >= 1 && grad_out && d_cands, "psfm_reduce_candidates_backward: bad argument");
  PSFM_BLOCK_CHECK(op != PSFM_REDUCE_MIN || argmin, "psfm_reduce_candidates_backward: argmin is NULL");
  const size_t plane = (size_t)H * W;
  CandPtrs cp;
  for (int k = 0; k < kMaxCand; ++k) { cp.p[k] = nullptr; cp.g[k] = k < K ? d_cands[k] : nullptr; }
  const double scale = op == PSFM_REDUCE_MIN ? 1.0 / ((double)B * plane) : 1.0 / ((double)K * B * Cg * plane);
  count_launches(1);
  PSFM_DISPATCH(dt, (cand_reduce_bwd_kernel<real><<<dim3(nblk_for(plane), B), kBT, 0, (cudaStream_t)stream>>>(
                        K, Cg, plane, op, (real)scale, (const real*)grad_out, argmin, cp)));
  return launch_result("psfm_reduce_candidates_backward");
}

int psfm_abs_mean(size_t n, int dt, const void* x, void* out, void* ws, size_t ws_bytes, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(n >= 1 && x && out, "psfm_abs_mean: bad argument");
  if (!ws || ws_bytes < psfm_block_workspace_bytes(1, n)) return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_abs_mean: workspace too small");
  const unsigned nb = nblk_for(n);
  cudaStream_t st = (cudaStream_t)stream;
  count_launches(2);
  PSFM_DISPATCH(dt, {
    rowsum_kernel<real><<<dim3(nb, 1), kBT, 0, st>>>(n, 2, (const real*)x, nullptr, (double*)ws);
    rowsum_finish_kernel<real><<<1, 256, 0, st>>>((const double*)ws, (int)nb, 1.0 / (double)n, (real*)out);
  });
  return launch_result("psfm_abs_mean");
}

int psfm_abs_mean_backward(size_t n, int dt, const void* x, const void* grad_out, void* dx, void* stream) {
  PSFM_DT_CHECK(dt);
  PSFM_BLOCK_CHECK(n >= 1 && x && grad_out && dx, "psfm_abs_mean_backward: bad argument");
  count_launches(1);
  PSFM_DISPATCH(dt, (abs_mean_bwd_kernel<real><<<nblk_for(n), kBT, 0, (cudaStream_t)stream>>>(
                        n, (real)(1.0 / (double)n), (const real*)x, (const real*)grad_out, (real*)dx)));
  return launch_result("psfm_abs_mean_backward");
}

}  // extern "C"
