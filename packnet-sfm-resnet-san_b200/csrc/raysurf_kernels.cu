// raysurf_kernels.cu -- projection of the generic (ray-surface) camera model, forward and backward.
//
// Reference: GenericCamera.project (geometry/camera_generic.py:86-208).  A 3-D point is "projected"
// by a soft arg-max: the cosine between the point's direction and the camera's viewing rays in a 41 x 41 patch around
// the pixel is turned into weights by a softmax with an annealed temperature, and the projection is the weighted mean
// of the patch's pixel coordinates.  The reference materialises the [h w, 1681, 3] patch tensor and runs bmm / softmax /
// bmm on it (it is batch-1 and calls .cuda() on the way); here a warp owns a pixel, walks the patch in the ray image
// itself (L1 / L2 resident: neighbouring pixels share 40/41 of their patch) and keeps its 53 logits in registers.
//   forward  : dir [B,3,h,w] (un-normalised point directions), rays [B,3,h,w] -> coords [B,h,w,2] = (row, col),
//              dhat [B,3,h,w] (normalised directions), stats [B,h,w,2] = (max logit, sum of exp)
//   backward : grad_coords -> d_dir (through the softmax and the normalisation) and d_rays.  d_rays is GATHERED: a warp
//              owns a ray and visits the pixels whose patch holds it, in a fixed order -- no atomics, deterministic.
#include <cuda_runtime.h>
#include "sweep_device.cuh"
#include "loss_tile.cuh"
#include "profile.h"

extern int psfm_fail(int code, const char* fmt, ...);

namespace psfm {

constexpr int kPatch = 20;                       // camera_generic.py:104
constexpr int kSide = 2 * kPatch + 1;            // 41
constexpr int kRays = kSide * kSide;             // 1681
constexpr int kPerLane = (kRays + 31) / 32;      // 53

// first index of the patch around i, shifted to stay inside [0, n) (camera_generic.py:133-147)
__device__ __forceinline__ int patch_origin(int i, int n) { return min(max(i - kPatch, 0), n - kSide); }

template <typename real> __device__ __forceinline__ real warp_allmax(real v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = rmax(v, __shfl_xor_sync(kFull, v, o));
  return v;
}

template <typename real>
__global__ void __launch_bounds__(256) raysurf_fwd_kernel(int B, int h, int w, real inv_t, const real* __restrict__ dir,
                                                          const real* __restrict__ rays, real* __restrict__ coords,
                                                          real* __restrict__ dhat, real* __restrict__ stats) {
  const int lane = threadIdx.x & 31;
  const size_t px = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5), plane = (size_t)h * w;
  if (px >= (size_t)B * plane) return;
  const int b = (int)(px / plane), r = (int)(px - (size_t)b * plane), i = r / w, j = r - i * w;
  const real* dp = dir + (size_t)b * 3 * plane + r;
  const real d0 = dp[0], d1 = dp[plane], d2 = dp[2 * plane];
  const real inv = real(1) / rroot(d0 * d0 + d1 * d1 + d2 * d2);
  const real n0 = d0 * inv, n1 = d1 * inv, n2 = d2 * inv;
  const int pi = patch_origin(i, h), pj = patch_origin(j, w);
  const real* rp = rays + (size_t)b * 3 * plane;
  real l[kPerLane];
  real m = -real_inf<real>();
#pragma unroll
  for (int t = 0; t < kPerLane; ++t) {
    const int k = t * 32 + lane;
    l[t] = -real_inf<real>();
    if (k < kRays) {
      const int a = k / kSide, c = k - a * kSide;
      const size_t o = (size_t)(pi + a) * w + (pj + c);
      l[t] = n0 * rp[o] + n1 * rp[plane + o] + n2 * rp[2 * plane + o];
      m = rmax(m, l[t]);
    }
  }
  m = warp_allmax<real>(m);
  real Z = real(0), Sr = real(0), Sc = real(0);
#pragma unroll
  for (int t = 0; t < kPerLane; ++t) {
    const int k = t * 32 + lane;
    if (k < kRays) {
      const int a = k / kSide, c = k - a * kSide;
      const real e = rexp((l[t] - m) * inv_t);
      Z += e; Sr += e * (real)(pi + a); Sc += e * (real)(pj + c);
    }
  }
  Z = warp_allsum<real>(Z); Sr = warp_allsum<real>(Sr); Sc = warp_allsum<real>(Sc);
  if (lane == 0) {
    coords[px * 2] = Sr / Z; coords[px * 2 + 1] = Sc / Z;
    stats[px * 2] = m; stats[px * 2 + 1] = Z;
    real* hp = dhat + (size_t)b * 3 * plane + r;
    hp[0] = n0; hp[plane] = n1; hp[2 * plane] = n2;
  }
}

// d loss / d dir: per pixel, c_k = p_k / T ((row_k - row) g_r + (col_k - col) g_c), d dhat = sum_k c_k ray_k, then the
// adjoint of the normalisation dhat = dir / |dir|
template <typename real>
__global__ void __launch_bounds__(256) raysurf_bwd_dir_kernel(int B, int h, int w, real inv_t, const real* __restrict__ dir,
                                                              const real* __restrict__ rays, const real* __restrict__ coords,
                                                              const real* __restrict__ dhat, const real* __restrict__ stats,
                                                              const real* __restrict__ gcoords, real* __restrict__ d_dir) {
  const int lane = threadIdx.x & 31;
  const size_t px = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5), plane = (size_t)h * w;
  if (px >= (size_t)B * plane) return;
  const int b = (int)(px / plane), r = (int)(px - (size_t)b * plane), i = r / w, j = r - i * w;
  const real* hp = dhat + (size_t)b * 3 * plane + r;
  const real n0 = hp[0], n1 = hp[plane], n2 = hp[2 * plane];
  const real row = coords[px * 2], col = coords[px * 2 + 1], m = stats[px * 2], iz = real(1) / stats[px * 2 + 1];
  const real gr = gcoords[px * 2] * inv_t, gc = gcoords[px * 2 + 1] * inv_t;
  const int pi = patch_origin(i, h), pj = patch_origin(j, w);
  const real* rp = rays + (size_t)b * 3 * plane;
  real g0 = real(0), g1 = real(0), g2 = real(0);
  for (int k = lane; k < kRays; k += 32) {
    const int a = k / kSide, c = k - a * kSide;
    const size_t o = (size_t)(pi + a) * w + (pj + c);
    const real r0 = rp[o], r1 = rp[plane + o], r2 = rp[2 * plane + o];
    const real p = rexp(((n0 * r0 + n1 * r1 + n2 * r2) - m) * inv_t) * iz;
    const real ck = p * (((real)(pi + a) - row) * gr + ((real)(pj + c) - col) * gc);
    g0 += ck * r0; g1 += ck * r1; g2 += ck * r2;
  }
  g0 = warp_allsum<real>(g0); g1 = warp_allsum<real>(g1); g2 = warp_allsum<real>(g2);
  if (lane == 0) {
    const real* dp = dir + (size_t)b * 3 * plane + r;
    const real inv = real(1) / rroot(dp[0] * dp[0] + dp[plane] * dp[plane] + dp[2 * plane] * dp[2 * plane]);
    const real dot = g0 * n0 + g1 * n1 + g2 * n2;
    real* o = d_dir + (size_t)b * 3 * plane + r;
    o[0] = (g0 - dot * n0) * inv; o[plane] = (g1 - dot * n1) * inv; o[2 * plane] = (g2 - dot * n2) * inv;
  }
}

// d loss / d rays: per ray (y, x), the pixels (i, j) whose patch holds it: rows i with patch_origin(i) <= y <= patch_origin(i) + 40
template <typename real>
__global__ void __launch_bounds__(256) raysurf_bwd_rays_kernel(int B, int h, int w, real inv_t, const real* __restrict__ rays,
                                                               const real* __restrict__ coords, const real* __restrict__ dhat,
                                                               const real* __restrict__ stats, const real* __restrict__ gcoords,
                                                               real* __restrict__ d_rays) {
  const int lane = threadIdx.x & 31;
  const size_t px = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5), plane = (size_t)h * w;
  if (px >= (size_t)B * plane) return;
  const int b = (int)(px / plane), r = (int)(px - (size_t)b * plane), y = r / w, x = r - y * w;
  const real* rp = rays + (size_t)b * 3 * plane + r;
  const real r0 = rp[0], r1 = rp[plane], r2 = rp[2 * plane];
  const int i0 = y <= 2 * kPatch ? 0 : y - kPatch, i1 = y >= h - kSide ? h - 1 : y + kPatch;
  const int j0 = x <= 2 * kPatch ? 0 : x - kPatch, j1 = x >= w - kSide ? w - 1 : x + kPatch;
  const int nj = j1 - j0 + 1, cnt = (i1 - i0 + 1) * nj;
  const real* hp = dhat + (size_t)b * 3 * plane;
  const size_t pb = (size_t)b * plane;
  real g0 = real(0), g1 = real(0), g2 = real(0);
  for (int k = lane; k < cnt; k += 32) {
    const int a = k / nj, i = i0 + a, j = j0 + (k - a * nj);
    const int pi = patch_origin(i, h), pj = patch_origin(j, w);
    if (y < pi || y > pi + 2 * kPatch || x < pj || x > pj + 2 * kPatch) continue;
    const size_t q = (size_t)i * w + j;
    const real n0 = hp[q], n1 = hp[plane + q], n2 = hp[2 * plane + q];
    const real p = rexp(((n0 * r0 + n1 * r1 + n2 * r2) - stats[(pb + q) * 2]) * inv_t) / stats[(pb + q) * 2 + 1];
    const real ck = p * inv_t * (((real)y - coords[(pb + q) * 2]) * gcoords[(pb + q) * 2] +
                                 ((real)x - coords[(pb + q) * 2 + 1]) * gcoords[(pb + q) * 2 + 1]);
    g0 += ck * n0; g1 += ck * n1; g2 += ck * n2;
  }
  g0 = warp_allsum<real>(g0); g1 = warp_allsum<real>(g1); g2 = warp_allsum<real>(g2);
  if (lane == 0) {
    real* o = d_rays + (size_t)b * 3 * plane + r;
    o[0] = g0; o[plane] = g1; o[2 * plane] = g2;
  }
}

}  // namespace psfm

using namespace psfm;

extern "C" {

int psfm_raysurf_project(int B, int h, int w, int dt, double temperature, const void* dir, const void* rays, void* coords,
                         void* dhat, void* stats, void* stream) {
  if (dt != PSFM_F32 && dt != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "psfm_raysurf_project: dtype must be f32 | f64");
  if (B < 1 || h < kSide || w < kSide)
    return psfm_fail(PSFM_ERR_INVALID, "psfm_raysurf_project: the ray image must be at least %d x %d (got %d x %d)", kSide, kSide, h, w);
  if (!(temperature > 0.0) || !dir || !rays || !coords || !dhat || !stats) return psfm_fail(PSFM_ERR_INVALID, "psfm_raysurf_project: bad argument");
  const unsigned nb = (unsigned)(((size_t)B * h * w + 7) / 8);
  count_launches(1);
  if (dt == PSFM_F64)
    raysurf_fwd_kernel<double><<<nb, 256, 0, (cudaStream_t)stream>>>(B, h, w, 1.0 / temperature, (const double*)dir, (const double*)rays,
                                                                    (double*)coords, (double*)dhat, (double*)stats);
  else
    raysurf_fwd_kernel<float><<<nb, 256, 0, (cudaStream_t)stream>>>(B, h, w, (float)(1.0 / temperature), (const float*)dir, (const float*)rays,
                                                                   (float*)coords, (float*)dhat, (float*)stats);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? PSFM_OK : psfm_fail(PSFM_ERR_CUDA, "psfm_raysurf_project: %s", cudaGetErrorString(e));
}

int psfm_raysurf_project_backward(int B, int h, int w, int dt, double temperature, const void* dir, const void* rays,
                                  const void* coords, const void* dhat, const void* stats, const void* grad_coords, void* d_dir,
                                  void* d_rays, void* stream) {
  if (dt != PSFM_F32 && dt != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "psfm_raysurf_project_backward: dtype must be f32 | f64");
  if (B < 1 || h < kSide || w < kSide || !(temperature > 0.0) || !dir || !rays || !coords || !dhat || !stats || !grad_coords || !d_dir || !d_rays)
    return psfm_fail(PSFM_ERR_INVALID, "psfm_raysurf_project_backward: bad argument");
  const unsigned nb = (unsigned)(((size_t)B * h * w + 7) / 8);
  cudaStream_t st = (cudaStream_t)stream;
  count_launches(2);
  if (dt == PSFM_F64) {
    raysurf_bwd_dir_kernel<double><<<nb, 256, 0, st>>>(B, h, w, 1.0 / temperature, (const double*)dir, (const double*)rays, (const double*)coords,
                                                      (const double*)dhat, (const double*)stats, (const double*)grad_coords, (double*)d_dir);
    raysurf_bwd_rays_kernel<double><<<nb, 256, 0, st>>>(B, h, w, 1.0 / temperature, (const double*)rays, (const double*)coords, (const double*)dhat,
                                                       (const double*)stats, (const double*)grad_coords, (double*)d_rays);
  } else {
    const float it = (float)(1.0 / temperature);
    raysurf_bwd_dir_kernel<float><<<nb, 256, 0, st>>>(B, h, w, it, (const float*)dir, (const float*)rays, (const float*)coords,
                                                     (const float*)dhat, (const float*)stats, (const float*)grad_coords, (float*)d_dir);
    raysurf_bwd_rays_kernel<float><<<nb, 256, 0, st>>>(B, h, w, it, (const float*)rays, (const float*)coords, (const float*)dhat,
                                                      (const float*)stats, (const float*)grad_coords, (float*)d_rays);
  }
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? PSFM_OK : psfm_fail(PSFM_ERR_CUDA, "psfm_raysurf_project_backward: %s", cudaGetErrorString(e));
}

}  // extern "C"
