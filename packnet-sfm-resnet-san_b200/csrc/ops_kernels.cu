// ops_kernels.cu -- stand-alone operators of the path:
//   warp_fwd / warp_bwd   view_synthesis (geometry/camera_utils.py:27-59)
//   resize                interpolate_image / match_scales (utils/image.py:117-214)
//   scale_inplace         x *= *scale
#include <cuda_runtime.h>
#include "psfm_device.cuh"
#include "profile.h"

namespace psfm {

constexpr int kOpThreads = 256;

// ------------------------------------------------------------------------------------------
// view synthesis forward: one thread per pixel.  K / ref_K are the intrinsics of this
// resolution (sf = 1); depth is a depth map.
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t>
__global__ void __launch_bounds__(kOpThreads) warp_fwd_kernel(int H, int W, int padding, const img_t* ref,
                                                              const real* depth, const real* K, const real* refK,
                                                              const real* pose, real* out) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W;
  __shared__ CamPair<real> cam;
  __shared__ real Ts[12];
  if (threadIdx.x == 0) setup_cam<real>(cam, K, refK, b, real(1));
  if (threadIdx.x < 12) Ts[threadIdx.x] = pose[(size_t)b * 16 + (threadIdx.x / 4) * 4 + threadIdx.x % 4];
  __syncthreads();
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= plane) return;
  const int v = (int)(i / W), u = (int)(i - (size_t)v * W);
  PixGeom<real> g;
  pixel_geometry_depth<real>(cam, Ts, (real)u, (real)v, depth[(size_t)b * plane + i], H, W, g);
  real w[3];
  bilinear3<real, img_t, false>(ref + (size_t)b * 3 * plane, plane, H, W, padding, g.ix, g.iy, w, nullptr, nullptr);
#pragma unroll
  for (int c = 0; c < 3; ++c) out[((size_t)b * 3 + c) * plane + i] = w[c];
}

// atomicAdd for double exists on sm_60+; one overload set for both reals
template <typename real> __device__ __forceinline__ void atomic_add(real* p, real v) { atomicAdd(p, v); }

// ------------------------------------------------------------------------------------------
// view synthesis backward: d depth per pixel, d pose via block partials, optional d ref_image
// (scatter with atomics; only the fp32/fp64 image case, bf16 images never require grad).
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t>
__global__ void __launch_bounds__(kOpThreads) warp_bwd_kernel(int H, int W, int padding, const img_t* ref,
                                                              const real* depth, const real* K, const real* refK,
                                                              const real* pose, const real* gout, real* d_depth,
                                                              real* pose_part, real* d_ref) {
  const int b = blockIdx.y;
  const size_t plane = (size_t)H * W;
  __shared__ CamPair<real> cam;
  __shared__ real Ts[12];
  __shared__ real scratch[12 * (kOpThreads / 32)];
  if (threadIdx.x == 0) setup_cam<real>(cam, K, refK, b, real(1));
  if (threadIdx.x < 12) Ts[threadIdx.x] = pose[(size_t)b * 16 + (threadIdx.x / 4) * 4 + threadIdx.x % 4];
  __syncthreads();
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  real acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = real(0);
  if (i < plane) {
    const int v = (int)(i / W), u = (int)(i - (size_t)v * W);
    PixGeom<real> g;
    pixel_geometry_depth<real>(cam, Ts, (real)u, (real)v, depth[(size_t)b * plane + i], H, W, g);
    real w[3], gx[3], gy[3], go[3];
    bilinear3<real, img_t, true>(ref + (size_t)b * 3 * plane, plane, H, W, padding, g.ix, g.iy, w, gx, gy);
#pragma unroll
    for (int c = 0; c < 3; ++c) go[c] = gout[((size_t)b * 3 + c) * plane + i];
    const real gix = go[0] * gx[0] + go[1] * gx[1] + go[2] * gx[2];
    const real giy = go[0] * gy[0] + go[1] * gy[1] + go[2] * gy[2];
    d_depth[(size_t)b * plane + i] = geometry_adjoint_depth<real>(cam, Ts, g, gix, giy, acc);
    if (d_ref != nullptr) {
      real ix = g.ix, iy = g.iy;
      if (padding == PSFM_PAD_BORDER) {
        ix = rmin(rmax(ix, real(0)), real(W - 1));
        iy = rmin(rmax(iy, real(0)), real(H - 1));
      }
      real fx = rfloor(ix), fy = rfloor(iy);
      const real wx1 = ix - fx, wx0 = (fx + real(1)) - ix, wy1 = iy - fy, wy0 = (fy + real(1)) - iy;
      fx = rmax(rmin(fx, real(1 << 30)), real(-(1 << 30)));
      fy = rmax(rmin(fy, real(1 << 30)), real(-(1 << 30)));
      const int x0 = (int)fx, y0 = (int)fy;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        real* base = d_ref + ((size_t)b * 3 + c) * plane;
        for (int dy = 0; dy < 2; ++dy)
          for (int dx = 0; dx < 2; ++dx) {
            const int x = x0 + dx, y = y0 + dy;
            if (x >= 0 && x < W && y >= 0 && y < H)
              atomic_add<real>(base + (size_t)y * W + x, go[c] * (dx ? wx1 : wx0) * (dy ? wy1 : wy0));
          }
      }
    }
  }
  block_reduce<real, 12>(acc, scratch);
  if (threadIdx.x < 12) pose_part[((size_t)b * gridDim.x + blockIdx.x) * 12 + threadIdx.x] = acc[threadIdx.x];
}

template <typename real>
__global__ void __launch_bounds__(32) warp_pose_finish_kernel(const real* pose_part, int nblk, real* d_pose) {
  const int b = blockIdx.x, t = threadIdx.x;
  if (t < 12) {
    double a = 0.0;
    for (int i = 0; i < nblk; ++i) a += (double)pose_part[((size_t)b * nblk + i) * 12 + t];
    d_pose[(size_t)b * 16 + t] = (real)a;
  } else if (t < 16) {
    d_pose[(size_t)b * 16 + t] = real(0);
  }
}

// ------------------------------------------------------------------------------------------
// resize of N planes: mode 0 bilinear align_corners=True (F.interpolate, utils/image.py:145),
// mode 1 nearest (torch 'nearest': src = floor(dst * in / out)).
// ------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ void st_val(T* p, double v);
template <> __device__ __forceinline__ void st_val<float>(float* p, double v) { *p = (float)v; }
template <> __device__ __forceinline__ void st_val<double>(double* p, double v) { *p = v; }
template <> __device__ __forceinline__ void st_val<__nv_bfloat16>(__nv_bfloat16* p, double v) { *p = __float2bfloat16((float)v); }

// One axis of torch's upsample kernels: the two source indices and weights of destination index d.
//   mode 0 bilinear align_corners=True : src = d (in - 1) / (out - 1)
//   mode 1 nearest                      : src = min(floor(d * (float)in / out), in - 1), one tap
//   mode 2 bilinear align_corners=False : src = max((d + 0.5) in / out - 0.5, 0)
template <typename acc_t>
__device__ __forceinline__ void axis_taps(int mode, int d, int in, int out, int& i0, int& i1, acc_t& w0, acc_t& w1) {
  if (mode == 1) {
    const float sc = (float)in / (float)out;
    i0 = (int)floorf(d * sc);
    i0 = i0 < in - 1 ? i0 : in - 1;
    i1 = i0; w0 = acc_t(1); w1 = acc_t(0);
    return;
  }
  acc_t f;
  if (mode == 0) f = (out > 1 ? (acc_t)(in - 1) / (acc_t)(out - 1) : acc_t(0)) * d;
  else { f = ((acc_t)d + acc_t(0.5)) * ((acc_t)in / (acc_t)out) - acc_t(0.5); f = f < acc_t(0) ? acc_t(0) : f; }
  i0 = (int)f;
  i0 = i0 < in - 1 ? i0 : in - 1;
  i1 = i0 + (i0 < in - 1 ? 1 : 0);
  w1 = f - i0; w0 = acc_t(1) - w1;
}

// Adjoint of the resize: one thread per SOURCE pixel gathers, in a fixed order, the gradients of the destination pixels
// that read it (no atomics: deterministic).  grad_out [N,Ho,Wo] -> grad_in [N,Hi,Wi], arithmetic type = element type.
template <typename acc_t>
__global__ void __launch_bounds__(kOpThreads) resize_bwd_kernel(int Hi, int Wi, int Ho, int Wo, int mode, const acc_t* gout,
                                                                acc_t* gin) {
  const size_t n = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)Hi * Wi) return;
  const int y = (int)(i / Wi), x = (int)(i - (size_t)y * Wi);
  // destination range whose taps can touch source index s (monotone maps; one extra index on both sides)
  auto range = [&](int s, int in, int out, int& lo, int& hi) {
    double a, b;
    if (mode == 1) { const double sc = (double)in / out; a = s / sc; b = (s + 1) / sc; }
    else if (mode == 0) {
      if (out <= 1 || in <= 1) { lo = 0; hi = out - 1; return; }
      const double r = (double)(in - 1) / (out - 1); a = (s - 1) / r; b = (s + 1) / r;
    } else { const double sc = (double)in / out; a = (s - 0.5) / sc - 0.5; b = (s + 1.5) / sc - 0.5; }
    lo = (int)floor(a) - 2; hi = (int)ceil(b) + 2;
    lo = lo < 0 ? 0 : lo; hi = hi > out - 1 ? out - 1 : hi;
  };
  int ylo, yhi, xlo, xhi;
  range(y, Hi, Ho, ylo, yhi);
  range(x, Wi, Wo, xlo, xhi);
  const acc_t* g = gout + n * (size_t)Ho * Wo;
  acc_t acc = acc_t(0);
  for (int dy = ylo; dy <= yhi; ++dy) {
    int a0, a1; acc_t u0, u1;
    axis_taps<acc_t>(mode, dy, Hi, Ho, a0, a1, u0, u1);
    const acc_t wy = (a0 == y ? u0 : acc_t(0)) + (a1 == y ? u1 : acc_t(0));      // a1 == a0 at the last row: both weights count
    if (wy == acc_t(0)) continue;
    acc_t row = acc_t(0);
    for (int dx = xlo; dx <= xhi; ++dx) {
      int b0, b1; acc_t v0, v1;
      axis_taps<acc_t>(mode, dx, Wi, Wo, b0, b1, v0, v1);
      const acc_t wx = (b0 == x ? v0 : acc_t(0)) + (b1 == x ? v1 : acc_t(0));
      if (wx != acc_t(0)) row += wx * g[(size_t)dy * Wo + dx];
    }
    acc += wy * row;
  }
  gin[n * (size_t)Hi * Wi + i] = acc;
}

template <typename T, typename acc_t>
__global__ void __launch_bounds__(kOpThreads) resize_kernel(int Hi, int Wi, int Ho, int Wo, int mode, const T* in, T* out) {
  const size_t n = blockIdx.y;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)Ho * Wo) return;
  const int y = (int)(i / Wo), x = (int)(i - (size_t)y * Wo);
  const T* src = in + n * (size_t)Hi * Wi;
  if (mode == 2) {
    int y0, y1, x0, x1; acc_t hy, ly, hx, lx;
    axis_taps<acc_t>(2, y, Hi, Ho, y0, y1, hy, ly);
    axis_taps<acc_t>(2, x, Wi, Wo, x0, x1, hx, lx);
    const acc_t v = hy * (hx * ldimg<acc_t>(src, (size_t)y0 * Wi + x0) + lx * ldimg<acc_t>(src, (size_t)y0 * Wi + x1)) +
                    ly * (hx * ldimg<acc_t>(src, (size_t)y1 * Wi + x0) + lx * ldimg<acc_t>(src, (size_t)y1 * Wi + x1));
    st_val<T>(out + n * (size_t)Ho * Wo + i, (double)v);
    return;
  }
  if (mode == 1) {
    // upsample_nearest2d: scale = in / out computed in float, src = min(floor(dst * scale), in - 1)
    const float sy = (float)Hi / (float)Ho, sx = (float)Wi / (float)Wo;
    int yy = (int)floorf(y * sy), xx = (int)floorf(x * sx);
    yy = yy < Hi - 1 ? yy : Hi - 1;
    xx = xx < Wi - 1 ? xx : Wi - 1;
    out[n * (size_t)Ho * Wo + i] = src[(size_t)yy * Wi + xx];
    return;
  }
  // upsample_bilinear2d, align_corners=True: src = dst * (in - 1) / (out - 1)
  const acc_t ry = Ho > 1 ? (acc_t)(Hi - 1) / (acc_t)(Ho - 1) : acc_t(0);
  const acc_t rx = Wo > 1 ? (acc_t)(Wi - 1) / (acc_t)(Wo - 1) : acc_t(0);
  const acc_t fy = ry * y, fx = rx * x;
  int y0 = (int)fy, x0 = (int)fx;
  const int y1 = y0 + (y0 < Hi - 1 ? 1 : 0), x1 = x0 + (x0 < Wi - 1 ? 1 : 0);
  const acc_t ly = fy - y0, lx = fx - x0, hy = acc_t(1) - ly, hx = acc_t(1) - lx;
  const acc_t v = hy * (hx * ldimg<acc_t>(src, (size_t)y0 * Wi + x0) + lx * ldimg<acc_t>(src, (size_t)y0 * Wi + x1)) +
                  ly * (hx * ldimg<acc_t>(src, (size_t)y1 * Wi + x0) + lx * ldimg<acc_t>(src, (size_t)y1 * Wi + x1));
  st_val<T>(out + n * (size_t)Ho * Wo + i, (double)v);
}

template <typename real>
__global__ void __launch_bounds__(kOpThreads) scale_kernel(real* x, const real* scale, size_t count) {
  const real s = *scale;
  if (s == real(1)) return;            // x * 1 == x bit for bit: the usual loss.backward() costs no memory traffic
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x)
    x[i] *= s;
}

// ------------------------------------------------------------------------------------------
// host launchers
// ------------------------------------------------------------------------------------------
template <typename real, typename img_t>
static cudaError_t warp_forward_t(int B, int H, int W, int padding, const void* ref, const void* depth, const void* K,
                                  const void* refK, const void* pose, void* out, cudaStream_t st) {
  const size_t plane = (size_t)H * W;
  dim3 grid((unsigned)((plane + kOpThreads - 1) / kOpThreads), B);
  count_launches(1);
  warp_fwd_kernel<real, img_t><<<grid, kOpThreads, 0, st>>>(H, W, padding, (const img_t*)ref, (const real*)depth,
                                                            (const real*)K, (const real*)refK, (const real*)pose,
                                                            (real*)out);
  return cudaGetLastError();
}

cudaError_t warp_forward_dispatch(int B, int H, int W, int img_dtype, int real_dtype, int padding, const void* ref,
                                  const void* depth, const void* K, const void* refK, const void* pose, void* out,
                                  cudaStream_t st) {
  if (real_dtype == PSFM_F64) return warp_forward_t<double, double>(B, H, W, padding, ref, depth, K, refK, pose, out, st);
  if (img_dtype == PSFM_BF16)
    return warp_forward_t<float, __nv_bfloat16>(B, H, W, padding, ref, depth, K, refK, pose, out, st);
  return warp_forward_t<float, float>(B, H, W, padding, ref, depth, K, refK, pose, out, st);
}

size_t warp_workspace_bytes(int B, int H, int W, int real_dtype) {
  const size_t plane = (size_t)H * W;
  const size_t nblk = (plane + kOpThreads - 1) / kOpThreads;
  return (size_t)B * nblk * 12 * (real_dtype == PSFM_F64 ? 8 : 4);
}

template <typename real, typename img_t>
static cudaError_t warp_backward_t(int B, int H, int W, int padding, const void* ref, const void* depth, const void* K,
                                   const void* refK, const void* pose, const void* gout, void* d_depth, void* d_pose,
                                   void* d_ref, void* ws, cudaStream_t st) {
  const size_t plane = (size_t)H * W;
  const unsigned nblk = (unsigned)((plane + kOpThreads - 1) / kOpThreads);
  count_launches(2);
  warp_bwd_kernel<real, img_t><<<dim3(nblk, B), kOpThreads, 0, st>>>(
      H, W, padding, (const img_t*)ref, (const real*)depth, (const real*)K, (const real*)refK, (const real*)pose,
      (const real*)gout, (real*)d_depth, (real*)ws, (real*)d_ref);
  warp_pose_finish_kernel<real><<<B, 32, 0, st>>>((const real*)ws, (int)nblk, (real*)d_pose);
  return cudaGetLastError();
}

cudaError_t warp_backward_dispatch(int B, int H, int W, int img_dtype, int real_dtype, int padding, const void* ref,
                                   const void* depth, const void* K, const void* refK, const void* pose,
                                   const void* gout, void* d_depth, void* d_pose, void* d_ref, void* ws,
                                   cudaStream_t st) {
  if (real_dtype == PSFM_F64)
    return warp_backward_t<double, double>(B, H, W, padding, ref, depth, K, refK, pose, gout, d_depth, d_pose, d_ref, ws, st);
  if (img_dtype == PSFM_BF16)
    return warp_backward_t<float, __nv_bfloat16>(B, H, W, padding, ref, depth, K, refK, pose, gout, d_depth, d_pose,
                                                 nullptr, ws, st);
  return warp_backward_t<float, float>(B, H, W, padding, ref, depth, K, refK, pose, gout, d_depth, d_pose, d_ref, ws, st);
}

cudaError_t resize_dispatch(int N, int Hi, int Wi, int Ho, int Wo, int dtype, int mode, const void* in, void* out,
                            cudaStream_t st) {
  dim3 grid((unsigned)(((size_t)Ho * Wo + kOpThreads - 1) / kOpThreads), N);
  count_launches(1);
  if (dtype == PSFM_F64)
    resize_kernel<double, double><<<grid, kOpThreads, 0, st>>>(Hi, Wi, Ho, Wo, mode, (const double*)in, (double*)out);
  else if (dtype == PSFM_BF16)
    resize_kernel<__nv_bfloat16, float><<<grid, kOpThreads, 0, st>>>(Hi, Wi, Ho, Wo, mode, (const __nv_bfloat16*)in,
                                                                     (__nv_bfloat16*)out);
  else
    resize_kernel<float, float><<<grid, kOpThreads, 0, st>>>(Hi, Wi, Ho, Wo, mode, (const float*)in, (float*)out);
  return cudaGetLastError();
}

cudaError_t resize_backward_dispatch(int N, int Hi, int Wi, int Ho, int Wo, int dtype, int mode, const void* gout, void* gin,
                                     cudaStream_t st) {
  dim3 grid((unsigned)(((size_t)Hi * Wi + kOpThreads - 1) / kOpThreads), N);
  count_launches(1);
  if (dtype == PSFM_F64)
    resize_bwd_kernel<double><<<grid, kOpThreads, 0, st>>>(Hi, Wi, Ho, Wo, mode, (const double*)gout, (double*)gin);
  else
    resize_bwd_kernel<float><<<grid, kOpThreads, 0, st>>>(Hi, Wi, Ho, Wo, mode, (const float*)gout, (float*)gin);
  return cudaGetLastError();
}

cudaError_t scale_dispatch(void* x, const void* scale, size_t count, int real_dtype, cudaStream_t st) {
  if (count == 0) return cudaSuccess;
  size_t nb = (count + kOpThreads - 1) / kOpThreads;
  if (nb > 148 * 8) nb = 148 * 8;
  count_launches(1);
  if (real_dtype == PSFM_F64)
    scale_kernel<double><<<(unsigned)nb, kOpThreads, 0, st>>>((double*)x, (const double*)scale, count);
  else
    scale_kernel<float><<<(unsigned)nb, kOpThreads, 0, st>>>((float*)x, (const float*)scale, count);
  return cudaGetLastError();
}

}  // namespace psfm
