// sweep_common.cuh -- declarations shared by the row-sweep loss kernels (loss_sweep.cu: prep / stats / final and the
// one-pixel-per-lane sweep; loss_sweep2.cu: the two-pixels-per-lane sweep).
#pragma once
#include <cuda_runtime.h>
#include "sweep_device.cuh"
#include "loss_params.h"
#include "loss_tile.cuh"

namespace psfm {

constexpr int kPrepVals = 3;         // sum d, sum |dx d| wx, sum |dy d| wy
constexpr int kPrepInner = 30;       // columns a prep warp produces (32 lanes - 1 halo each side)
constexpr int kPrepWarps = 4;        // independent warps per prep CTA
constexpr int kInner = 28;           // columns a sweep warp produces (32 lanes - 2 halo each side)
constexpr int kInner2 = 60;          // two-pixels-per-lane sweep (loss_sweep2.cu): 64 columns - 2 halo each side

constexpr int kPadLo = 2;            // zero border of the packed context images: taps reach -2 .. size + 2
constexpr int kPadHi = 3;

struct SweepSet {                    // scales of equal resolution sharing the same images
  int ns;
  int scale[PSFM_MAX_SCALES];
  int H, W;
  int nstrips, nchunks, chunk_h;     // sweep kernel decomposition
  int nstrips2, nchunks2, chunk_h2;  // two-pixels-per-lane sweep kernel decomposition
  int pstrips, pchunks, pchunk_h;    // prep kernel decomposition
  int Hp, Wp;                        // padded size of the packed context images
  int Wh;                            // (W + 1) / 2: pitch of a half row of the pixel records
  // pixel records, real4 [B][H][2][Wh][3]: a row holds its even columns, then its odd columns, three 4-vectors per
  // pixel (rec_ofs) -- the pixels of the lanes of the two-pixels-per-lane sweep sit in two contiguous runs, which two
  // bulk copies stage in shared memory, and the 48-byte pitch keeps 16-byte shared loads free of bank conflicts.
  // ra, rb = ra + 1, rc = ra + 2 (in 4-vectors) address the three parts with the same offset
  void* ra;                          // target pixel y x3, mask value
  void* rb;                          // mean(y) x3, min identity photometric value (automask)
  void* rc;                          // E[y^2]-mean(y)^2+C2 x3, candidate index of that min
  void* wxy;                         // real [B][H][W][2]: exp(-mean|dx I|), exp(-mean|dy I|)
  void* ray;                         // real [B][H][W][2]: fisheye only, x and y of the pixel's viewing ray (z = 1)
  void* ctxp[PSFM_MAX_CTX];          // real4 [B][Hp][Wp]: context image (r, g, b, 0), zero border
};

// element offset of pixel (row, u) in a record plane of one image
__device__ __forceinline__ int rec_ofs(int row, int u, int Wh) { return ((row * 2 + (u & 1)) * Wh + (u >> 1)) * 3; }

template <int V> struct IntC { static constexpr int value = V; };

// 4-vector of the arithmetic type with 16-byte vector loads / stores
template <typename real> struct alignas(16) Real4 { real x, y, z, w; };
__device__ __forceinline__ Real4<float> ld4(const Real4<float>* p) {
  const float4 v = __ldg((const float4*)p);
  return Real4<float>{v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ Real4<double> ld4(const Real4<double>* p) {
  const double2 a = __ldg((const double2*)p), b = __ldg((const double2*)p + 1);
  return Real4<double>{a.x, a.y, b.x, b.y};
}
__device__ __forceinline__ void st4(Real4<float>* p, float x, float y, float z, float w) {
  *(float4*)p = make_float4(x, y, z, w);
}
__device__ __forceinline__ void st4(Real4<double>* p, double x, double y, double z, double w) {
  ((double2*)p)[0] = make_double2(x, y);
  ((double2*)p)[1] = make_double2(z, w);
}

template <typename real> struct SweepConsts {   // hyper-parameters pre-converted to the arithmetic type
  real ssim_w3;                      // ssim_loss_weight / 3
  real k1;                           // (1 - ssim_loss_weight) / 3
  real C1, C2;
  real inv_count;                    // weight of one pixel in the photometric mean (same for a set)
  real premap_lo, premap_hi;
  real pad_;
  real coef_x[PSFM_MAX_SCALES], coef_y[PSFM_MAX_SCALES], sf[PSFM_MAX_SCALES];   // per scale id
};

struct SweepParams {
  LossParams P;
  SweepSet set;
  SweepConsts<float> cf;
  SweepConsts<double> cd;
  // per scale (indexed by the scale id)
  double* prep_part[PSFM_MAX_SCALES];   // [B][punits][kPrepVals]
  void* stats[PSFM_MAX_SCALES];         // real [B][4]: clamped mean, coupling constant, SX, SY
  double* loss_part[PSFM_MAX_SCALES];   // [B][nunits] photometric sums
  void* pose_part[PSFM_MAX_SCALES];     // real [B][nunits][C][12]
};

// ------------------------------------------------------------------------------------------
// prep kernel: grid (ceil(punits / kPrepWarps), B), 32 * kPrepWarps threads; every warp is an
// independent (strip, chunk) unit.  Lane l is column c0 - 1 + l; step t loads row t + 1 and finishes
// row p = t - 1.  One pass over the rows handles the CT contexts [j0, j0 + CT) together (packed copy,
// identity-reprojection term); the pass with j0 == 0 also emits the image-only outputs.
// Per-channel arithmetic runs on (channel 0, channel 1) pairs + a scalar channel 2 (pair_math.cuh).
// ------------------------------------------------------------------------------------------
// F.interpolate(mode='nearest') source index (utils/image.py:148-176 via models/model_utils.py:152-176)
__device__ __forceinline__ int nearest_src(int dst, float scale, int size, bool active) {
  return active ? min((int)floorf((float)dst * scale), size - 1) : dst;
}

__device__ __forceinline__ int value_bits(float v) { return __float_as_int(v); }
__device__ __forceinline__ int value_bits(double v) { return __double2hiint(v); }

template <typename real> struct SweepCam {
  real ki[9];   // reference `Kinv` (camera.py:73-81)
  real rk[9];   // scaled ref_K
  real T[12];   // rows of the 3x4 [R|t] (pose.py:80-86)
};

template <typename real>
__device__ __forceinline__ void sweep_cam(SweepCam<real>& c, const LossParams& P, int b, int j, real sf) {
  CamPair<real> cp;
  setup_cam<real>(cp, (const real*)P.K, (const real*)P.refK, b, sf);
#pragma unroll
  for (int i = 0; i < 9; ++i) { c.ki[i] = cp.ki[i]; c.rk[i] = cp.rk[i]; }
  const real* T = (const real*)P.pose[j] + (size_t)b * 16;
#pragma unroll
  for (int i = 0; i < 12; ++i) c.T[i] = T[i];
}


template <typename real> __device__ __forceinline__ const SweepConsts<real>& sweep_consts(const SweepParams& SP);
template <> __device__ __forceinline__ const SweepConsts<float>& sweep_consts<float>(const SweepParams& SP) { return SP.cf; }
template <> __device__ __forceinline__ const SweepConsts<double>& sweep_consts<double>(const SweepParams& SP) { return SP.cd; }

__device__ __forceinline__ void load_pair(const float* p, float& a, float& b) {
  const float2 v = __ldg((const float2*)p);
  a = v.x; b = v.y;
}
__device__ __forceinline__ void load_pair(const double* p, double& a, double& b) {
  const double2 v = __ldg((const double2*)p);
  a = v.x; b = v.y;
}

// mag * sign(x) with sign(0) = 0
__device__ __forceinline__ float signed_by(float mag, float x) {
  const float s = __int_as_float(__float_as_int(mag) ^ (__float_as_int(x) & 0x80000000));
  return x == 0.0f ? 0.0f : s;
}
__device__ __forceinline__ double signed_by(double mag, double x) {
  return x > 0.0 ? mag : (x < 0.0 ? -mag : 0.0);
}

// loss_sweep2.cu
struct SweepParams;
bool sweep2_eligible(const psfm_loss_cfg* cfg);
template <typename real> cudaError_t launch_sweep2(const SweepParams& SP, bool upsampled, bool debug, cudaStream_t st);

}  // namespace psfm
