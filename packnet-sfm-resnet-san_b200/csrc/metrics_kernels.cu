// metrics_kernels.cu -- compute_depth_metrics (reference utils/depth.py:258-447; scale_depth :450-481;
// SURVEY.md section 8(f) N4): the evaluation metrics (abs_rel, sq_rel, rmse, rmse_log, a1, a2, a3) of a batch of
// predicted depth maps against ground truth, with the reference's per-image median scaling, garg crop and the
// two ways of matching the prediction to the ground-truth size -- all on the device, no host read-back.
//
// The reference gathers the valid pixels of every image (boolean indexing), sorts them twice (torch.median) and
// reduces about twenty temporaries.  Here nothing is gathered:
//   init     zeroes the digit histograms and the selection state
//   hist     (per 11-bit digit, most significant first) digit histogram of the order-preserving integer keys of
//            gt and of the resized prediction over the valid pixels whose higher digits equal the prefix found so far
//            (shared-memory histogram per CTA, integer atomics into the per-image histogram: deterministic)
//   select   one warp per (image, array): finds the digit holding the k-th smallest key, k = (n - 1) / 2 -- the
//            lower median, what torch.median returns -- and extends the prefix.  After the last digit the prefix IS
//            the median's bit pattern: exact selection, no sort, 3 passes for fp32 and 6 for fp64.
//   sums     eight per-image sums over the valid pixels with pred * (gt_median / pred_median), per-CTA partials
//   image    fixed-order sum of the partials -> the seven metrics of each image (state block)
//   batch    batch average in image order (divided by B like the reference; images without a valid pixel
//            contribute nothing)
// The prediction is never materialised at the ground-truth size: `pred_at` evaluates F.interpolate(bilinear,
// align_corners=True) ('resize') or the zero-padded 'top-center' placement at each pixel.
// Algorithmic bytes: (passes + 1) x (4 B gt + <= 4 B pred) per pixel; everything else is O(B).
#include <cuda_runtime.h>
#include <stdint.h>
#include "psfm_device.cuh"
#include "profile.h"
#include "../../include/psfm.h"

namespace psfm {

constexpr int kMT = 256;          // threads per CTA
constexpr int kDigitBits = 11;
constexpr int kBins = 1 << kDigitBits;
constexpr int kMetricVals = 8;    // n, |d|/gt, d^2/gt, d^2, dlog^2, a1, a2, a3
constexpr int kStateVals = 12;
constexpr unsigned int kNoBin = 0xffffffffu;
#ifndef PSFM_METRICS_MATCH
#define PSFM_METRICS_MATCH 0   // 1: warp-aggregated atomics for the leading digit (measured slower on B200, profiles/)
#endif

struct MetricsCfg {
  int B, H, W, h, w;
  int scale_mode;                 // 0 'resize', 1 'top-center'
  int top, left;                  // top-center placement
  int y1, y2, x1, x2;             // crop rectangle (whole image when no crop)
  int use_gt_scale;
  double min_depth, max_depth;
};

struct SelState {                 // per (image, array): array 0 = gt, 1 = prediction
  unsigned long long prefix;
  unsigned int k;                 // rank still to find inside the prefix' bucket
  unsigned int n;                 // valid pixels of the image
};

template <typename real> struct KeyOf;
template <> struct KeyOf<float> {
  typedef unsigned int type;
  static constexpr int bits = 32;
  __device__ static __forceinline__ type enc(float v) {
    const unsigned int u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  }
  __device__ static __forceinline__ float dec(unsigned long long k) {
    const unsigned int u = (unsigned int)k;
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
  }
};
template <> struct KeyOf<double> {
  typedef unsigned long long type;
  static constexpr int bits = 64;
  __device__ static __forceinline__ type enc(double v) {
    const unsigned long long u = (unsigned long long)__double_as_longlong(v);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
  }
  __device__ static __forceinline__ double dec(unsigned long long u) {
    return __longlong_as_double((long long)((u >> 63) ? (u & 0x7fffffffffffffffull) : ~u));
  }
};

template <int bits> __host__ __device__ constexpr int num_passes() { return (bits + kDigitBits - 1) / kDigitBits; }
// digit p (0 = most significant) occupies bits [shift, shift + width)
template <int bits> __host__ __device__ __forceinline__ void digit_of(int p, int& shift, int& width) {
  const int np = num_passes<bits>();
  shift = bits - kDigitBits * (p + 1);
  width = kDigitBits;
  if (p == np - 1) { width = bits - kDigitBits * (np - 1); shift = 0; }
}

// the prediction at ground-truth pixel (y, x) of image b (scale_depth, utils/depth.py:450-481)
template <typename real>
__device__ __forceinline__ real pred_at(const MetricsCfg& c, const real* pred, int y, int x) {
  if (c.scale_mode == 1) {                      // zeros outside the placed prediction
    const int yy = y - c.top, xx = x - c.left;
    if (yy < 0 || yy >= c.h || xx < 0 || xx >= c.w) return real(0);
    return __ldg(pred + (size_t)yy * c.w + xx);
  }
  if (c.h == c.H && c.w == c.W) return __ldg(pred + (size_t)y * c.W + x);   // interpolate_image returns its input
  // upsample_bilinear2d, align_corners=True: src = dst * (in - 1) / (out - 1)
  const real ry = c.H > 1 ? (real)(c.h - 1) / (real)(c.H - 1) : real(0);
  const real rx = c.W > 1 ? (real)(c.w - 1) / (real)(c.W - 1) : real(0);
  const real fy = ry * y, fx = rx * x;
  const int y0 = (int)fy, x0 = (int)fx;
  const int yb = y0 + (y0 < c.h - 1 ? 1 : 0), xb = x0 + (x0 < c.w - 1 ? 1 : 0);
  const real ly = fy - y0, lx = fx - x0, hy = real(1) - ly, hx = real(1) - lx;
  return hy * (hx * __ldg(pred + (size_t)y0 * c.w + x0) + lx * __ldg(pred + (size_t)y0 * c.w + xb)) +
         ly * (hx * __ldg(pred + (size_t)yb * c.w + x0) + lx * __ldg(pred + (size_t)yb * c.w + xb));
}

template <typename real>
__device__ __forceinline__ bool valid_at(const MetricsCfg& c, real g, int y, int x) {
  return g > (real)c.min_depth && g < (real)c.max_depth && y >= c.y1 && y < c.y2 && x >= c.x1 && x < c.x2;
}

__global__ void metrics_init_kernel(unsigned int* hist, size_t nhist, SelState* sel, int nsel) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nhist) hist[i] = 0u;
  if (i < (size_t)nsel) { sel[i].prefix = 0ull; sel[i].k = 0u; sel[i].n = 0u; }
}

// Q consecutive pixels of a row per thread and step: one 16-byte load per map for Q = 4 (fp64: two)
template <typename real, int Q> __device__ __forceinline__ void ld_quad(const real* p, real (&v)[Q]);
template <> __device__ __forceinline__ void ld_quad<float, 1>(const float* p, float (&v)[1]) { v[0] = __ldg(p); }
template <> __device__ __forceinline__ void ld_quad<double, 1>(const double* p, double (&v)[1]) { v[0] = __ldg(p); }
template <> __device__ __forceinline__ void ld_quad<float, 4>(const float* p, float (&v)[4]) {
  const float4 t = __ldg(reinterpret_cast<const float4*>(p));
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <> __device__ __forceinline__ void ld_quad<double, 4>(const double* p, double (&v)[4]) {
  const double2 t = __ldg(reinterpret_cast<const double2*>(p)), u = __ldg(reinterpret_cast<const double2*>(p) + 1);
  v[0] = t.x; v[1] = t.y; v[2] = u.x; v[3] = u.y;
}

// ground truth and prediction of the Q pixels starting at pixel i of image b (the prediction straight from memory when
// it already has the ground truth's size, SAME, else evaluated per pixel)
template <typename real, int Q, bool SAME>
__device__ __forceinline__ void load_pixels(const MetricsCfg& c, const real* g_b, const real* p_b, unsigned int i, int y, int x,
                                            real (&g)[Q], real (&q)[Q]) {
  ld_quad<real, Q>(g_b + i, g);
  if (SAME) {
    ld_quad<real, Q>(p_b + i, q);
  } else {
#pragma unroll
    for (int k = 0; k < Q; ++k) q[k] = valid_at<real>(c, g[k], y, x + k) ? pred_at<real>(c, p_b, y, x + k) : real(0);
  }
}

// grid (chunks, B).  Pass p of the radix selection for both arrays at once.
template <typename real, int Q, bool SAME>
__global__ void __launch_bounds__(kMT) metrics_hist_kernel(MetricsCfg c, int p, const real* gt, const real* pred,
                                                           const SelState* sel, unsigned int* hist) {
  typedef KeyOf<real> KO;
  __shared__ unsigned int sh[2 * kBins];
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < 2 * kBins; i += kMT) sh[i] = 0u;
  __syncthreads();
  int shift, width;
  digit_of<KO::bits>(p, shift, width);
  const unsigned long long pre_g = sel[2 * b].prefix, pre_p = sel[2 * b + 1].prefix;
  const unsigned int mask = (1u << width) - 1u;
  const unsigned int plane = (unsigned int)c.H * (unsigned int)c.W;       // < 2^30 (checked by the caller)
  const unsigned int units = plane / Q;                                   // Q divides W
  const real* g_b = gt + (size_t)b * plane;
  const real* p_b = pred + (size_t)b * c.h * c.w;
  // uniform trip count over the CTA: the warp votes below need every lane
  for (unsigned int u0 = blockIdx.x * kMT; u0 < units; u0 += gridDim.x * kMT) {
    const unsigned int u = u0 + threadIdx.x, i = u * Q;
    const unsigned int y = i / (unsigned int)c.W, x = i - y * (unsigned int)c.W;
    real g[Q], q[Q];
#pragma unroll
    for (int k = 0; k < Q; ++k) g[k] = q[k] = real(0);
    if (u < units) load_pixels<real, Q, SAME>(c, g_b, p_b, i, (int)y, (int)x, g, q);
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      // bins of this lane (kNoBin: nothing to count).  p == 0: every key matches the empty prefix (and a shift by the
      // full width is undefined)
      unsigned int bg = kNoBin, bp = kNoBin;
      if (u < units && valid_at<real>(c, g[k], (int)y, (int)x + k)) {
        const unsigned long long kg = KO::enc(g[k]), kp = KO::enc(q[k]);
        if (p == 0 || (kg >> (shift + width)) == pre_g) bg = (unsigned int)(kg >> shift) & mask;
        if (p == 0 || (kp >> (shift + width)) == pre_p) bp = kBins + ((unsigned int)(kp >> shift) & mask);
      }
      if (PSFM_METRICS_MATCH && p == 0) {
        // the leading digit of depths takes few values: lanes with equal bins elect one to add their count
        // (one shared-memory atomic per distinct bin and warp instead of one per pixel)
        const unsigned int pg = __match_any_sync(0xffffffffu, bg), pp = __match_any_sync(0xffffffffu, bp);
        if (bg != kNoBin && (threadIdx.x & 31) == __ffs(pg) - 1) atomicAdd(&sh[bg], (unsigned int)__popc(pg));
        if (bp != kNoBin && (threadIdx.x & 31) == __ffs(pp) - 1) atomicAdd(&sh[bp], (unsigned int)__popc(pp));
      } else {                 // later digits: few keys carry the prefix, and their digits are spread out
        if (bg != kNoBin) atomicAdd(&sh[bg], 1u);
        if (bp != kNoBin) atomicAdd(&sh[bp], 1u);
      }
    }
  }
  __syncthreads();
  unsigned int* h_b = hist + (size_t)b * 2 * kBins;
  for (int i = threadIdx.x; i < 2 * kBins; i += kMT)
    if (sh[i]) atomicAdd(h_b + i, sh[i]);
}

// grid B, 64 threads: warp a of block b extends the prefix of array a of image b by one digit and clears the histogram.
__global__ void __launch_bounds__(64) metrics_select_kernel(int p, int width, unsigned int* hist, SelState* sel) {
  const int b = blockIdx.x, a = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned int* h = hist + ((size_t)b * 2 + a) * kBins;
  constexpr int per = kBins / 32;
  unsigned int cnt[per];
  unsigned int local = 0u;
#pragma unroll
  for (int j = 0; j < per; ++j) { cnt[j] = h[lane * per + j]; local += cnt[j]; }
#pragma unroll
  for (int j = 0; j < per; ++j) h[lane * per + j] = 0u;
  unsigned int incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const unsigned int total = __shfl_sync(0xffffffffu, incl, 31);
  SelState s = sel[2 * b + a];
  if (p == 0) { s.n = total; s.k = total ? (total - 1u) / 2u : 0u; }
  if (total == 0u) {                       // no valid pixel: nothing to select (uniform over the warp)
    if (lane == 0) sel[2 * b + a] = s;
    return;
  }
  const unsigned int excl = incl - local;
  const unsigned int owner = __ballot_sync(0xffffffffu, incl > s.k);
  const int first = __ffs(owner) - 1;
  if (lane == first) {
    unsigned int run = excl;
    int bin = 0;
#pragma unroll
    for (int j = 0; j < per; ++j) {
      if (run + cnt[j] > s.k) { bin = lane * per + j; break; }
      run += cnt[j];
    }
    s.k -= run;
    s.prefix = (s.prefix << width) | (unsigned long long)bin;
    sel[2 * b + a] = s;
  }
}

// true when r lies within 8 ulps of 1.25, 1.25^2 or 1.25^3 (g * (1/q) is within 2 ulps of g / q)
template <typename real> __device__ __forceinline__ bool near_threshold(real r) {
  const real tol = (sizeof(real) == 4 ? real(1.1920929e-7) : real(2.220446049250313e-16)) * real(16);
  return rabs(r - real(1.25)) < tol || rabs(r - real(1.5625)) < tol || rabs(r - real(1.953125)) < tol;
}

// grid (chunks, B): per-CTA partial sums of the metric terms (reference :415-431)
template <typename real, int Q, bool SAME>
__global__ void __launch_bounds__(kMT) metrics_sums_kernel(MetricsCfg c, const real* gt, const real* pred,
                                                           const SelState* sel, double* part) {
  typedef KeyOf<real> KO;
  __shared__ real scratch[kMetricVals * (kMT / 32)];
  const int b = blockIdx.y;
  real scale = real(1);
  if (c.use_gt_scale) scale = KO::dec(sel[2 * b].prefix) / KO::dec(sel[2 * b + 1].prefix);   // gt_median / pred_median
  real v[kMetricVals];
#pragma unroll
  for (int k = 0; k < kMetricVals; ++k) v[k] = real(0);
  const unsigned int plane = (unsigned int)c.H * (unsigned int)c.W;
  const unsigned int units = plane / Q;
  const real* g_b = gt + (size_t)b * plane;
  const real* p_b = pred + (size_t)b * c.h * c.w;
#pragma unroll 2
  for (unsigned int u = blockIdx.x * kMT + threadIdx.x; u < units; u += gridDim.x * kMT) {
    const unsigned int i = u * Q, yu = i / (unsigned int)c.W;
    const int y = (int)yu, x = (int)(i - yu * (unsigned int)c.W);
    real gq[Q], qq[Q];
    load_pixels<real, Q, SAME>(c, g_b, p_b, i, y, x, gq, qq);
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      const real g = gq[k];
      if (!valid_at<real>(c, g, y, x + k)) continue;
      real q = qq[k];
      if (c.use_gt_scale) q = q * scale;
      // two reciprocals serve the four quotients of the reference; thresh = max(g/q, q/g) < t <=> both below t (false on
      // NaN), and a ratio within a few ulps of a threshold is re-formed as the reference's exact quotient, so the counts
      // a1..a3 are those of IEEE division
      const real ig = real(1) / g, iq = real(1) / q;
      real r1 = g * iq, r2 = q * ig;
      const real d = g - q, dl = log(r1);              // log(g) - log(q) up to rounding
      if (near_threshold<real>(r1)) r1 = g / q;
      if (near_threshold<real>(r2)) r2 = q / g;
      v[0] += real(1);
      v[1] += rabs(d) * ig;
      v[2] += d * d * ig;
      v[3] += d * d;
      v[4] += dl * dl;
      v[5] += (r1 < real(1.25) && r2 < real(1.25)) ? real(1) : real(0);
      v[6] += (r1 < real(1.5625) && r2 < real(1.5625)) ? real(1) : real(0);
      v[7] += (r1 < real(1.953125) && r2 < real(1.953125)) ? real(1) : real(0);
    }
  }
  block_reduce<real, kMetricVals>(v, scratch);
  if (threadIdx.x < kMetricVals)
    part[((size_t)b * gridDim.x + blockIdx.x) * kMetricVals + threadIdx.x] = (double)v[threadIdx.x];
}

// grid B: warp w of CTA b sums value w of image b over the chunks (lane-strided, then a fixed shuffle tree), lane 0 of
// warp 0 forms the image's metrics into state[b].
template <typename real>
__global__ void __launch_bounds__(kMT) metrics_image_kernel(MetricsCfg c, int nchunk, const double* part,
                                                            const SelState* sel, double* state) {
  typedef KeyOf<real> KO;
  __shared__ double s[kMetricVals];
  const int b = blockIdx.x, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double a = 0.0;
  for (int i = lane; i < nchunk; i += 32) a += part[((size_t)b * nchunk + i) * kMetricVals + w];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
  if (lane == 0) s[w] = a;
  __syncthreads();
  if (threadIdx.x != 0) return;
  const double n = s[0];
  double m[7] = {0, 0, 0, 0, 0, 0, 0};
  if (n > 0) {
    m[0] = s[1] / n; m[1] = s[2] / n; m[2] = sqrt(s[3] / n); m[3] = sqrt(s[4] / n);
    // (thresh < t).float().mean(): a float32 quotient whatever the type of the maps (reference :417-419)
    m[4] = (double)((float)s[5] / (float)n); m[5] = (double)((float)s[6] / (float)n); m[6] = (double)((float)s[7] / (float)n);
  }
  double* st = state + (size_t)b * kStateVals;
  const bool scaled = c.use_gt_scale && n > 0;
  st[0] = n;
  st[1] = scaled ? (double)KO::dec(sel[2 * b].prefix) : 0.0;
  st[2] = scaled ? (double)KO::dec(sel[2 * b + 1].prefix) : 0.0;
  st[3] = scaled ? (double)(KO::dec(sel[2 * b].prefix) / KO::dec(sel[2 * b + 1].prefix)) : 1.0;
  for (int k = 0; k < 7; ++k) st[4 + k] = m[k];
  st[11] = 0.0;
}

// one warp: batch average in image order (lane k owns metric k).  The first four accumulate in double, a1..a3 in float32
// like the reference's float32 tensors (:417-419, :445-447).
template <typename real>
__global__ void __launch_bounds__(32) metrics_batch_kernel(int B, const double* state, real* out) {
  const int k = threadIdx.x;
  if (k >= 7) return;
  if (k < 4) {
    double a = 0.0;
    for (int b = 0; b < B; ++b) a += state[(size_t)b * kStateVals + 4 + k];
    out[k] = (real)(a / (double)B);
  } else {
    float a = 0.f;
    for (int b = 0; b < B; ++b) a += (float)state[(size_t)b * kStateVals + 4 + k];
    out[k] = (real)(a / (float)B);
  }
}

static inline size_t al256m(size_t x) { return (x + 255) / 256 * 256; }
static inline int metrics_chunks(int B, int H, int W) {
  // about 4 CTAs per SM over the batch, at least 16 pixels per thread
  const size_t plane = (size_t)H * W;
  size_t want = (148u * 4u + (size_t)B - 1) / (size_t)B;
  const size_t most = (plane + (size_t)kMT * 16 - 1) / ((size_t)kMT * 16);
  if (want > most) want = most;
  return (int)(want < 1 ? 1 : want);
}

}  // namespace psfm

int psfm_fail(int code, const char* fmt, ...);      // psfm_abi.cu
using namespace psfm;

extern "C" {

size_t psfm_depth_metrics_workspace_bytes(int B, int H, int W) {
  if (B < 1 || H < 1 || W < 1) return 0;
  return al256m((size_t)B * 2 * kBins * sizeof(unsigned int)) + al256m((size_t)B * 2 * sizeof(SelState)) +
         al256m((size_t)B * metrics_chunks(B, H, W) * kMetricVals * sizeof(double));
}

int psfm_depth_metrics(int B, int H, int W, int h, int w, int dt, int scale_mode, int crop, int use_gt_scale,
                       double min_depth, double max_depth, const void* gt, const void* pred, void* out, void* state,
                       void* workspace, size_t workspace_bytes, void* stream) {
  if (dt != PSFM_F32 && dt != PSFM_F64) return psfm_fail(PSFM_ERR_INVALID, "psfm_depth_metrics: real_dtype must be f32|f64");
  if (B < 1 || H < 1 || W < 1 || h < 1 || w < 1 || !gt || !pred || !out || !state)
    return psfm_fail(PSFM_ERR_INVALID, "psfm_depth_metrics: bad argument");
  if (scale_mode != 0 && scale_mode != 1)
    return psfm_fail(PSFM_ERR_INVALID, "psfm_depth_metrics: scale function %d not implemented (0 resize, 1 top-center)", scale_mode);
  if ((size_t)H * W >= ((size_t)1 << 30) || (size_t)h * w >= ((size_t)1 << 30))
    return psfm_fail(PSFM_ERR_INVALID, "psfm_depth_metrics: maps of up to 2^30 pixels are supported");
  if (scale_mode == 1 && (h > H || w > W))
    return psfm_fail(PSFM_ERR_INVALID, "psfm_depth_metrics: top-center needs a prediction no larger than the ground truth (%dx%d vs %dx%d)", h, w, H, W);
  if (!workspace || workspace_bytes < psfm_depth_metrics_workspace_bytes(B, H, W))
    return psfm_fail(PSFM_ERR_WORKSPACE, "psfm_depth_metrics: workspace too small");
  MetricsCfg c;
  c.B = B; c.H = H; c.W = W; c.h = h; c.w = w;
  c.scale_mode = scale_mode;
  c.top = H - h; c.left = (W - w) / 2;
  c.y1 = 0; c.y2 = H; c.x1 = 0; c.x2 = W;
  if (crop) {                         // garg crop (reference :335-339), truncation of the double products like int()
    c.y1 = (int)(0.40810811 * H); c.y2 = (int)(0.99189189 * H);
    c.x1 = (int)(0.03594771 * W); c.x2 = (int)(0.96405229 * W);
  }
  c.use_gt_scale = use_gt_scale ? 1 : 0;
  c.min_depth = min_depth; c.max_depth = max_depth;
  char* wsp = (char*)workspace;
  unsigned int* hist = (unsigned int*)wsp;
  wsp += al256m((size_t)B * 2 * kBins * sizeof(unsigned int));
  SelState* sel = (SelState*)wsp;
  wsp += al256m((size_t)B * 2 * sizeof(SelState));
  double* part = (double*)wsp;
  const int nchunk = metrics_chunks(B, H, W);
  const dim3 grid((unsigned)nchunk, (unsigned)B);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t nhist = (size_t)B * 2 * kBins;
  const int npass = use_gt_scale ? (dt == PSFM_F64 ? num_passes<64>() : num_passes<32>()) : 0;
  count_launches(4 + 2 * npass);
  metrics_init_kernel<<<(unsigned)((nhist + 255) / 256), 256, 0, st>>>(hist, nhist, sel, 2 * B);
  // 4 pixels per thread and 16-byte loads when rows hold whole quads and the maps are aligned for them
  const bool same = scale_mode == 0 && h == H && w == W;
  const bool quad = W % 4 == 0 && ((uintptr_t)gt & 15) == 0 && (!same || ((uintptr_t)pred & 15) == 0);
#define PSFM_METRICS_LAUNCH(KERNEL, ...)                                                                         \
  do {                                                                                                            \
    if (dt == PSFM_F64) {                                                                                         \
      typedef double real;                                                                                        \
      if (quad && same) KERNEL<real, 4, true><<<grid, kMT, 0, st>>>(__VA_ARGS__);                                  \
      else if (quad) KERNEL<real, 4, false><<<grid, kMT, 0, st>>>(__VA_ARGS__);                                    \
      else KERNEL<real, 1, false><<<grid, kMT, 0, st>>>(__VA_ARGS__);                                              \
    } else {                                                                                                      \
      typedef float real;                                                                                         \
      if (quad && same) KERNEL<real, 4, true><<<grid, kMT, 0, st>>>(__VA_ARGS__);                                  \
      else if (quad) KERNEL<real, 4, false><<<grid, kMT, 0, st>>>(__VA_ARGS__);                                    \
      else KERNEL<real, 1, false><<<grid, kMT, 0, st>>>(__VA_ARGS__);                                              \
    }                                                                                                             \
  } while (0)
  for (int p = 0; p < npass; ++p) {
    int shift, width;
    if (dt == PSFM_F64) digit_of<64>(p, shift, width); else digit_of<32>(p, shift, width);
    PSFM_METRICS_LAUNCH(metrics_hist_kernel, c, p, (const real*)gt, (const real*)pred, sel, hist);
    metrics_select_kernel<<<B, 64, 0, st>>>(p, width, hist, sel);
  }
  PSFM_METRICS_LAUNCH(metrics_sums_kernel, c, (const real*)gt, (const real*)pred, sel, part);
#undef PSFM_METRICS_LAUNCH
  if (dt == PSFM_F64) {
    metrics_image_kernel<double><<<B, kMT, 0, st>>>(c, nchunk, part, sel, (double*)state);
    metrics_batch_kernel<double><<<1, 32, 0, st>>>(B, (const double*)state, (double*)out);
  } else {
    metrics_image_kernel<float><<<B, kMT, 0, st>>>(c, nchunk, part, sel, (double*)state);
    metrics_batch_kernel<float><<<1, 32, 0, st>>>(B, (const double*)state, (float*)out);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return psfm_fail(PSFM_ERR_CUDA, "psfm_depth_metrics: %s", cudaGetErrorString(e));
  return PSFM_OK;
}

}  // extern "C"
