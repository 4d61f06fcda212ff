// loss_sweep2.cu -- the hot kernel of the photometric loss, second generation: TWO pixels per lane.
//
// Same function, same inputs (prep_kernel's records and packed context images) and same outputs as sweep_kernel of
// loss_sweep.cu -- loss AND every gradient in one pass (reference losses/multiview_photometric_loss.py:199-410 with
// geometry/camera_utils.py:27-59, geometry/camera.py:111-190) -- with a different mapping onto
// the machine, chosen from the round-1 profile (issue slots 56 % busy at 2 warps per scheduler, LSU data pipe 52 %,
// DRAM 6 %: the kernel is bound by instruction issue and by the shuffle / shared-memory pipe, not by bytes):
//
//  * a lane owns two adjacent columns, a warp a strip of 64 (60 produced + 2 halo per side: 6 % halo instead of
//    12.5 %); per-channel values are three packed pairs per lane (px2_math.cuh), so all per-channel arithmetic is
//    FADD2 / FMUL2 / FFMA2 with no idle half (the one-pixel kernel packs channels 0, 1 and leaves channel 2 scalar),
//    the geometry of the two pixels is packed as well, a 3-tap horizontal window costs two shuffles per PAIR, and
//    the per-row bookkeeping (addresses, predicates, barrier, loop) is paid once per 64 pixels;
//  * the vertical 3-row windows are kept as (sum of the two previous rows, previous row) -- 2 register rows per
//    window instead of 3 -- and nothing is rotated;
//  * the warped pixel, its bilinear gradients and the target pixel wait for the adjoint in a shared-memory ring as
//    16-byte vectors (6 per lane and row); depth, projection and the border masks are recomputed from the inverse
//    depth in the adjoint by the same inlined function instead of being parked.
//
// One warp = one context of a strip; the C warps of a CTA meet once per row for the candidate min.
// Restrictions (anything else runs on sweep_kernel): pinhole cameras, no fork pre-map, C in {1, 2, 4}.
#include <cuda_runtime.h>
#include "sweep_common.cuh"
#include "px2_math.cuh"
#include "profile.h"

namespace psfm {

constexpr int kSlotVec = 6;          // 16-byte vectors per lane and ring slot

template <typename real> struct Geo2 {
  P2<real> D, T0, T1, T2, p2, iz, ix, iy;    // T = A0 + A1 v (d p / d D)
};

// decision word of the debug instantiation (include/psfm.h): bits 0-11 floor(ix) + 8, 12-22 floor(iy) + 8 (after the
// mode's clipping), 23-28 L1 signs, 29-31 SSIM clamp-range flags
__device__ __forceinline__ uint32_t decision_word(int x0, int y0) {
  return (uint32_t)min(max(x0 + 8, 0), 4095) | ((uint32_t)min(max(y0 + 8, 0), 2047) << 12);
}

// ---- staging of the pixel records: 1-D bulk copies (TMA) into shared memory, completion on an mbarrier ----------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t a, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t a, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t a, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(a), "r"(parity)
                 : "memory");
  } while (!ok);
}
constexpr int kRecStages = 4;        // record rows staged in shared memory: t - 1, t, and t + 1, t + 2 in flight

template <typename real, int PAD, int VAR, int CT, bool DBG>
__global__ void __launch_bounds__(32 * CT, sizeof(real) == 4 ? 8 / CT : 1) sweep2_kernel(const SweepParams SP) {
  typedef Real4<real> vec4;
  typedef P2<real> pr;
  typedef C3<real> c3;
  constexpr int C = CT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const LossParams& P = SP.P;
  const SweepSet& S = SP.set;
  const SweepConsts<real>& K = sweep_consts<real>(SP);
  const int lane = threadIdx.x & 31, j = threadIdx.x >> 5;
  const int unit = blockIdx.x, nunits = gridDim.x;
  const int strip = unit % S.nstrips2, chunk = unit / S.nstrips2;
  const int si = blockIdx.y, s = S.scale[si], b = blockIdx.z;
  const int H = S.H, W = S.W, plane = H * W;
  const int u0 = strip * kInner2 - 2 + 2 * lane, u1 = u0 + 1;
  const int su0 = reflect1(u0, W), su1 = reflect1(u1, W);
  const bool inner = (lane >= 1) && (lane <= 30);
  const bool real0 = (u0 >= 0) && (u0 < W), real1 = (u1 >= 0) && (u1 < W);
  const bool ok0 = inner && real0, ok1 = inner && real1;
  const bool edge_strip = (strip == 0) || (strip * kInner2 + kInner2 + 2 >= W);
  const int r0 = chunk * S.chunk_h2, r1 = min(r0 + S.chunk_h2, H);

  vec4* ring = (vec4*)smem_raw + (size_t)j * 4 * kSlotVec * 32 + lane;      // [4 slots][kSlotVec][32]
  pr* photo_s = (pr*)((vec4*)smem_raw + (size_t)C * 4 * kSlotVec * 32);      // [2][C][32]
  pr* dpart_s = photo_s + 2 * C * 32;                                        // [2][C][32]
  vec4* rec_s = (vec4*)(dpart_s + 2 * C * 32);                               // [kRecStages][2 halves][32][3 parts]
  const uint32_t mbar0 = smem_u32(rec_s + kRecStages * 192);                 // kRecStages mbarriers

  const ScaleDesc& SD = P.sc[s];
  const vec4* ctxp = (const vec4*)S.ctxp[j] + (size_t)b * S.Hp * S.Wp;
  // pixel records (sweep_common.cuh), [B][H][2][Wh][3]: the strip's segment of a row -- 32 even and 32 odd columns,
  // clipped to the image -- is staged in shared memory by two bulk copies
  const int Wh = S.Wh;
  const vec4* ra = (const vec4*)S.ra + ((size_t)b * H * 2 * Wh + max(strip * (kInner2 / 2) - 1, 0)) * 3;   // + the segment start
  asm volatile("" : "+l"(ra));
  const int h0 = max(strip * (kInner2 / 2) - 1, 0);                // first half-row index of the segment
  const int ne = min(strip * (kInner2 / 2) + 31, (W + 1) / 2) - h0, no = min(strip * (kInner2 / 2) + 31, W / 2) - h0;
  const int pos0 = ((su0 & 1) * 32 + min(max((su0 >> 1) - h0, 0), ((su0 & 1) ? no : ne) - 1)) * 3;
  const int pos1 = ((su1 & 1) * 32 + min(max((su1 >> 1) - h0, 0), ((su1 & 1) ? no : ne) - 1)) * 3;
  asm volatile("" : "+l"(ctxp));   // opaque: one IMAD.WIDE per load off a 32-bit offset
  constexpr bool ups = VAR == 1;                                   // inverse depths nearest-upsampled on load
  const real* invd = (const real*)SD.invd + (size_t)b * SD.Hd * SD.Wd;
  const int Wd = SD.Wd;
  const int sud0 = nearest_src(su0, SD.cs, SD.Wd, ups), sud1 = nearest_src(su1, SD.cs, SD.Wd, ups);
  // `edge`: the row may lie outside the image (reflected / clamped); interior steps skip that arithmetic
  auto load_d = [&](int row, bool edge, int dep) {
    const int o = nearest_src(edge ? reflect1(row, H) : row, SD.rs, SD.Hd, ups) * Wd + dep;
    return pr(__ldg(invd + (o + sud0)), __ldg(invd + (o + sud1)));
  };
  const real* wxy = (const real*)S.wxy + (size_t)b * plane * 2;
  asm volatile("" : "+l"(invd), "+l"(wxy));
  const int Wp = S.Wp;
  const bool smooth = P.smooth_w > 0.0;

  // ---- per-lane geometry constants: p = (A0 + A1 v) D + m with A = ref_K R Kinv (u, v, 1) ----------------
  pr A0[3];
  real A1[3], mt[3];
  {
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
    real a[2][3];
#pragma unroll
    for (int px = 0; px < 2; ++px) {
      const real fu = (real)(px ? su1 : su0);
      const real b0 = cam.ki[0] * fu + cam.ki[2], b1 = cam.ki[3] * fu + cam.ki[5], b2 = cam.ki[6] * fu + cam.ki[8];
#pragma unroll
      for (int i = 0; i < 3; ++i) a[px][i] = M[i * 3] * b0 + M[i * 3 + 1] * b1 + M[i * 3 + 2] * b2;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      A0[i] = pr(a[0][i], a[1][i]);
      A1[i] = M[i * 3] * cam.ki[1] + M[i * 3 + 1] * cam.ki[4] + M[i * 3 + 2] * cam.ki[7];
    }
  }
  // depth, projection: the ONE function both the gather and the adjoint evaluate (identical instructions, so the
  // border decisions of the two agree bit for bit)
  auto geometry = [&](real fv, pr dd) {
    Geo2<real> g;
    g.D = pr(rrcp(rmax(dd.lo(), real(1e-6))), rrcp(rmax(dd.hi(), real(1e-6))));      // utils/depth.py:120
    const pr fvp(fv);
    g.T0 = vfma(pr(A1[0]), fvp, A0[0]);
    g.T1 = vfma(pr(A1[1]), fvp, A0[1]);
    g.T2 = vfma(pr(A1[2]), fvp, A0[2]);
    const pr p0 = vfma(g.T0, g.D, pr(mt[0])), p1 = vfma(g.T1, g.D, pr(mt[1]));
    g.p2 = vfma(g.T2, g.D, pr(mt[2]));
    g.iz = pr(rrcp(rmax(g.p2.lo(), real(1e-5))), rrcp(rmax(g.p2.hi(), real(1e-5))));   // camera.py:179
    g.ix = p0 * g.iz;
    g.iy = p1 * g.iz;
    return g;
  };
  const int mycand = P.automask ? 2 * j : j;
  real inv_m = real(1), cb = real(0);
  if (smooth) {
    const real* st = (const real*)SP.stats[s] + (size_t)b * 4;
    inv_m = real(1) / st[0];
    cb = st[1];
  }

  // ---- window state: (sum of the two previous rows, previous row) of the horizontal sums ------------------
  // forward: m = 0 sums of x, 1 of x^2, 2 of x y; adjoint: the three gradient coefficient planes
  c3 Aw[3], hp[3], AF[3], hFp[3];
#pragma unroll
  for (int m = 0; m < 3; ++m) { Aw[m] = c3(real(0)); hp[m] = c3(real(0)); AF[m] = c3(real(0)); hFp[m] = c3(real(0)); }
  {
    const vec4 z{real(0), real(0), real(0), real(0)};
#pragma unroll
    for (int k = 0; k < 4 * kSlotVec; ++k) ring[k * 32] = z;
  }
  pr S0a[3];                                                       // per pixel: the epilogue weighs them by the column
  real Sva[3], S1a[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { S0a[i] = pr(real(0)); Sva[i] = real(0); S1a[i] = real(0); }
  real loss_acc = real(0);
  pr w_cur(real(0)), w_old(real(0)), ls_p(real(0));
  pr wy_prev(real(0)), dsum_prev(real(0));
  int prev_q = -1;
  real* d_out = (real*)SD.d_full + (size_t)b * plane;
  asm volatile("" : "+l"(d_out));

  // ---- gather pipeline: geometry of `row`, 8 tap loads go in flight; consumed later in the same step -------
  vec4 tap[8];
  pr twx(real(0)), twy(real(0));
  auto start_row = [&](int row, bool edge, pr draw, int dep) {
    const real fv = (real)((edge ? reflect1(row, H) : row) + dep);
    const Geo2<real> g = geometry(fv, draw);
    // tap offset in the packed context image and the bilinear weights (sweep_device.cuh: sample4_issue)
    auto tap_ofs = [&](real ix, real iy, real& wx, real& wy) {
      if (PAD == PSFM_PAD_BORDER) {
        ix = !(ix > real(0)) ? real(0) : rmin(ix, real(W - 1));
        iy = !(iy > real(0)) ? real(0) : rmin(iy, real(H - 1));
      } else {       // fully outside beyond [-1, size]: all four taps land in the zero border
        ix = rmin(rmax(ix, real(-2)), real(W + 1));
        iy = rmin(rmax(iy, real(-2)), real(H + 1));
      }
      const real fx = rfloor(ix), fy = rfloor(iy);
      wx = ix - fx; wy = iy - fy;
      return ((int)fy + kPadLo) * Wp + ((int)fx + kPadLo);
    };
    real wx0, wy0, wx1, wy1;
    const int o0 = tap_ofs(g.ix.lo(), g.iy.lo(), wx0, wy0), o1 = tap_ofs(g.ix.hi(), g.iy.hi(), wx1, wy1);
    tap[0] = ld4(ctxp + o0); tap[1] = ld4(ctxp + o0 + 1); tap[2] = ld4(ctxp + o0 + Wp); tap[3] = ld4(ctxp + o0 + Wp + 1);
    tap[4] = ld4(ctxp + o1); tap[5] = ld4(ctxp + o1 + 1); tap[6] = ld4(ctxp + o1 + Wp); tap[7] = ld4(ctxp + o1 + Wp + 1);
    // the next row's bottom taps are (mostly) one image row further down: pull that row into L1 now
    asm volatile("prefetch.global.L1 [%0];" ::"l"(ctxp + o0 + 2 * Wp));
    twx = pr(wx0, wx1);
    twy = pr(wy0, wy1);
    if (DBG) {
      if (row >= r0 && row < r1 && SD.dbg) {
        uint32_t* dw = SD.dbg + ((size_t)j * P.B + b) * plane + (size_t)row * W;
        auto fl = [&](real ix, real iy) {
          real cx = ix, cy = iy;
          if (PAD == PSFM_PAD_BORDER) {
            cx = !(cx > real(0)) ? real(0) : rmin(cx, real(W - 1));
            cy = !(cy > real(0)) ? real(0) : rmin(cy, real(H - 1));
          } else {
            cx = rmin(rmax(cx, real(-2)), real(W + 1));
            cy = rmin(rmax(cy, real(-2)), real(H + 1));
          }
          return decision_word((int)rfloor(cx), (int)rfloor(cy));
        };
        if (ok0) dw[u0] = fl(g.ix.lo(), g.iy.lo());
        if (ok1) dw[u1] = fl(g.ix.hi(), g.iy.hi());
      }
    }
  };
  c3 xcar(real(0));
  auto finish_row = [&](int row) {
    // per pixel: channels 0, 1 as a pair; channel 2 rides with the (always zero) 4th component through the first
    // subtraction, which keeps that component's register allocated until the load has landed (loss_sweep.cu)
    pr xa[2], gxa[2], gya[2];
    real xb[2], gxb[2], gyb[2];
#pragma unroll
    for (int px = 0; px < 2; ++px) {
      const vec4* tp = tap + 4 * px;
      const real wx = px ? twx.hi() : twx.lo(), wy = px ? twy.hi() : twy.lo();
      bilerp_v<pr>(pr(tp[0].x, tp[0].y), pr(tp[1].x, tp[1].y), pr(tp[2].x, tp[2].y), pr(tp[3].x, tp[3].y), pr(wx), pr(wy),
                   xa[px], gxa[px], gya[px]);
      const pr dt = pr(tp[1].z, tp[1].w) - pr(tp[0].z, tp[0].w), db = pr(tp[3].z, tp[3].w) - pr(tp[2].z, tp[2].w);
      const real dxt = dt.lo() + dt.hi(), dxb = db.lo() + db.hi();       // the high lanes are 0 - 0: consumed on purpose
      const real top = vfma(dxt, wx, tp[0].z), bot = vfma(dxb, wx, tp[2].z);
      gyb[px] = bot - top;
      xb[px] = vfma(gyb[px], wy, top);
      gxb[px] = vfma(dxb - dxt, wy, dxt);
    }
    xcar = c3(xa[0], xa[1], pr(xb[0], xb[1]));
    vec4* rs = ring + (row & 3) * kSlotVec * 32;
    rs[0 * 32] = vec4{xa[0].lo(), xa[0].hi(), xa[1].lo(), xa[1].hi()};
    // slot vector 1 = (x of channel 2, y of channel 2): the y half is written when the row's target pixel arrives
    ((pr*)(rs + 1 * 32))[0] = pr(xb[0], xb[1]);
    rs[3 * 32] = vec4{gxa[0].lo(), gxa[0].hi(), gxa[1].lo(), gxa[1].hi()};
    rs[4 * 32] = vec4{gxb[0], gxb[1], gyb[0], gyb[1]};
    rs[5 * 32] = vec4{gya[0].lo(), gya[0].hi(), gya[1].lo(), gya[1].hi()};
  };

  pr wxn(real(0)), wyn(real(0));                                   // edge weights of row q + 1 (context 0 only)
  auto load_w = [&](int row, bool edge, int dep) {
    const int o = (edge ? min(max(row, 0), H - 1) : row) * W + dep;
    real a0, b0, a1, b1;
    load_pair(wxy + (o + su0) * 2, a0, b0);
    load_pair(wxy + (o + su1) * 2, a1, b1);
    wxn = pr(a0, a1);
    wyn = pr(b0, b1);
  };
  const int t0 = r0 - 2;
  // record row `row` -> stage (row - (t0 - 1)) & 3; issued by one thread, two steps ahead of its use
  const uint32_t rec_s32 = smem_u32(rec_s);
  auto stage_row = [&](int row, bool mine) {
    if (mine) {
      const int k = row - (t0 - 1), st = k & (kRecStages - 1);
      const uint32_t mb = mbar0 + st * 8;
      const vec4* src = ra + (reflect1(row, H) * 2 * Wh) * 3;
      const uint32_t dst = rec_s32 + st * (192 * (int)sizeof(vec4));
      mbar_expect_tx(mb, (uint32_t)(3 * (ne + no) * sizeof(vec4)));
      bulk_g2s(dst, src, (uint32_t)(3 * ne * sizeof(vec4)), mb);
      bulk_g2s(dst + 96 * (int)sizeof(vec4), src + Wh * 3, (uint32_t)(3 * no * sizeof(vec4)), mb);
    }
  };
  auto wait_row = [&](int row) {
    const int k = row - (t0 - 1);
    mbar_wait(mbar0 + (k & (kRecStages - 1)) * 8, (k / kRecStages) & 1);
    return (const vec4*)(rec_s + (k & (kRecStages - 1)) * 192);
  };
  if (threadIdx.x == 0) {
#pragma unroll
    for (int st = 0; st < kRecStages; ++st) mbar_init(mbar0 + st * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  stage_row(t0 - 1, threadIdx.x == 0);
  stage_row(t0, threadIdx.x == 0);
  stage_row(t0 + 1, threadIdx.x == 0);
  if (j == 0 && smooth) load_w(t0 - 2, true, 0);
  pr dnx = load_d(t0, true, 0);
  pr dpre = load_d(t0 + 1, true, 0);
  start_row(t0, true, dnx, 0);
  finish_row(t0);
  const vec4* rec_prev = wait_row(t0 - 1);                         // staged records of row t - 1
  const int never = H >> 31;                                       // 0, unknown to the compiler

  auto finalize_row = [&](int par) {      // context 0 adds the other contexts' parts and stores the row
    if (j == 0 && prev_q >= 0) {
      pr tot = dsum_prev;
#pragma unroll
      for (int jj = 1; jj < C; ++jj) tot = tot + dpart_s[(par * C + jj) * 32 + lane];
      if (ok0) d_out[prev_q * W + su0] = tot.lo();
      if (ok1) d_out[prev_q * W + su1] = tot.hi();
    }
    prev_q = -1;
  };

  // RE: the step touches rows outside [0, H) (reflection, clamps, doubled border rows); steps 3 <= t <= H - 3 do not
  auto step = [&](const int t, auto re) __attribute__((always_inline)) {
    constexpr bool RE = decltype(re)::value;
    const int p = t - 1, q = t - 2;
    const int par = t & 1;
    const bool p_row = !RE || ((p >= 0) && (p < H));
    const bool p_real0 = p_row && real0, p_real1 = p_row && real1;
    // ---------------- staged records: target pixel of row t, statistics / mask / identity candidate of row p ------
    if (C == 1) __syncwarp();
    // its stage held row t - 2, last read before the barrier of step t - 1; the warps take turns at issuing
    if (t + 2 <= r1 + 1) stage_row(t + 2, lane == 0 && j == (C == 1 ? 0 : (t & (C - 1))));
    const vec4* rec_cur = wait_row(t);
    c3 y, my, qy;
    pr idmin, idcand, m;
    {
      const vec4 ya = rec_cur[pos0], yb = rec_cur[pos1];
      y = c3(pr(ya.x, ya.y), pr(yb.x, yb.y), pr(ya.z, yb.z));
      const vec4 ma = rec_prev[pos0], mb = rec_prev[pos1];          // mask of row p in the 4th component
      m = pr(p_real0 ? ma.w : real(0), p_real1 ? mb.w : real(0));   // 1 without a mask
      const vec4 ba = rec_prev[pos0 + 1], bb = rec_prev[pos1 + 1];
      my = c3(pr(ba.x, ba.y), pr(bb.x, bb.y), pr(ba.z, bb.z));
      idmin = pr(ba.w, bb.w);
      const vec4 ca = rec_prev[pos0 + 2], cb2 = rec_prev[pos1 + 2];
      qy = c3(pr(ca.x, ca.y), pr(cb2.x, cb2.y), pr(ca.z, cb2.z));
      idcand = pr(ca.w, cb2.w);
    }
    rec_prev = rec_cur;
    dnx = dpre;
    const pr wxq = wxn, wyq = wyn;
    // All loads of a step share scoreboards, and a wait on one waits for everything in flight on it: every value of
    // the previous batch is therefore consumed BEFORE the next batch is issued -- `dep` (always 0, unknown to the
    // compiler) enters every new address (loss_sweep.cu, prep kernel: without it the one-row prefetch is void).
    const int dep = (value_bits(dnx.lo()) | value_bits(dnx.hi()) | value_bits(wxq.lo()) | value_bits(wxq.hi())) & never;
    // ---------------- prefetch of the next steps' operands, gather of row t + 1 ---------------------------
    dpre = load_d(t + 2, RE, dep);
    if (j == 0 && smooth) load_w(q + 1, RE, dep);
    // inverse depths of the adjoint's row (and, for the smoothness term of context 0, of the rows around it)
    const pr d2 = load_d(q, RE, dep);
    pr d1(real(0)), d3(real(0));
    if (j == 0 && smooth) { d1 = load_d(q + 1, RE, dep); d3 = load_d(q - 1, RE, dep); }
    start_row(t + 1, RE, dnx, dep);
    // ---------------- horizontal window sums of row t, L1 term of row t -----------------------------------
    const c3 x = xcar;
    c3 V[3];                                                       // 3 x 3 window sums at row p
    {
      auto hsums = [&](c3* dst) {
        const c3 z = x * y;
        const Nb3<real> nx = neighbours(x), nz = neighbours(z);
        dst[0] = hsum(x, nx);
        dst[1] = hsumsq(x, nx);
        dst[2] = hsum(z, nz);
      };
      c3 hn[3];
      hsums(hn);
#pragma unroll
      for (int k = 0; k < 3; ++k) { V[k] = Aw[k] + hn[k]; Aw[k] = hp[k] + hn[k]; hp[k] = hn[k]; }
      vec4* rs = ring + (t & 3) * kSlotVec * 32;
      rs[2 * 32] = vec4{y.a0.lo(), y.a0.hi(), y.a1.lo(), y.a1.hi()};
      ((pr*)(rs + 1 * 32))[1] = y.b;
    }
    const pr ls_t = chsum(cabs(x - y));
    if (DBG) {         // L1 signs (bits 23-28, two per channel: 0 negative, 1 zero, 2 positive) of this context at row t
      if (SD.dbg && t >= r0 && t < r1) {
        uint32_t* dw = SD.dbg + ((size_t)j * P.B + b) * plane + (size_t)t * W;
        const c3 e = x - y;
        auto sg = [](real v) { return v > real(0) ? 2u : (v < real(0) ? 0u : 1u); };
        if (ok0) dw[su0] |= (sg(e.a0.lo()) | (sg(e.a0.hi()) << 2) | (sg(e.b.lo()) << 4)) << 23;
        if (ok1) dw[su1] |= (sg(e.a1.lo()) | (sg(e.a1.hi()) << 2) | (sg(e.b.hi()) << 4)) << 23;
      }
    }
    // ---------------- finish the gather of row t + 1 ---------------------------------------------------
    finish_row(t + 1);
    // ---------------- SSIM + L1 at row p, candidate min across the contexts --------------------------
    c3 F[3];
    {
      Ssim3<real> st;
      const c3 dist = ssim_fwd3<real>(V[0], V[1], V[2], my, qy, K.C1, K.C2, st);
      // gradient coefficients for a unit selection weight, formed before the contexts meet: 3 values per channel
      // cross the barrier instead of the 8 SSIM factors
      {
        const pr m9(real(-1) / real(9));
        pr fa0[3], fa1[3], fb[3];
        ssim_grad_w<real, pr>(st.a0, my.a0, m9, fa0);
        ssim_grad_w<real, pr>(st.a1, my.a1, m9, fa1);
        ssim_grad_w<real, pr>(st.b, my.b, m9, fb);
#pragma unroll
        for (int k = 0; k < 3; ++k) F[k] = c3(fa0[k], fa1[k], fb[k]);
      }
      if (DBG) {       // SSIM clamp-range flags (bits 29-31) of this context at row p
        if (SD.dbg && p >= r0 && p < r1 && p_row) {
          uint32_t* dw = SD.dbg + ((size_t)j * P.B + b) * plane + (size_t)p * W;
          auto in01 = [](real d) { return d >= real(0) && d <= real(1); };
          if (ok0) dw[su0] |= ((in01(dist.a0.lo()) ? 1u : 0u) | (in01(dist.a0.hi()) ? 2u : 0u) | (in01(dist.b.lo()) ? 4u : 0u)) << 29;
          if (ok1) dw[su1] |= ((in01(dist.a1.lo()) ? 1u : 0u) | (in01(dist.a1.hi()) ? 2u : 0u) | (in01(dist.b.hi()) ? 4u : 0u)) << 29;
        }
      }
      const pr ds = chsum(c3(pclamp01(dist.a0), pclamp01(dist.a1), pclamp01(dist.b)));
      const pr photo = vfma(pr(K.ssim_w3), ds, pr(K.k1) * ls_p);
      ls_p = ls_t;
      real best[2] = {photo.lo() * m.lo(), photo.hi() * m.hi()};
      int sel[2] = {mycand, mycand};
      if (C > 1) {
        photo_s[(par * C + j) * 32 + lane] = photo;
        asm volatile("bar.sync 1, %0;" ::"n"(32 * C) : "memory");
        best[0] = best[1] = real_inf<real>();
        sel[0] = sel[1] = 0;
#pragma unroll
        for (int jj = 0; jj < C; ++jj) {
          const pr v = photo_s[(par * C + jj) * 32 + lane] * m;
          const int cand = P.automask ? 2 * jj : jj;
          if (v.lo() < best[0]) { best[0] = v.lo(); sel[0] = cand; }
          if (v.hi() < best[1]) { best[1] = v.hi(); sel[1] = cand; }
        }
      }
      // identity candidates (automask): the first minimum wins (reference :269-297, torch.min); +inf without
      const pr vi = idmin * m;
      if (p_real0) {
        const int ki = (int)idcand.lo();
        if (vi.lo() < best[0] || (vi.lo() == best[0] && ki < sel[0])) { best[0] = vi.lo(); sel[0] = ki; }
      }
      if (p_real1) {
        const int ki = (int)idcand.hi();
        if (vi.hi() < best[1] || (vi.hi() == best[1] && ki < sel[1])) { best[1] = vi.hi(); sel[1] = ki; }
      }
      w_old = w_cur;
      w_cur = pr((p_real0 && sel[0] == mycand) ? K.inv_count * m.lo() : real(0),
                 (p_real1 && sel[1] == mycand) ? K.inv_count * m.hi() : real(0));
      if (j == 0 && p >= r0 && p < r1) {
        if (p_real0 && ok0) loss_acc += best[0];
        if (p_real1 && ok1) loss_acc += best[1];
        if (DBG) {
          if (SD.sel) {
            uint8_t* so = SD.sel + (size_t)b * plane + (size_t)p * W;
            if (p_real0 && ok0) so[u0] = (uint8_t)sel[0];
            if (p_real1 && ok1) so[u1] = (uint8_t)sel[1];
          }
        }
      }
      const c3 wf = bc(w_cur * pr(K.ssim_w3));
#pragma unroll
      for (int k = 0; k < 3; ++k) F[k] = F[k] * wf;
    }
    finalize_row(par ^ 1);
    // ---------------- adjoint at row q ----------------------------------------------------------------
    {
      c3 sq[3];
      const real mv2 = (RE && q == H - 2) ? real(2) : real(1), mv0n = (RE && q == 0) ? real(2) : real(1);
      const c3 mv2c(mv2), mv0c(mv0n);
      // 3-tap horizontal sums of the coefficient planes (the reflected border columns count twice)
      auto hfsums = [&](c3* dst) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const Nb3<real> nf = neighbours(F[k]);
          dst[k] = hsum(F[k], nf);
          if (edge_strip) {
            const real e0 = (u1 == 1) ? real(1) : real(0);            // column 0 once more into column 1
            const real e1 = (u0 == W - 2) ? real(1) : real(0);        // column W-1 (own pixel 1) once more into pixel 0
            const real e2 = (u1 == W - 2) ? real(1) : real(0);        // column W-1 (right neighbour) once more into pixel 1
            dst[k].a1 = vfma(pr(e0), F[k].a0, dst[k].a1);
            dst[k].a0 = vfma(pr(e1), F[k].a1, dst[k].a0);
            dst[k].a1 = vfma(pr(e2), nf.ra, dst[k].a1);
            dst[k].b = dst[k].b + pr(e1 * F[k].b.hi(), e0 * F[k].b.lo() + e2 * nf.lrb.hi());
          }
        }
      };
      // vertical: s(q) = mv0 hF(q-1) + hF(q) + mv2 hF(q+1); the reflected border rows count twice
      {
        c3 hF[3];
        hfsums(hF);
#pragma unroll
        for (int k = 0; k < 3; ++k) { sq[k] = vfma(mv2c, hF[k], AF[k]); AF[k] = vfma(mv0c, hFp[k], hF[k]); hFp[k] = hF[k]; }
      }
      const bool q_row = (q >= r0) && (q < r1);
      const bool q_ok0 = q_row && ok0, q_ok1 = q_row && ok1;
      const vec4* rq = ring + (q & 3) * kSlotVec * 32;
      const vec4 v0 = rq[0 * 32], v1 = rq[1 * 32], v2 = rq[2 * 32], v3 = rq[3 * 32], v4 = rq[4 * 32], v5 = rq[5 * 32];
      const c3 xq(pr(v0.x, v0.y), pr(v0.z, v0.w), pr(v1.x, v1.y));
      const c3 yq(pr(v2.x, v2.y), pr(v2.z, v2.w), pr(v1.z, v1.w));
      const c3 gxq(pr(v3.x, v3.y), pr(v3.z, v3.w), pr(v4.x, v4.y));
      const c3 gyq(pr(v5.x, v5.y), pr(v5.z, v5.w), pr(v4.z, v4.w));
      const pr wk = w_old * pr(K.k1);
      const c3 e = xq - yq;
      const c3 sg(pr(signed_by(wk.lo(), e.a0.lo()), signed_by(wk.lo(), e.a0.hi())),
                  pr(signed_by(wk.hi(), e.a1.lo()), signed_by(wk.hi(), e.a1.hi())),
                  pr(signed_by(wk.lo(), e.b.lo()), signed_by(wk.hi(), e.b.hi())));
      const c3 dl = vfma(yq, sq[2], vfma(xq, sq[1], sq[0])) + sg;
      pr gix = chsum(dl * gxq), giy = chsum(dl * gyq);
      gix = pr(q_ok0 ? gix.lo() : real(0), q_ok1 ? gix.hi() : real(0));
      giy = pr(q_ok0 ? giy.lo() : real(0), q_ok1 ? giy.hi() : real(0));
      const real fq = (real)q;
      const Geo2<real> g = geometry(fq, d2);
      pr izx = g.iz, izy = g.iz;
      if (PAD == PSFM_PAD_BORDER) {       // zero coordinate gradient where the sample position was clipped
        auto clipped = [](real v, real hi) { return !(v > real(0)) || v >= hi; };
        izx = pr(clipped(g.ix.lo(), real(W - 1)) ? real(0) : g.iz.lo(), clipped(g.ix.hi(), real(W - 1)) ? real(0) : g.iz.hi());
        izy = pr(clipped(g.iy.lo(), real(H - 1)) ? real(0) : g.iz.lo(), clipped(g.iy.hi(), real(H - 1)) ? real(0) : g.iz.hi());
      }
      const pr dp0 = gix * izx, dp1 = giy * izy;
      const pr t2 = vfma(dp0, g.ix, dp1 * g.iy);
      const pr dp2((g.p2.lo() >= real(1e-5)) ? -t2.lo() : real(0), (g.p2.hi() >= real(1e-5)) ? -t2.hi() : real(0));
      const pr dD = vfma(g.T2, dp2, vfma(g.T1, dp1, g.T0 * dp0));
      const pr mD2 = g.D * g.D;
      const pr dph = mD2 * dD;
      const pr dphoto((q_ok0 && d2.lo() >= real(1e-6)) ? -dph.lo() : real(0), (q_ok1 && d2.hi() >= real(1e-6)) ? -dph.hi() : real(0));
      {
        const pr e0 = dp0 * g.D, e1 = dp1 * g.D, e2 = dp2 * g.D;
        S0a[0] = S0a[0] + e0; S0a[1] = S0a[1] + e1; S0a[2] = S0a[2] + e2;
        Sva[0] = vfma(e0.lo() + e0.hi(), fq, Sva[0]); Sva[1] = vfma(e1.lo() + e1.hi(), fq, Sva[1]);
        Sva[2] = vfma(e2.lo() + e2.hi(), fq, Sva[2]);
        S1a[0] += dp0.lo() + dp0.hi(); S1a[1] += dp1.lo() + dp1.hi(); S1a[2] += dp2.lo() + dp2.hi();
      }
      if (j == 0) {
        pr gs(real(0));
        if (smooth) {
          const real dl0 = shfl_up1(d2.hi()), dr1 = shfl_dn1(d2.lo());      // left of pixel 0, right of pixel 1
          const real wxl0 = shfl_up1(wxq.hi());                              // edge weight wx of the left neighbour of pixel 0
          const real cx = K.coef_x[s], cy = K.coef_y[s];
          real qq0 = real(0), qq1 = real(0);
          if (u0 + 1 < W) qq0 += signed_by(cx * wxq.lo(), d2.lo() - d2.hi());
          if (u0 >= 1) qq0 -= signed_by(cx * wxl0, dl0 - d2.lo());
          if (u1 + 1 < W) qq1 += signed_by(cx * wxq.hi(), d2.hi() - dr1);
          if (u1 >= 1) qq1 -= signed_by(cx * wxq.lo(), d2.lo() - d2.hi());
          if (!RE || q + 1 < H) {
            qq0 += signed_by(cy * wyq.lo(), d2.lo() - d1.lo());
            qq1 += signed_by(cy * wyq.hi(), d2.hi() - d1.hi());
          }
          if (!RE || q >= 1) {
            qq0 -= signed_by(cy * wy_prev.lo(), d3.lo() - d2.lo());
            qq1 -= signed_by(cy * wy_prev.hi(), d3.hi() - d2.hi());
          }
          gs = pr(qq0 * inv_m - cb, qq1 * inv_m - cb);
          wy_prev = wyq;
        }
        if (q_row) {
          if (C > 1) { dsum_prev = dphoto + gs; prev_q = q; }
          else {
            const pr tot = dphoto + gs;
            if (ok0) d_out[q * W + su0] = tot.lo();
            if (ok1) d_out[q * W + su1] = tot.hi();
          }
        }
      } else {
        dpart_s[(par * C + j) * 32 + lane] = dphoto;
      }
    }
  };
#pragma unroll 1
  for (int t = t0; t <= r1 + 1; ++t) {
    if (t >= 3 && t <= H - 3) step(t, IntC<0>());
    else step(t, IntC<1>());
  }
  if (C > 1) {
    __syncthreads();
    finalize_row((r1 + 1) & 1);
  }

  // ---- epilogue: pose partials (converted from the (u, v, 1) basis) and the loss partial -----------
  {
    SweepCam<real> cam;
    sweep_cam<real>(cam, P, b, j, K.sf[s]);
    const real fu0 = (real)su0, fu1 = (real)su1;
    real G[12];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const real eu = fu0 * S0a[i].lo() + fu1 * S0a[i].hi(), ev = Sva[i], e1 = S0a[i].lo() + S0a[i].hi();
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
// host side
// ------------------------------------------------------------------------------------------
bool sweep2_eligible(const psfm_loss_cfg* cfg) {
  if (cfg->camera_model != PSFM_CAM_PINHOLE || cfg->premap) return false;
  return cfg->C == 1 || cfg->C == 2 || cfg->C == 4;
}

template <typename real, int PAD, int VAR, int CT, bool DBG>
static cudaError_t launch2(const SweepParams& SP, dim3 grid, cudaStream_t st) {
  auto kern = sweep2_kernel<real, PAD, VAR, CT, DBG>;
  const size_t smem = (size_t)CT * 4 * kSlotVec * 32 * sizeof(Real4<real>) + (size_t)4 * CT * 32 * sizeof(P2<real>) +
                      (size_t)kRecStages * 192 * sizeof(Real4<real>) + kRecStages * 8;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  kern<<<grid, 32 * CT, smem, st>>>(SP);
  return cudaSuccess;
}

template <typename real, int PAD, int VAR, bool DBG>
static cudaError_t launch2_c(const SweepParams& SP, dim3 grid, cudaStream_t st) {
  switch (SP.P.C) {
    case 1: return launch2<real, PAD, VAR, 1, DBG>(SP, grid, st);
    case 2: return launch2<real, PAD, VAR, 2, DBG>(SP, grid, st);
    default: return launch2<real, PAD, VAR, 4, DBG>(SP, grid, st);
  }
}

template <typename real>
cudaError_t launch_sweep2(const SweepParams& SP, bool upsampled, bool debug, cudaStream_t st) {
  const SweepSet& S = SP.set;
  const dim3 grid(S.nstrips2 * S.nchunks2, S.ns, SP.P.B);
#ifdef PSFM_SWEEP2_DEV      // development: the headline instantiation only (seconds to compile)
  return launch2<float, PSFM_PAD_ZEROS, 0, 2, false>(SP, grid, st);
#else
  const bool border = SP.P.padding == PSFM_PAD_BORDER;
  if (debug) {
    if (upsampled) return cudaErrorNotSupported;
    return border ? launch2_c<real, PSFM_PAD_BORDER, 0, true>(SP, grid, st) : launch2_c<real, PSFM_PAD_ZEROS, 0, true>(SP, grid, st);
  }
  if (upsampled)
    return border ? launch2_c<real, PSFM_PAD_BORDER, 1, false>(SP, grid, st) : launch2_c<real, PSFM_PAD_ZEROS, 1, false>(SP, grid, st);
  return border ? launch2_c<real, PSFM_PAD_BORDER, 0, false>(SP, grid, st) : launch2_c<real, PSFM_PAD_ZEROS, 0, false>(SP, grid, st);
#endif
}

template cudaError_t launch_sweep2<float>(const SweepParams&, bool, bool, cudaStream_t);
template cudaError_t launch_sweep2<double>(const SweepParams&, bool, bool, cudaStream_t);

}  // namespace psfm
