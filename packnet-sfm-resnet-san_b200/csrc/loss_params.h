// loss_params.h -- by-value kernel parameter block and workspace layout of the loss kernels.
#pragma once
#include <stdint.h>
#include <stddef.h>
#include "../../include/psfm.h"

namespace psfm {

// Output tile of one CTA in the loss kernels.
constexpr int kTileW = 32;
constexpr int kTileH = 8;
constexpr int kThreads = 256;

struct ScaleDesc {
  const void* image;                 // [B,3,H,W]
  const void* ctx[PSFM_MAX_CTX];     // [B,3,H,W]
  const void* invd;                  // [B,1,H,W] real
  const void* mask;                  // [B,1,H,W] or null
  uint8_t* sel;                      // [B,H,W]
  uint32_t* dbg;                     // [C][B][H][W] decision words of the debug instantiations, or null
  void* d_invd;                      // [B,1,H,W] real
  int H, W;
  int Hd, Wd;                        // size of the inverse-depth map in memory (== H, W unless upsampled in-kernel)
  float rs, cs;                      // Hd / H, Wd / W: nearest source index = min(floor(dst * scale), size - 1)
  void* d_full;                      // where the kernels write d/d inv_depth at (H, W): d_invd, or a workspace buffer
  int tiles_x, tiles_y;
  double sf;                         // W_s / W (python float division, utils: warp_ref_image)
  double inv_count;                  // 1 / (n * B * H * W): weight of one pixel in the photometric mean
  double coef_x, coef_y;             // smooth_w / (n * 2^s * B*H*(W-1)), smooth_w / (n * 2^s * B*(H-1)*W)
  size_t part_off;                   // offset (in doubles) of this scale's loss partials [B][tiles][kLossVals]
  size_t pose_off;                   // offset (in reals) of this scale's pose partials [B][tiles][C][12]
  size_t clip_off;                   // offset (in doubles) of this scale's clip statistics [B][tiles][groups][2]
};

constexpr int kLossVals = 3;         // photometric sum, smoothness-x sum, smoothness-y sum

struct LossParams {
  int B, C, n;
  int H, W;                          // full resolution
  int padding, reduce_op, automask, premap;
  int camera;                        // psfm_camera
  int ncand;                         // candidates per pixel (C or 2C; x3 when ssim_w == 0)
  double ssim_w, C1, C2, clip_loss, smooth_w;
  double premap_lo, premap_hi;
  ScaleDesc sc[PSFM_MAX_SCALES];
  const void* K;
  const void* refK;
  const void* pose[PSFM_MAX_CTX];
  void* d_pose[PSFM_MAX_CTX];
  // workspace views
  void* means;                       // real [n][B]     per-image mean of the raw inverse-depth input
  void* smooth_img;                  // real [n][B][2]  per-image sums of |gx wx|, |gy wy|
  double* loss_part;                 // double partials
  void* pose_part;                   // real partials
  void* clip_thr;                    // real [n][ncand] clip thresholds (+inf when disabled)
  double* clip_part;                 // double [n][B*tiles][ncand][2] (sum, sum of squares)
  const void* grad_loss;             // real[1] upstream gradient (backward only)
  void* loss_out;                    // real[3]
  void* one;                         // real[1] == 1, written by the loss finish kernel
};

}  // namespace psfm
