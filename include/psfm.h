/*
 * psfm.h -- C ABI of libpsfm_b200.so: the multi-view photometric loss with differentiable
 * view synthesis (forward + backward) as hand-written CUDA kernels for sm_100a.
 *
 * The reference (YoungseokOh/packnet-sfm-resnet-san) is 100 % Python/PyTorch and has no
 * FFI of its own; each entry point below therefore cites the reference *Python* interface
 * it replaces (file:line relative to the reference root).  INTEGRATION.md shows the ctypes
 * binding a reference maintainer would add.
 *
 * Conventions
 *   - every data pointer is a DEVICE pointer to a caller-owned, contiguous NCHW buffer;
 *   - nothing is allocated or freed by the library; scratch lives in a caller-provided
 *     workspace of psfm_loss_workspace_bytes() bytes;
 *   - all work is enqueued asynchronously on `stream` (a cudaStream_t passed as void*);
 *     the library never synchronises;
 *   - every function returns 0 on success and a negative psfm_status on failure;
 *     psfm_last_error() returns a thread-local human-readable message;
 *   - `real` below is the arithmetic type selected by cfg->real_dtype (fp32 in production,
 *     fp64 for the math-parity tests).  Inverse depths, K, poses, losses and gradients are
 *     `real`; images are stored as cfg->img_dtype (fp32, bf16 or fp64) and converted on load.
 *   - reductions are deterministic: per-block partials in fixed slots + fixed-order finish.
 */
#ifndef PSFM_H_
#define PSFM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PSFM_ABI_VERSION 4
#define PSFM_MAX_SCALES 4
#define PSFM_MAX_CTX 8

typedef enum psfm_status {
  PSFM_OK = 0,
  PSFM_ERR_INVALID = -1,      /* bad argument / unsupported configuration */
  PSFM_ERR_CUDA = -2,         /* a CUDA runtime call failed (see psfm_last_error) */
  PSFM_ERR_WORKSPACE = -3     /* workspace too small */
} psfm_status;

/* PSFM_U8: image bytes as decoded from the file; the kernels use byte / 255 in fp32 -- bit-identical to
 * torchvision's to_tensor, which the reference's data pipeline applies on the CPU (datasets/transforms.py) -- so the
 * host->device copy moves 1 byte per sample instead of 4.  A mask of this type holds 0 / 1 bytes.  Accepted by
 * psfm_loss_forward_backward when every scale is at image resolution (sweep kernels, fp32 arithmetic). */
typedef enum psfm_dtype { PSFM_F32 = 0, PSFM_BF16 = 1, PSFM_F64 = 2, PSFM_U8 = 3 } psfm_dtype;
typedef enum psfm_padding { PSFM_PAD_ZEROS = 0, PSFM_PAD_BORDER = 1 } psfm_padding;
typedef enum psfm_reduce { PSFM_REDUCE_MIN = 0, PSFM_REDUCE_MEAN = 1 } psfm_reduce;
/* Camera model of the view synthesis.  PINHOLE: geometry/camera.py:15-190 (Camera).  FISHEYE: the fork's
 * VADAS model, geometry/camera.py:194-394 (FisheyeCamera), as its warp_ref_image builds the cameras
 * (losses/multiview_photometric_loss.py:131-195): K / ref_K then hold [B,11] rows
 * (k0..k6, s, div, ux, uy) -- the 'distortion_coeffs' dict of datasets/ncdb_dataset.py:577-584. */
typedef enum psfm_camera { PSFM_CAM_PINHOLE = 0, PSFM_CAM_FISHEYE = 1 } psfm_camera;

/* Hyper-parameters and shapes of one loss evaluation.
 * Mirrors MultiViewPhotometricLoss.__init__ (losses/multiview_photometric_loss.py:92-118)
 * and the shapes of MultiViewPhotometricLoss.forward (:331-360). */
typedef struct psfm_loss_cfg {
  int32_t B;                      /* batch                                                    */
  int32_t C;                      /* number of context frames (1..PSFM_MAX_CTX)               */
  int32_t n;                      /* number of scales in use (1..PSFM_MAX_SCALES)             */
  int32_t H, W;                   /* size of the full-resolution image (K/ref_K refer to it)  */
  int32_t Hs[PSFM_MAX_SCALES];    /* inv_depths[i].shape[2]                                   */
  int32_t Ws[PSFM_MAX_SCALES];    /* inv_depths[i].shape[3]                                   */
  int32_t Hd[PSFM_MAX_SCALES];    /* 0, or: the decoder's native map is [B,1,Hd,Wd] and is     */
  int32_t Wd[PSFM_MAX_SCALES];    /* nearest-upsampled to (Hs,Ws) INSIDE the kernels (what     */
                                  /* SfmModel.compute_depth_net -> upsample_output does before */
                                  /* the loss, models/SfmModel.py:86-88, model_utils.py:152-176);*/
                                  /* inv_depth[i] / d_inv_depth[i] are then [B,1,Hd,Wd]        */
  int32_t img_dtype;              /* psfm_dtype of images / mask                              */
  int32_t real_dtype;             /* PSFM_F32 or PSFM_F64                                     */
  int32_t padding;                /* psfm_padding  (padding_mode)                             */
  int32_t reduce_op;              /* psfm_reduce   (photometric_reduce_op)                    */
  int32_t automask;               /* automask_loss                                            */
  int32_t premap;                 /* 0: inv_depths are inverse depths (upstream / north star) */
                                  /* 1: fork's sigmoid pre-map (:362-369)                     */
  int32_t camera_model;           /* psfm_camera; FISHEYE: psfm_loss_forward_backward only    */
  double ssim_w;                  /* ssim_loss_weight                                         */
  double smooth_w;                /* smooth_loss_weight                                       */
  double C1, C2;                  /* SSIM constants                                           */
  double clip_loss;               /* clip_loss (0 disables)                                   */
  double min_depth, max_depth;    /* only used when premap == 1                               */
} psfm_loss_cfg;

/* Device pointers of one loss evaluation.  Per-scale entries beyond cfg->n and per-context
 * entries beyond cfg->C are ignored.  In "U mode" (all scales full resolution) image[i] and
 * ctx[i][j] are simply the same pointers for every i. */
typedef struct psfm_loss_io {
  const void* image[PSFM_MAX_SCALES];               /* [B,3,Hs,Ws] img_dtype: target at scale i     */
  const void* ctx[PSFM_MAX_SCALES][PSFM_MAX_CTX];   /* [B,3,Hs,Ws] img_dtype: context j at scale i  */
  const void* inv_depth[PSFM_MAX_SCALES];           /* [B,1,Hs,Ws] real                              */
  const void* mask[PSFM_MAX_SCALES];                /* [B,1,Hs,Ws] img_dtype or NULL                 */
  const void* K;                                    /* [B,3,3] real, full-resolution intrinsics      */
  const void* ref_K;                                /* [B,3,3] real ([B,11] each for PSFM_CAM_FISHEYE) */
  const void* pose[PSFM_MAX_CTX];                   /* [B,4,4] real, Pose.mat of context j           */
  uint8_t* sel[PSFM_MAX_SCALES];                    /* [B,Hs,Ws] u8: arg-min candidate (fwd out, bwd in) */
  void* d_inv_depth[PSFM_MAX_SCALES];               /* [B,1,Hs,Ws] real (bwd out, overwritten)       */
  void* d_pose[PSFM_MAX_CTX];                       /* [B,4,4] real (bwd out, overwritten; row 3 = 0) */
  uint32_t* dbg[PSFM_MAX_SCALES];                   /* NULL, or [C,B,Hs,Ws] u32 (test hook of psfm_loss_forward_backward,
                                                     * pinhole sweep kernels): the discrete decisions the loss took at
                                                     * each pixel -- bits 0-11 floor(ix) + 8, bits 12-22 floor(iy) + 8 of
                                                     * the bilinear tap of context c (after the padding mode's clipping),
                                                     * bits 23-28 sign of warped - target per channel (L1 term; two bits
                                                     * each: 0 negative, 1 zero, 2 positive), bits 29-31 SSIM distance of
                                                     * the channel inside its clamp range;
                                                     * io->sel[i] then receives the arg-min candidate as in
                                                     * psfm_loss_forward.  Lets the parity tests name the pixels whose
                                                     * tap or arg-min ACTUALLY differs from the fp64 oracle's
                                                     * (SURVEY.md section 8(c), T2). */
} psfm_loss_io;

int psfm_abi_version(void);
const char* psfm_last_error(void);

/* Measurement hooks (no reference counterpart; used by bench.py).  psfm_launch_count: kernels
 * this library has launched since load.  psfm_profile_enable(1) makes the loss entry points
 * bracket their kernels with CUDA events on the caller's stream; psfm_profile_read waits for
 * them, returns up to max_records (elapsed ms, tag) pairs and clears the list.
 * tags: 0 mean, 1 loss forward, 2 loss backward, 3 finish passes, 4 fused forward+backward. */
unsigned long long psfm_launch_count(void);
int psfm_profile_enable(int on);
int psfm_profile_read(float* ms, int* tags, int max_records);

/* Scratch bytes psfm_loss_forward/backward need for this configuration (upper bound: every scale is assumed to
 * have its own images). */
size_t psfm_loss_workspace_bytes(const psfm_loss_cfg* cfg);
/* The same for these pointers: scales that share their image / context / mask pointers (all of them in the shipped
 * upsample_depth_maps mode) share the image-sized scratch regions -- 0.11 GiB instead of 0.33 GiB at 8 x 192 x 640.
 * A workspace of this size is accepted by the loss entry points when they are called with the same cfg / io. */
size_t psfm_loss_workspace_bytes_io(const psfm_loss_cfg* cfg, const psfm_loss_io* io);

/* MultiViewPhotometricLoss.forward (losses/multiview_photometric_loss.py:331-410) with the
 * pinhole warp_ref_image (:131-195 upstream form): view synthesis for every (scale, context)
 * (geometry/camera_utils.py:27-59, geometry/camera.py:111-190, geometry/pose.py:80-86),
 * calc_photometric_loss (:218-267), automask identity terms (:394-399),
 * reduce_photometric_loss (:269-297) and calc_smoothness_loss (:301-327).
 * loss_out: device real[3] = {loss, photometric_loss, smoothness_loss}.
 * Also fills io->sel (arg-min maps) and workspace state that psfm_loss_backward consumes. */
int psfm_loss_forward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* workspace,
                      size_t workspace_bytes, void* loss_out, void* stream);

/* Backward of the above (replaces the autograd replay of SURVEY.md §3.3): recomputes the
 * warp, writes d loss / d inv_depths[i] and d loss / d pose[j].mat, both scaled by the
 * upstream gradient *grad_loss (device real[1]).  Must follow psfm_loss_forward with the
 * same cfg/io/workspace. */
int psfm_loss_backward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* workspace,
                       size_t workspace_bytes, const void* grad_loss, void* stream);

/* Single-pass variant: loss and the gradients for upstream gradient 1 in one sweep over the
 * inputs (no arg-min maps, no recompute).  Outputs as forward + backward(grad_loss = 1). */
int psfm_loss_forward_backward(const psfm_loss_cfg* cfg, const psfm_loss_io* io, void* workspace,
                               size_t workspace_bytes, void* loss_out, void* stream);

/* y[i] = x[i] * (*scale) for i < count; used to apply the upstream gradient to the outputs
 * of psfm_loss_forward_backward. */
int psfm_scale_inplace(void* x, const void* scale, size_t count, int real_dtype, void* stream);

/* view_synthesis (geometry/camera_utils.py:27-59): out[B,3,H,W] = grid_sample(ref_image,
 * project(reconstruct(depth))) with align_corners=True.  K / ref_K are the intrinsics of THIS
 * resolution (already scaled); pose = ref_cam.Tcw.mat [B,4,4].  depth is a depth map. */
int psfm_warp_forward(int B, int H, int W, int img_dtype, int real_dtype, int padding,
                      const void* ref_image, const void* depth, const void* K, const void* ref_K,
                      const void* pose, void* out, void* stream);

/* Backward of psfm_warp_forward w.r.t. depth [B,1,H,W] and pose [B,4,4] (and, when
 * d_ref_image != NULL, w.r.t. the reference image, accumulated with atomics into a
 * pre-zeroed buffer).  grad_out: [B,3,H,W] real.  workspace: psfm_warp_workspace_bytes(). */
size_t psfm_warp_workspace_bytes(int B, int H, int W, int real_dtype);
int psfm_warp_backward(int B, int H, int W, int img_dtype, int real_dtype, int padding,
                       const void* ref_image, const void* depth, const void* K, const void* ref_K,
                       const void* pose, const void* grad_out, void* d_depth, void* d_pose,
                       void* d_ref_image, void* workspace, size_t workspace_bytes, void* stream);

/* interpolate_image / interpolate_scales / match_scales (utils/image.py:117-214) and the model's upsample_output
 * (models/model_utils.py:152-193): resize of [N,Hi,Wi] planes to [N,Ho,Wo] with torch's F.interpolate arithmetic --
 * mode 0 bilinear align_corners=True, mode 1 nearest, mode 2 bilinear align_corners=False.
 * psfm_resize_backward is its adjoint (grad_out [N,Ho,Wo] -> grad_in [N,Hi,Wi], f32 / f64, gathered per source
 * pixel in a fixed order: deterministic). */
int psfm_resize(int N, int Hi, int Wi, int Ho, int Wo, int dtype, int mode, const void* in, void* out,
                void* stream);
int psfm_resize_backward(int N, int Hi, int Wi, int Ho, int Wo, int dtype, int mode, const void* grad_out,
                         void* grad_in, void* stream);

/* GenericCamera.project (geometry/camera_generic.py:86-208): the ray-surface camera "projects" a
 * point by a soft arg-max over the 41 x 41 patch of viewing rays around the pixel -- softmax, with temperature
 * `temperature` (max(1e-8, 1e-4 / exp(0.1 progress)), :107-110,:183-185), of the cosines between the point's direction
 * and the rays; the result is the weighted mean of the patch's (row, col) indices.  dir [B,3,h,w]: point directions
 * (un-normalised; the half-resolution, pose-transformed points of :172-180), rays [B,3,h,w]: the ray surface at the
 * same resolution; h, w >= 41.  Outputs: coords [B,h,w,2] = (row, col); dhat [B,3,h,w] and stats [B,h,w,2] are state
 * for the backward.  The reference is batch-1 (.squeeze()) and materialises the [h w, 1681, 3] patch tensor.
 * Backward: grad_coords [B,h,w,2] -> d_dir, d_rays [B,3,h,w]; d_rays is gathered per ray in a fixed order (no atomics). */
int psfm_raysurf_project(int B, int h, int w, int dt, double temperature, const void* dir, const void* rays, void* coords,
                         void* dhat, void* stats, void* stream);
int psfm_raysurf_project_backward(int B, int h, int w, int dt, double temperature, const void* dir, const void* rays,
                                  const void* coords, const void* dhat, const void* stats, const void* grad_coords,
                                  void* d_dir, void* d_rays, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Building-block operators: the reference functions of this path for callers that use them
 * directly (not through MultiViewPhotometricLoss.forward).  All tensors are contiguous and of the
 * arithmetic type `dt` (PSFM_F32 | PSFM_F64).  Scratch: psfm_block_workspace_bytes(B, N) bytes,
 * N = elements per batch item the reduction runs over.
 * ------------------------------------------------------------------------------------------- */
size_t psfm_block_workspace_bytes(int B, size_t N);

/* Camera.reconstruct (geometry/camera.py:111-147): out[B,3,H,W] = Twc (Kinv (u,v,1) * depth);
 * Twc = [B,4,4] world<-camera matrix or NULL for frame 'c' / an identity pose.
 * Backward: d_depth[B,1,H,W] from grad_out[B,3,H,W]. */
int psfm_reconstruct(int B, int H, int W, int dt, const void* depth, const void* K, const void* Twc, void* out,
                     void* stream);
int psfm_reconstruct_backward(int B, int H, int W, int dt, const void* K, const void* Twc, const void* grad_out,
                              void* d_depth, void* stream);

/* Camera.project (geometry/camera.py:149-190): out[B,H,W,2] = normalised coordinates of
 * K (Tcw X), Z clamped at 1e-5; Tcw = [B,4,4] or NULL.  Backward: dX[B,3,H,W] and, when Tcw and
 * d_Tcw are given, d_Tcw[B,4,4] (rows 0-2). */
int psfm_project(int B, int H, int W, int dt, const void* X, const void* K, const void* Tcw, void* out, void* stream);
int psfm_project_backward(int B, int H, int W, int dt, const void* X, const void* K, const void* Tcw,
                          const void* grad_out, void* dX, void* d_Tcw, void* workspace, size_t workspace_bytes,
                          void* stream);

/* FisheyeCamera.reconstruct / FisheyeCamera.project in the camera frame (geometry/camera.py:242-305, :311-394; VADAS
 * model; intr = [B,11] rows k0..k6, s, div, ux, uy).  The world-frame variants of the reference are these composed
 * with psfm_transform_points (as shipped they raise).
 * reconstruct: out[B,3,H,W] = (tan(r_d)/r_d x_d, tan(r_d)/r_d y_d, 1) * depth; backward: d_depth[B,1,H,W].
 * project: X[B,3,N] -> out[B,N,2] normalised with (Hn, Wn) (N = H*W, Hn = H, Wn = W for maps; image_size for clouds);
 * Z clamped at sys.float_info.epsilon; backward: dX[B,3,N]. */
int psfm_fisheye_reconstruct(int B, int H, int W, int dt, const void* depth, const void* intr, void* out, void* stream);
int psfm_fisheye_reconstruct_backward(int B, int H, int W, int dt, const void* intr, const void* grad_out, void* d_depth,
                                      void* stream);
int psfm_fisheye_project(int B, size_t N, int Hn, int Wn, int dt, const void* X, const void* intr, void* out, void* stream);
int psfm_fisheye_project_backward(int B, size_t N, int Hn, int Wn, int dt, const void* X, const void* intr,
                                  const void* grad_out, void* dX, void* stream);

/* funct.grid_sample(image, grid, mode='bilinear', padding_mode, align_corners=True) as view_synthesis calls it
 * (geometry/camera_utils.py:58-59): image[B,3,Hi,Wi], grid[B,Ho,Wo,2] -> out[B,3,Ho,Wo], all of type dt.
 * Backward: d_grid[B,Ho,Wo,2] (the image is an input frame; its gradient is not produced). */
int psfm_grid_sample(int B, int Hi, int Wi, int Ho, int Wo, int dt, int padding, const void* image, const void* grid,
                     void* out, void* stream);
int psfm_grid_sample_backward(int B, int Hi, int Wi, int Ho, int Wo, int dt, int padding, const void* image,
                              const void* grid, const void* grad_out, void* d_grid, void* stream);

/* SupervisedLoss.calculate_loss for one scale (losses/supervised_loss.py:292-417) with the loss
 * functions get_loss_func returns for '...l1' / '...mse' / '...silog' (:84-92, SilogLoss :67-82) and '...ssi-silog'
 * (SSISilogLoss, losses/ssi_silog_loss.py:53-429; its multi-scale gradient term: psfm_sobel_gradient_loss below).
 * sparse != 0: the 'sparse-*' methods -- valid = gt > 0 (and mask != 0 when a [N] byte mask is given); pred and gt are
 * filled with 1e-6 elsewhere and the mean still runs over all N elements, as the reference does; ssi-silog averages
 * over the valid elements only.  pred, gt, d_pred: [N] of type dt, 16-byte aligned.  hp: HOST double[6] = {silog ratio2, ssi alpha, ssi weight,
 * silog weight, depth clamp min, depth clamp max} (the last four: ssi-silog only).  loss_out: real[1].
 * state: DEVICE double[16], written by the forward call, read by the backward call and by callers that log parts:
 * {sum|d|, sum d^2, sum l, sum l^2, valid count, sum d, N, loss, ssi, silog, mean d, var d, silog1, silog2, sign, const}.
 * Backward: d_pred[N] = grad_loss[0] * d loss / d pred.
 * ('...berhu' raises inside the reference itself, torch.cat of a 4-D and a 1-D tensor at :65, and is not provided.) */
typedef enum psfm_sup_method { PSFM_SUP_L1 = 0, PSFM_SUP_MSE = 1, PSFM_SUP_SILOG = 2, PSFM_SUP_SSI_SILOG = 3 } psfm_sup_method;
int psfm_supervised_loss(int method, int sparse, size_t N, int dt, const void* pred, const void* gt, const void* mask,
                         const double* hp, void* loss_out, void* state, void* workspace, size_t workspace_bytes, void* stream);
int psfm_supervised_loss_backward(int method, int sparse, size_t N, int dt, const void* pred, const void* gt, const void* mask,
                                  const double* hp, const void* state, const void* grad_loss, void* d_pred, void* stream);

/* The multi-scale Sobel gradient term of SSISilogLoss (losses/ssi_silog_loss.py:12-50 Gradient2D, :116-174
 * compute_gradient_loss; `gradient_weight` / `gradient_scales` of the loss section, set by
 * configs/train_resnet_san_ncdb_dual_head_640x384.yaml).  pred_inv, gt_inv: [B,1,H,W] inverse depths
 * of type dt; valid = gt_inv > 0 (and mask != 0 when a [B*H*W] byte mask is given), as the sparse caller builds it
 * (losses/supervised_loss.py:314-333).  Per scale s < scales: depth maps resized by 0.5^s (bilinear, align_corners=False),
 * mask by nearest, mean |Sobel(pred) - Sobel(gt)| (x plus y) over the masked interior; average over the scales that have a
 * masked pixel (0 when none).  loss_out: real[1].  state: DEVICE double[20] = {loss, scales used, 0, 0, count per scale [8],
 * loss per scale [8]}.  codes: DEVICE bytes (psfm_sobel_gradient_code_bytes: one sign code per pyramid pixel, then the eight
 * per-scale coefficients of the backward as doubles), written by the forward call and read by the backward call.  Backward: d_pred[B,1,H,W] = grad_loss[0] * d loss / d pred_inv. */
size_t psfm_sobel_gradient_code_bytes(int B, int H, int W, int scales);
size_t psfm_sobel_gradient_workspace_bytes(int B, int H, int W, int scales);
int psfm_sobel_gradient_loss(int B, int H, int W, int dt, int scales, const void* pred_inv, const void* gt_inv,
                             const void* mask, void* loss_out, void* state, void* codes, void* workspace,
                             size_t workspace_bytes, void* stream);
int psfm_sobel_gradient_loss_backward(int B, int H, int W, int dt, int scales, const void* pred_inv, const void* gt_inv,
                                      const void* mask, const void* state, const void* codes, const void* grad_loss,
                                      void* d_pred, void* stream);

/* DualHeadDepthLoss.forward (losses/dual_head_depth_loss.py:23-201 with decompose_depth / dual_head_to_depth,
 * networks/layers/resnet/layers.py:76-149): integer_pred, fractional_pred (sigmoid outputs) and
 * depth_gt are [N] of type dt (depth_gt already at the prediction's size).  hp: HOST double[5] = {max_depth, min_depth,
 * integer_weight, fractional_weight, consistency_weight}.  loss_out: real[4] = {loss, integer_loss, fractional_loss,
 * consistency_loss}; state: DEVICE double[8] = {sum_i, sum_f, sum_c, valid count, #(integer error < 1 m), sum frac_err^2,
 * 0, 1 if no pixel is valid}.  Backward: d_integer[N], d_fractional[N] = grad_loss[0] * d loss / d (.). */
int psfm_dual_head_loss(size_t N, int dt, const void* integer_pred, const void* fractional_pred, const void* depth_gt,
                        const double* hp, void* loss_out, void* state, void* workspace, size_t workspace_bytes, void* stream);
int psfm_dual_head_loss_backward(size_t N, int dt, const void* integer_pred, const void* fractional_pred, const void* depth_gt,
                                 const double* hp, const void* state, const void* grad_loss, void* d_integer, void* d_fractional,
                                 void* stream);

/* compute_depth_metrics(config, gt, pred, use_gt_scale) (utils/depth.py:258-447 with scale_depth :450-481; SURVEY.md
 * section 8(f) N4): gt[B,1,H,W], pred[B,1,h,w] of type dt.  scale_mode: 0 = config.scale_output 'resize' (bilinear,
 * align_corners=True, evaluated per pixel; the prediction itself when the sizes agree), 1 = 'top-center' (prediction placed
 * at the bottom, centred, zeros elsewhere); anything else is PSFM_ERR_INVALID like the reference's NotImplementedError.
 * crop != 0: config.crop == 'garg'.  valid = min_depth < gt < max_depth (and inside the crop).  use_gt_scale != 0: the
 * prediction of every image is multiplied by median(gt valid) / median(pred valid), torch.median's lower median, found by
 * exact radix selection on the device.  out: real[7] = (abs_rel, sq_rel, rmse, rmse_log, a1, a2, a3), each the sum over the
 * images that have a valid pixel divided by B.  state: DEVICE double[B * 12], per image {valid count, gt median, pred
 * median, scale, the seven metrics of the image, 0}.  No host synchronisation (the reference reads valid.sum() back). */
size_t psfm_depth_metrics_workspace_bytes(int B, int H, int W);
int psfm_depth_metrics(int B, int H, int W, int h, int w, int dt, int scale_mode, int crop, int use_gt_scale,
                       double min_depth, double max_depth, const void* gt, const void* pred, void* out, void* state,
                       void* workspace, size_t workspace_bytes, void* stream);

/* Pose.transform_points (geometry/pose.py:80-86): out[B,3,N] = R X + t.  Backward: dX, d_mat[B,4,4]. */
int psfm_transform_points(int B, size_t N, int dt, const void* mat, const void* X, void* out, void* stream);
int psfm_transform_points_backward(int B, size_t N, int dt, const void* mat, const void* X, const void* grad_out,
                                   void* dX, void* d_mat, void* workspace, size_t workspace_bytes, void* stream);

/* Pose.from_vec -> pose_vec2mat -> euler2mat (geometry/pose.py:39-46, geometry/pose_utils.py:8-51):
 * vec[B,6] = [tx, ty, tz, rx, ry, rz] -> mat[B,4,4] = [Rx(rx) Ry(ry) Rz(rz) | t; 0 0 0 1].
 * Backward: d_vec[B,6] from grad_mat[B,4,4] (rows 0-2 are read). */
int psfm_pose_from_vec(int B, int dt, const void* vec, void* mat, void* stream);
int psfm_pose_from_vec_backward(int B, int dt, const void* vec, const void* grad_mat, void* d_vec, void* stream);

/* SSIM / calc_photometric_loss (losses/multiview_photometric_loss.py:15-54,199-267) of x, y [B,3,H,W]:
 *   mode 0 SSIM index [B,3,H,W]; mode 1 clamp((1-SSIM)/2,0,1) [B,3,H,W];
 *   mode 2 ssim_w * mean_c(mode 1) + (1-ssim_w) * mean_c|x-y| [B,1,H,W]; mode 3 |x-y| [B,3,H,W].
 * Backward: dx (and dy when not NULL) from grad_out of the output's shape. */
int psfm_photometric_map(int B, int H, int W, int dt, int mode, double ssim_w, double C1, double C2, const void* x,
                         const void* y, void* out, void* stream);
int psfm_photometric_map_backward(int B, int H, int W, int dt, int mode, double ssim_w, double C1, double C2,
                                  const void* x, const void* y, const void* grad_out, void* dx, void* dy, void* stream);

/* inv_depths_normalize (utils/depth.py:146-162): out = d / max(mean over the N elements of each
 * batch item, 1e-6); mean_out[B] is kept for the backward. */
int psfm_normalize(int B, size_t N, int dt, const void* d, void* out, void* mean_out, void* workspace,
                   size_t workspace_bytes, void* stream);
int psfm_normalize_backward(int B, size_t N, int dt, const void* d, const void* mean, const void* grad_out, void* dd,
                            void* workspace, size_t workspace_bytes, void* stream);

/* calc_smoothness maps (utils/depth.py:165-198, utils/image.py:85-113): sx[B,1,H,W-1], sy[B,1,H-1,W]
 * from normalised inverse depths dn[B,1,H,W] and image[B,3,H,W].  Backward w.r.t. dn. */
int psfm_smoothness_maps(int B, int H, int W, int dt, const void* dn, const void* image, void* sx, void* sy,
                         void* stream);
int psfm_smoothness_maps_backward(int B, int H, int W, int dt, const void* image, const void* gsx, const void* gsy,
                                  void* ddn, void* stream);

/* reduce_photometric_loss for one scale (losses/multiview_photometric_loss.py:269-297): K candidates
 * (host array of device pointers), each [B,Cg,H,W]; op PSFM_REDUCE_MIN | PSFM_REDUCE_MEAN; out real[1];
 * argmin int32[B,H,W] (min only).  Backward writes d_cands[k] for upstream grad_out real[1]. */
int psfm_reduce_candidates(int K, int B, int Cg, int H, int W, int dt, int op, const void* const* cands, void* out,
                           int32_t* argmin, void* workspace, size_t workspace_bytes, void* stream);
int psfm_reduce_candidates_backward(int K, int B, int Cg, int H, int W, int dt, int op, const int32_t* argmin,
                                    const void* grad_out, void* const* d_cands, void* stream);

/* mean |x| over n elements (calc_smoothness_loss, :320-322) and its backward. */
int psfm_abs_mean(size_t n, int dt, const void* x, void* out, void* workspace, size_t workspace_bytes, void* stream);
int psfm_abs_mean_backward(size_t n, int dt, const void* x, const void* grad_out, void* dx, void* stream);

#ifdef __cplusplus
}
#endif
#endif  /* PSFM_H_ */
