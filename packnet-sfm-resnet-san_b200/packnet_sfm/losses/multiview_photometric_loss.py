"""Self-supervised multi-view photometric loss (reference: losses/multiview_photometric_loss.py).

`MultiViewPhotometricLoss.forward` is ONE autograd node over the C ABI: view synthesis for
every (scale, context), SSIM + L1, auto-masking identity terms, per-pixel min / mean, edge-aware
smoothness and the deterministic mean all run in sm_100a kernels (csrc/loss_kernels.cu); the
backward recomputes the warp.  The building-block methods (`warp_ref_image`, `SSIM`,
`calc_photometric_loss`, `reduce_photometric_loss`, `calc_smoothness_loss`) keep the reference
signatures and are kernel-backed individually for callers that use them directly.
"""
import torch

from packnet_sfm.geometry.camera import Camera, FisheyeCamera
from packnet_sfm.geometry.camera_utils import view_synthesis
from packnet_sfm.geometry.pose import Pose
from packnet_sfm.losses.loss_base import LossBase, ProgressiveScaling
from packnet_sfm.utils.depth import calc_smoothness, inv2depth
from packnet_sfm.utils.image import match_scales
from packnet_sfm import ops


def SSIM(x, y, C1=1e-4, C2=9e-4, kernel_size=3, stride=1):
    """3x3 structural similarity index of two [B,3,H,W] images with reflection padding
    (reference :15-54).  Returns the SSIM map (not the distance)."""
    if kernel_size != 3 or stride != 1:
        raise NotImplementedError('the SSIM kernel is specialised for kernel_size=3, stride=1')
    return ops.ssim_op(x, y, C1, C2)


class MultiViewPhotometricLoss(LossBase):
    """Drop-in for the reference class (:58-410); same constructor arguments and defaults."""

    def __init__(self, num_scales=4, ssim_loss_weight=0.85, occ_reg_weight=0.1, smooth_loss_weight=0.1,
                 C1=1e-4, C2=9e-4, photometric_reduce_op='mean', disp_norm=True, clip_loss=0.5,
                 progressive_scaling=0.0, padding_mode='zeros',
                 automask_loss=False, min_depth=0.05, max_depth=80.0, **kwargs):
        super().__init__()
        self.n = num_scales
        self.ssim_loss_weight = ssim_loss_weight
        self.occ_reg_weight = occ_reg_weight
        self.smooth_loss_weight = smooth_loss_weight
        self.C1 = C1
        self.C2 = C2
        self.photometric_reduce_op = photometric_reduce_op
        self.disp_norm = disp_norm
        self.clip_loss = clip_loss
        self.padding_mode = padding_mode
        self.automask_loss = automask_loss
        self.min_depth = min_depth
        self.max_depth = max_depth
        self.progressive_scaling = ProgressiveScaling(progressive_scaling, self.n)
        # B200-build options (not in the reference): the fork's sigmoid pre-map (:362-369) is opt-in,
        # the default takes true inverse depths (upstream / north-star semantics)
        self.sigmoid_premap = bool(kwargs.get('sigmoid_premap', False))
        self.kernel_mode = kwargs.get('kernel_mode', None)      # None -> ops.DEFAULT_MODE
        # cfg.model.loss.upsample_depth_maps (configs/default_config.py:24): in the reference the MODEL
        # nearest-upsamples every scale to full resolution before this loss sees it
        # (models/SfmModel.py:86-88 -> models/model_utils.py:152-176).  With this flag the loss takes the
        # decoder's native pyramid and does that upsampling (and its adjoint) inside the kernels.
        self.upsample_depth_maps = bool(kwargs.get('upsample_depth_maps', False))
        self.debug_decisions = bool(kwargs.get('debug_decisions', False))      # test hook: ops.LAST_DECISIONS
        if self.automask_loss:
            assert self.photometric_reduce_op == 'min', \
                'For automasking only the min photometric_reduce_op is supported.'

    @property
    def logs(self):
        return {'num_scales': self.n}

    def _hyper(self):
        return dict(ssim_loss_weight=self.ssim_loss_weight, smooth_loss_weight=self.smooth_loss_weight,
                    C1=self.C1, C2=self.C2, photometric_reduce_op=self.photometric_reduce_op,
                    clip_loss=self.clip_loss, padding_mode=self.padding_mode, automask_loss=self.automask_loss,
                    premap=int(self.sigmoid_premap), min_depth=self.min_depth, max_depth=self.max_depth,
                    upsample_depth_maps=self.upsample_depth_maps, debug_decisions=self.debug_decisions)

    def warp_ref_image(self, inv_depths, ref_image, K, ref_K, pose, image_size=None):
        """Warp `ref_image` into the target view at every scale (reference :131-195).  Returns a list of
        [B,3,Hs,Ws].  K / ref_K: [B,3,3] pinhole matrices (upstream body), or the fork's fisheye intrinsics
        dicts (the fork's body: FisheyeCamera per scale, principal point rescaled, k / s / div passed through)."""
        B, _, H, W = ref_image.shape
        ref_images = match_scales(ref_image, inv_depths, self.n)
        out = []
        if isinstance(K, dict):
            for i in range(self.n):
                _, _, DH, DW = inv_depths[i].shape
                sw, sh = DW / float(W), DH / float(H)
                scaled = lambda d: {'k': d['k'], 's': d['s'], 'div': d['div'],
                                    'ux': (d['ux'] + 0.5) * sw - 0.5, 'uy': (d['uy'] + 0.5) * sh - 0.5}
                cam = FisheyeCamera(intrinsics=scaled(K), image_size=(DH, DW))
                ref_cam = FisheyeCamera(intrinsics=scaled(ref_K), Tcw=pose, image_size=(DH, DW))
                out.append(view_synthesis(ref_images[i], inv2depth(inv_depths[i]), ref_cam, cam,
                                          padding_mode=self.padding_mode))
            return out
        for i in range(self.n):
            sf = inv_depths[i].shape[3] / float(W)
            cam = Camera(K=K).scaled(sf)
            ref_cam = Camera(K=ref_K, Tcw=pose).scaled(sf)
            out.append(view_synthesis(ref_images[i], inv2depth(inv_depths[i]), ref_cam, cam,
                                      padding_mode=self.padding_mode))
        return out

    def SSIM(self, x, y, kernel_size=3):
        """SSIM distance clamp((1 - SSIM) / 2, 0, 1) (reference :199-216)."""
        if kernel_size != 3:
            raise NotImplementedError('the SSIM kernel is specialised for kernel_size=3')
        return ops.ssim_op(x, y, self.C1, self.C2, distance=True)

    def calc_photometric_loss(self, t_est, images, masks_scaled=None):
        """Per-scale photometric maps alpha * SSIM + (1 - alpha) * L1 (reference :218-267)."""
        return [ops.photometric_map_op(t_est[i], images[i], self.ssim_loss_weight, self.C1, self.C2, self.clip_loss,
                                       None if masks_scaled is None else masks_scaled[i])
                for i in range(self.n)]

    def reduce_photometric_loss(self, photometric_losses):
        """min / mean over the candidates, mean over pixels, average over scales (reference :269-297)."""
        if self.photometric_reduce_op not in ('min', 'mean'):
            raise NotImplementedError('Unknown photometric_reduce_op: {}'.format(self.photometric_reduce_op))
        photometric_loss = sum(ops.reduce_candidates_op(photometric_losses[i], self.photometric_reduce_op)
                               for i in range(self.n)) / self.n
        self.add_metric('photometric_loss', photometric_loss)
        return photometric_loss

    def calc_smoothness_loss(self, inv_depths, images):
        """Edge-aware smoothness of the mean-normalised inverse depths (reference :301-327)."""
        smoothness_x, smoothness_y = calc_smoothness(inv_depths, images, self.n)
        smoothness_loss = sum([(ops.abs_mean_op(smoothness_x[i]) + ops.abs_mean_op(smoothness_y[i])) / 2 ** i
                               for i in range(self.n)]) / self.n
        smoothness_loss = self.smooth_loss_weight * smoothness_loss
        self.add_metric('smoothness_loss', smoothness_loss)
        return smoothness_loss

    def forward(self, image, context, inv_depths, K=None, ref_K=None, poses=None, return_logs=False, progress=0.0,
                mask=None, intrinsics=None, ref_intrinsics=None):
        """Photometric + smoothness loss (reference :331-410).

        image [B,3,H,W]; context: list of [B,3,H,W]; inv_depths: list of [B,1,Hs,Ws]; K, ref_K
        [B,3,3] (the fork's keyword names `intrinsics` / `ref_intrinsics` are accepted), or the fork's fisheye
        intrinsics dicts {'k': [B,7], 's', 'div', 'ux', 'uy': [B]} (reference :131-195 with
        geometry/camera.py:194-394; the pose is applied to the [B,3,H,W] points, which the reference's
        `frame='w'` branches fail to do);
        poses: list of Pose (target -> context).  Returns {'loss': [1], 'metrics': {...}}.
        """
        K = intrinsics if K is None else K
        ref_K = ref_intrinsics if ref_K is None else ref_K
        self.n = self.progressive_scaling(progress)
        mats = [p.mat if isinstance(p, Pose) else p for p in poses]
        out = ops.photometric_loss(self._hyper(), image, list(context), list(inv_depths[:self.n]), K, ref_K, mats,
                                   mask=mask, mode=self.kernel_mode)
        self.add_metric('photometric_loss', out[1])
        if self.smooth_loss_weight > 0.0:
            self.add_metric('smoothness_loss', out[2])
        return {'loss': out[0], 'metrics': self.metrics}
