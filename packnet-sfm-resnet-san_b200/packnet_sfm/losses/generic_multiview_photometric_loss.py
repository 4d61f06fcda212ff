"""Ray-surface (generic camera) multi-view photometric loss (reference: losses/generic_multiview_photometric_loss.py:92-400;
SURVEY.md 8(f) N3).

Same SSIM / L1 / automask / min-or-mean / smoothness tail as `MultiViewPhotometricLoss` (its kernel-backed building-block
methods are inherited); the view synthesis goes through `GenericCamera` -- ray surface = canonical template + annealed
residual, soft arg-max projection (csrc/raysurf_kernels.cu).  Gradients reach the inverse depths, the poses and the
ray-surface residual."""
import os

import numpy as np
import torch

from packnet_sfm.geometry.camera_generic import GenericCamera
from packnet_sfm.geometry.camera_utils import view_synthesis_generic
from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
from packnet_sfm.utils.depth import inv2depth
from packnet_sfm.utils.image import match_scales


class GenericMultiViewPhotometricLoss(MultiViewPhotometricLoss):
    """Drop-in for the reference class (same constructor arguments and defaults, :130-155).

    The reference loads its initial ray table from "omnicam_ray_template.npy" in the working directory (:146-148),
    a file it does not ship; pass `canonical_ray_surface=` ([1,3,H,W] tensor or array) or `ray_template=` (a path)."""

    def __init__(self, num_scales=1, ssim_loss_weight=0.85, occ_reg_weight=0.1, smooth_loss_weight=0.1,
                 C1=1e-4, C2=9e-4, photometric_reduce_op='mean', disp_norm=True, clip_loss=0.5,
                 progressive_scaling=0.0, padding_mode='zeros', automask_loss=False, **kwargs):
        template = kwargs.pop('canonical_ray_surface', None)
        path = kwargs.pop('ray_template', 'omnicam_ray_template.npy')
        super().__init__(num_scales=num_scales, ssim_loss_weight=ssim_loss_weight, occ_reg_weight=occ_reg_weight,
                         smooth_loss_weight=smooth_loss_weight, C1=C1, C2=C2, photometric_reduce_op=photometric_reduce_op,
                         disp_norm=disp_norm, clip_loss=clip_loss, progressive_scaling=progressive_scaling,
                         padding_mode=padding_mode, automask_loss=automask_loss, **kwargs)
        if template is None:
            if not os.path.exists(path):
                raise FileNotFoundError('ray template %s not found: pass canonical_ray_surface= or ray_template=' % path)
            template = np.load(path)
        self.canonical_ray_surface = torch.as_tensor(template)

    def ray_surface(self, raysurf_residual, progress):
        """Template + annealed residual, unit length per pixel (reference :193-195)."""
        coeff = float(np.min([((100.0 * progress) ** (4 / 3.) / 100.), 1.]))
        R = self.canonical_ray_surface.to(device=raysurf_residual.device, dtype=raysurf_residual.dtype) + coeff * raysurf_residual
        return R / torch.norm(R, dim=1, keepdim=True)

    def warp_ref_image(self, inv_depths, ref_image, raysurf_residual, pose, progress):
        """Warped reference image per scale (reference :163-212)."""
        Rmat = self.ray_surface(raysurf_residual, progress)
        ref_images = match_scales(ref_image, inv_depths, self.n)
        out = []
        for i in range(self.n):
            cam, ref_cam = GenericCamera(R=Rmat), GenericCamera(R=Rmat, Tcw=pose)
            out.append(view_synthesis_generic(ref_images[i], inv2depth(inv_depths[i]), ref_cam, cam,
                                              padding_mode=self.padding_mode, progress=progress))
        return out

    def forward(self, image, context, inv_depths, ray_surface, K=None, ref_K=None, poses=None, return_logs=False,
                progress=0.0):
        """Photometric + smoothness loss (reference :337-400).  ray_surface: {('raysurf', 0): residual [B,3,H,W]}."""
        self.n = self.progressive_scaling(progress)
        photometric_losses = [[] for _ in range(self.n)]
        images = match_scales(image, inv_depths, self.n)
        residual = ray_surface[('raysurf', 0)] if isinstance(ray_surface, dict) else ray_surface
        for ref_image, pose in zip(context, poses):
            ref_warped = self.warp_ref_image(inv_depths, ref_image, residual, pose, progress=progress)
            photometric_loss = self.calc_photometric_loss(ref_warped, images)
            for i in range(self.n):
                photometric_losses[i].append(photometric_loss[i])
            if self.automask_loss:
                ref_images = match_scales(ref_image, inv_depths, self.n)
                unwarped_image_loss = self.calc_photometric_loss(ref_images, images)
                for i in range(self.n):
                    photometric_losses[i].append(unwarped_image_loss[i])
        loss = self.reduce_photometric_loss(photometric_losses)
        if self.smooth_loss_weight > 0.0:
            loss = loss + self.calc_smoothness_loss(inv_depths, images)
        return {'loss': loss.unsqueeze(0), 'metrics': self.metrics}
