"""Generic (ray-surface) camera (reference: geometry/camera_generic.py:17-208).

`project` -- a soft arg-max over a 41 x 41 patch of viewing rays per pixel -- runs as CUDA kernels
(csrc/raysurf_kernels.cu) with a hand-written backward w.r.t. the points and the ray surface; the bilinear
half-resolution / full-resolution resizes around it are the library's differentiable resize kernels."""
import numpy as np
import torch

from packnet_sfm.geometry.pose import Pose


class GenericCamera(torch.nn.Module):
    """Drop-in for the reference class: `R` is the camera's ray surface [B,3,H,W], `Tcw` the camera -> world pose.
    Works for any batch size (the reference squeezes a batch of one) and does not assume a CUDA default device."""

    def __init__(self, R, Tcw=None):
        super().__init__()
        self.ray_surface = R
        self.Tcw = Pose.identity(1, device=R.device, dtype=R.dtype) if Tcw is None else Tcw
        self._identity_pose = Tcw is None

    def to(self, *args, **kwargs):
        self.ray_surface = self.ray_surface.to(*args, **kwargs)
        self.Tcw = self.Tcw.to(*args, **kwargs)
        return self

    @property
    def Twc(self):
        """World -> camera pose (inverse of Tcw)."""
        return self.Tcw.inverse()

    def reconstruct(self, depth, frame='w'):
        """P(x, y) = d(x, y) r(x, y) (reference :54-83); depth [B,1,H,W] -> points [B,3,H,W]."""
        B, C, H, W = depth.shape
        assert C == 1
        Xc = self.ray_surface * depth
        if frame == 'c':
            return Xc
        if frame == 'w':
            return Xc if self._identity_pose else self.Twc @ Xc
        raise ValueError('Unknown reference frame {}'.format(frame))

    def project(self, X, progress, downsample=True, frame='c'):
        """Soft arg-max projection (reference :86-208): [B,3,H,W] points -> normalised sampling grid [B,H,W,2]."""
        from packnet_sfm.block_ops import raysurf_project_op
        from packnet_sfm.ops import resize_op
        B, C, H, W = X.shape
        assert C == 3
        if frame == 'w':
            X = self.Tcw @ X
        rays = self.ray_surface.expand(B, 3, H, W) if self.ray_surface.shape[0] != B else self.ray_surface
        direction = X
        if downsample:
            H, W = int(H / 2.), int(W / 2.)
            rays = resize_op(rays, (H, W), 'bilinear')                # F.interpolate(scale_factor=0.5, align_corners=True)
            direction = resize_op(direction, (H, W), 'bilinear')
        temperature = np.maximum(1e-8, 0.0001 / np.exp(0.1 * progress))
        coords = raysurf_project_op(direction, rays, temperature)     # [B,h,w,2] = (row, col)
        xn = 2 * coords[..., 0] / (H - 1) - 1.
        yn = 2 * coords[..., 1] / (W - 1) - 1.
        if downsample:
            xn = resize_op(xn.unsqueeze(1), (2 * H, 2 * W), 'bilinear').squeeze(1)
            yn = resize_op(yn.unsqueeze(1), (2 * H, 2 * W), 'bilinear').squeeze(1)
        return torch.stack([yn, xn], dim=-1)
