"""Pinhole camera (reference: geometry/camera.py:15-190) and the fork's fisheye (VADAS) camera
(reference :194-394).

`reconstruct` and `project` keep the reference signatures and run as CUDA kernels through
the C ABI."""
import torch
import torch.nn as nn

from packnet_sfm.geometry.pose import Pose
from packnet_sfm.geometry.camera_utils import scale_intrinsics


class Camera(nn.Module):
    """Differentiable pinhole camera: intrinsics K [B,3,3] and camera->world pose Tcw."""

    def __init__(self, K, Tcw=None):
        super().__init__()
        self.K = K
        self._identity_pose = Tcw is None
        self.Tcw = Pose.identity(len(K), device=K.device if self._identity_pose else None) if Tcw is None else Tcw
        self._Twc = None
        self._Kinv = None

    def __len__(self):
        return len(self.K)

    def to(self, *args, **kwargs):
        self.K = self.K.to(*args, **kwargs)
        self.Tcw = self.Tcw.to(*args, **kwargs)
        self._Twc = self._Kinv = None
        return self

    @property
    def fx(self):
        return self.K[:, 0, 0]

    @property
    def fy(self):
        return self.K[:, 1, 1]

    @property
    def cx(self):
        return self.K[:, 0, 2]

    @property
    def cy(self):
        return self.K[:, 1, 2]

    @property
    def Twc(self):
        """World -> camera pose (inverse of Tcw); cached like the reference's lru_cache (:66-69)."""
        if self._Twc is None:
            self._Twc = self.Tcw.inverse()
        return self._Twc

    @property
    def Kinv(self):
        """The reference's lifting matrix (:73-81): K with fx, fy, cx, cy entries replaced by
        1/fx, 1/fy, -cx/fx, -cy/fy; the remaining entries are copied, not inverted."""
        if self._Kinv is None:
            Kinv = self.K.clone()
            Kinv[:, 0, 0] = 1. / self.fx
            Kinv[:, 1, 1] = 1. / self.fy
            Kinv[:, 0, 2] = -1. * self.cx / self.fx
            Kinv[:, 1, 2] = -1. * self.cy / self.fy
            self._Kinv = Kinv
        return self._Kinv

    def scaled(self, x_scale, y_scale=None):
        """Camera of the resized image (:84-108); one argument scales both axes by it."""
        if y_scale is None:
            y_scale = x_scale
        if x_scale == 1. and y_scale == 1.:
            return self
        cam = Camera(scale_intrinsics(self.K.clone(), x_scale, y_scale), Tcw=self.Tcw)
        cam._identity_pose = self._identity_pose
        return cam

    def reconstruct(self, depth, frame='w'):
        """Depth map [B,1,H,W] -> 3D points [B,3,H,W] in the camera ('c') or world ('w') frame (:111-147)."""
        B, C, H, W = depth.shape
        assert C == 1
        from packnet_sfm.ops import reconstruct_op
        if frame == 'c':
            return reconstruct_op(depth, self.K, None)
        elif frame == 'w':
            return reconstruct_op(depth, self.K, None if self._identity_pose else self.Twc.mat)
        else:
            raise ValueError('Unknown reference frame {}'.format(frame))

    def project(self, X, frame='w'):
        """3D points [B,3,H,W] -> normalised image coordinates [B,H,W,2] (:149-190)."""
        B, C, H, W = X.shape
        assert C == 3
        from packnet_sfm.ops import project_op
        if frame == 'c':
            return project_op(X, self.K, None)
        elif frame == 'w':
            return project_op(X, self.K, None if self._identity_pose else self.Tcw.mat)
        else:
            raise ValueError('Unknown reference frame {}'.format(frame))


class FisheyeCamera(nn.Module):
    """Differentiable fisheye camera (VADAS model), reference :194-394.

    intrinsics: dict with 'k' [B,7] (distortion polynomial), 's', 'div', 'ux', 'uy' [B]
    (datasets/ncdb_dataset.py:577-584).  `reconstruct` keeps the reference's placeholder inverse of the
    polynomial (theta = r_d, :278-283).  The world-frame branches, which raise in the reference because
    `Pose.transform_points` is handed a [B,3,HW] tensor (:305, :344), apply the pose to the point tensor here."""

    def __init__(self, intrinsics, Tcw=None, image_size=None):
        super().__init__()
        self.k = intrinsics['k']
        self.s = intrinsics['s']
        self.div = intrinsics['div']
        self.ux = intrinsics['ux']
        self.uy = intrinsics['uy']
        self._identity_pose = Tcw is None
        self.Tcw = Pose.identity(len(self.k), device=self.k.device, dtype=self.k.dtype) if Tcw is None else Tcw
        self.image_size = image_size
        self._Twc = None

    def __len__(self):
        return len(self.k)

    def to(self, *args, **kwargs):
        for name in ('k', 's', 'div', 'ux', 'uy'):
            setattr(self, name, getattr(self, name).to(*args, **kwargs))
        self.Tcw = self.Tcw.to(*args, **kwargs)
        self._Twc = None
        return self

    @property
    def intrinsics(self):
        return {'k': self.k, 's': self.s, 'div': self.div, 'ux': self.ux, 'uy': self.uy}

    @property
    def Twc(self):
        """World -> camera pose (inverse of Tcw)."""
        if self._Twc is None:
            self._Twc = self.Tcw.inverse()
        return self._Twc

    def _packed(self, like):
        from packnet_sfm.ops import pack_fisheye
        return pack_fisheye(self.intrinsics, like.shape[0], like.dtype, like.device)

    def reconstruct(self, depth, frame='w'):
        """Depth map [B,1,H,W] -> 3D points [B,3,H,W] in the camera ('c') or world ('w') frame (:242-309)."""
        B, C, H, W = depth.shape
        assert C == 1
        from packnet_sfm.block_ops import fisheye_reconstruct_op
        Xc = fisheye_reconstruct_op(depth, self._packed(depth))
        if frame == 'c':
            return Xc
        elif frame == 'w':
            return Xc if self._identity_pose else self.Twc @ Xc
        else:
            raise ValueError('Unknown reference frame {}'.format(frame))

    def project(self, X, frame='w'):
        """3D points [B,3,H,W] (or [B,3,N] with `image_size`) -> normalised coordinates [B,H,W,2] ([B,N,2]) (:311-394)."""
        if X.dim() not in (3, 4):
            raise ValueError('Input X must be of shape [B,3,H,W] or [B,3,N]')
        assert X.shape[1] == 3
        from packnet_sfm.block_ops import fisheye_project_op
        if frame == 'w':
            Xc = X if self._identity_pose else self.Tcw @ X
        elif frame == 'c':
            Xc = X
        else:
            raise ValueError('Unknown reference frame {}'.format(frame))
        return fisheye_project_op(Xc, self._packed(X), self.image_size)
