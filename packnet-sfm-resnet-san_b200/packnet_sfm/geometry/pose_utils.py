"""Pose-vector helpers (reference: geometry/pose_utils.py).  These act on [B,6] / [B,4,4]
tensors -- O(B) work that stays in PyTorch so autograd carries d pose.mat (produced by the
loss kernels) back to the pose network's 6-DoF output."""
import numpy as np
import torch


def euler2mat(angle):
    """[B,3] Euler angles (x, y, z) -> [B,3,3] rotation, R = Rx @ Ry @ Rz (reference :8-37)."""
    x, y, z = angle[:, 0], angle[:, 1], angle[:, 2]
    one, zero = torch.ones_like(x), torch.zeros_like(x)
    cx, sx, cy, sy, cz, sz = torch.cos(x), torch.sin(x), torch.cos(y), torch.sin(y), torch.cos(z), torch.sin(z)
    rz = torch.stack([cz, -sz, zero, sz, cz, zero, zero, zero, one], dim=1).view(-1, 3, 3)
    ry = torch.stack([cy, zero, sy, zero, one, zero, -sy, zero, cy], dim=1).view(-1, 3, 3)
    rx = torch.stack([one, zero, zero, zero, cx, -sx, zero, sx, cx], dim=1).view(-1, 3, 3)
    return rx.bmm(ry).bmm(rz)


def pose_vec2mat(vec, mode='euler'):
    """[B,6] = [t | euler] -> [B,3,4] (reference :41-51)."""
    if mode is None:
        return vec
    if mode != 'euler':
        raise ValueError('Rotation mode not supported {}'.format(mode))
    return torch.cat([euler2mat(vec[:, 3:]), vec[:, :3].unsqueeze(-1)], dim=2)


def invert_pose(T):
    """Rigid inverse of [B,4,4] (reference :55-60)."""
    Rt = T[:, :3, :3].transpose(-2, -1)
    Tinv = torch.eye(4, device=T.device, dtype=T.dtype).repeat([len(T), 1, 1])
    Tinv[:, :3, :3] = Rt
    Tinv[:, :3, -1] = torch.bmm(-1. * Rt, T[:, :3, -1].unsqueeze(-1)).squeeze(-1)
    return Tinv


def invert_pose_numpy(T):
    """Rigid inverse of a [4,4] numpy pose (reference :64-69)."""
    Tinv = np.copy(T)
    R, t = Tinv[:3, :3], Tinv[:3, 3]
    Tinv[:3, :3], Tinv[:3, 3] = R.T, -np.matmul(R.T, t)
    return Tinv
