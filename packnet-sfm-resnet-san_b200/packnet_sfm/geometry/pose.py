"""Pose: a batch of [4,4] rigid transforms (reference: geometry/pose.py)."""
import torch

from packnet_sfm.geometry.pose_utils import invert_pose, pose_vec2mat


class Pose:
    """Wraps `mat` [B,4,4].  `mat` is the autograd leaf/edge the loss kernels return d/d pose into."""

    def __init__(self, mat):
        assert tuple(mat.shape[-2:]) == (4, 4)
        if mat.dim() == 2:
            mat = mat.unsqueeze(0)
        assert mat.dim() == 3
        self.mat = mat

    def __len__(self):
        return len(self.mat)

    @classmethod
    def identity(cls, N=1, device=None, dtype=torch.float):
        return cls(torch.eye(4, device=device, dtype=dtype).repeat([N, 1, 1]))

    @classmethod
    def from_vec(cls, vec, mode):
        """[B,6] pose vector -> Pose (reference :39-46)."""
        if mode == 'euler' and vec.is_cuda and vec.dim() == 2 and vec.shape[1] == 6 and \
                vec.dtype in (torch.float32, torch.float64):
            from packnet_sfm.block_ops import pose_from_vec_op
            return cls(pose_from_vec_op(vec))                                 # one CUDA kernel each way
        top = pose_vec2mat(vec, mode)                                         # [B,3,4]
        mat = torch.eye(4, device=vec.device, dtype=vec.dtype).repeat([len(vec), 1, 1])
        mat[:, :3, :3] = top[:, :3, :3]
        mat[:, :3, -1] = top[:, :3, -1]
        return cls(mat)

    @property
    def shape(self):
        return self.mat.shape

    def item(self):
        return self.mat

    def repeat(self, *args, **kwargs):
        self.mat = self.mat.repeat(*args, **kwargs)
        return self

    def inverse(self):
        return Pose(invert_pose(self.mat))

    def to(self, *args, **kwargs):
        self.mat = self.mat.to(*args, **kwargs)
        return self

    def transform_pose(self, pose):
        """self * pose (reference :75-78)."""
        assert tuple(pose.shape[-2:]) == (4, 4)
        return Pose(self.mat.bmm(pose.item()))

    def transform_points(self, points):
        """R X + t on [B,3,H,W] points (reference :80-86), CUDA kernel through the C ABI."""
        assert points.shape[1] == 3
        from packnet_sfm.ops import transform_points_op
        return transform_points_op(self.mat, points)

    # north-star name for the same operation (BASELINE.json); the reference has no such method
    transform_pairs = transform_points

    def __matmul__(self, other):
        if isinstance(other, Pose):
            return self.transform_pose(other)
        elif isinstance(other, torch.Tensor):
            if other.shape[1] == 3 and other.dim() > 2:
                assert other.dim() == 3 or other.dim() == 4
                return self.transform_points(other)
            else:
                raise ValueError('Unknown tensor dimensions {}'.format(other.shape))
        else:
            raise NotImplementedError()
