"""Image helpers on the photometric-loss path (reference: utils/image.py:85-282)."""
from functools import lru_cache

import torch

from packnet_sfm.utils.misc import same_shape


def gradient_x(image):
    """image[..., :-1] - image[..., 1:] along x (reference :85-98)."""
    return image[:, :, :, :-1] - image[:, :, :, 1:]


def gradient_y(image):
    """image[..., :-1, :] - image[..., 1:, :] along y (reference :100-113)."""
    return image[:, :, :-1, :] - image[:, :, 1:, :]


def interpolate_image(image, shape, mode='bilinear', align_corners=True):
    """Resize to `shape`; returns `image` itself when the spatial size already matches
    (reference :117-146).  bilinear (both align_corners settings) and nearest run as CUDA kernels, with their adjoint
    under autograd."""
    if len(shape) > 2:
        shape = shape[-2:]
    if same_shape(image.shape[-2:], shape):
        return image
    from packnet_sfm.ops import resize_op
    if mode == 'bilinear':
        return resize_op(image, shape, 'bilinear' if align_corners else 'bilinear_ac_false')
    if mode == 'nearest':
        return resize_op(image, shape, 'nearest')
    raise NotImplementedError('interpolate_image supports the bilinear and nearest modes')


def interpolate_scales(images, shape=None, mode='bilinear', align_corners=False):
    """Resize a list of images to one shape, default the first image's (reference :148-176)."""
    if shape is None:
        shape = images[0].shape
    if len(shape) > 2:
        shape = shape[-2:]
    return [interpolate_image(image, shape, mode=mode, align_corners=align_corners) for image in images]


def match_scales(image, targets, num_scales, mode='bilinear', align_corners=True):
    """One resized copy of `image` per target resolution (reference :178-214)."""
    return [interpolate_image(image, targets[i].shape, mode=mode, align_corners=align_corners)
            for i in range(num_scales)]


@lru_cache(maxsize=None)
def meshgrid(B, H, W, dtype, device, normalized=False):
    """xs [B,H,W], ys [B,H,W] pixel (or [-1,1]) coordinates (reference :218-251)."""
    if normalized:
        xs = torch.linspace(-1, 1, W, device=device, dtype=dtype)
        ys = torch.linspace(-1, 1, H, device=device, dtype=dtype)
    else:
        xs = torch.linspace(0, W - 1, W, device=device, dtype=dtype)
        ys = torch.linspace(0, H - 1, H, device=device, dtype=dtype)
    ys, xs = torch.meshgrid([ys, xs], indexing='ij')
    return xs.repeat([B, 1, 1]), ys.repeat([B, 1, 1])


@lru_cache(maxsize=None)
def image_grid(B, H, W, dtype, device, normalized=False):
    """[B,3,H,W] grid of (x, y, 1) (reference :254-282).  The kernels generate these
    coordinates in registers; this is kept for API compatibility."""
    xs, ys = meshgrid(B, H, W, dtype, device, normalized=normalized)
    return torch.stack([xs, ys, torch.ones_like(xs)], dim=1)
