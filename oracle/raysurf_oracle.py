"""CPU restatement of the ray-surface (generic camera) photometric loss  --  TEST INFRASTRUCTURE.

Follows `GenericMultiViewPhotometricLoss` (losses/generic_multiview_photometric_loss.py:92-400), `GenericCamera`
(geometry/camera_generic.py:17-208) and `view_synthesis_generic` (geometry/camera_utils.py:64-96); pinned against
outputs of those classes (oracle/gen_golden_raysurf.py -> tests/golden/raysurf_*.npz, tests/test_oracle_golden.py).
The reference is batch-1 by construction; so is this.  Only tests/, smoke() and bench legs import it."""
import numpy as np
import torch
import torch.nn.functional as F

import photometric_oracle as O

PATCH = 20                      # camera_generic.py:104: patches of (2 * 20 + 1)^2 rays
RAYSURF_LOSS = dict(num_scales=1, ssim_loss_weight=0.85, smooth_loss_weight=0.001, photometric_reduce_op='min',
                    clip_loss=0.0, padding_mode='zeros', automask_loss=True)
RAYSURF_CASES = {
    'raysurf_96x128_p0': (dict(H=96, W=128, C=2, progress=0.0), {}),
    'raysurf_96x128_p05': (dict(H=96, W=128, C=2, progress=0.5), {}),
    'raysurf_100x140_border_mean': (dict(H=100, W=140, C=2, progress=0.2),
                                    dict(padding_mode='border', photometric_reduce_op='mean', automask_loss=False)),
}


def synth_ray_template(H, W):
    """[1,3,H,W] unit viewing rays of a wide-angle camera with a mild radial distortion (stands in for the
    omnicam_ray_template.npy the reference does not ship)."""
    v, u = torch.meshgrid(torch.arange(H, dtype=torch.float64), torch.arange(W, dtype=torch.float64), indexing='ij')
    x = (u - 0.5 * W + 0.3) / (0.9 * W)
    y = (v - 0.5 * H - 0.2) / (0.9 * W)
    r2 = x * x + y * y
    s = 1.0 + 0.15 * r2
    ray = torch.stack([x * s, y * s, torch.ones_like(x)], 0)
    return (ray / ray.norm(dim=0, keepdim=True)).unsqueeze(0)


def synth_raysurf(H, W, C, seed=1234):
    """image, context[C], inv_depth [1,1,H,W], ray-surface residual [1,3,H,W], pose vectors, template."""
    image, context, inv_depths, vecs, _ = O.synth_S3(1, H, W, C, torch.float64, seed=seed, pyramid=False, kd=15, n=1)
    g = torch.Generator('cpu').manual_seed(seed + 11)
    res = torch.rand(1, 3, H, W, generator=g, dtype=torch.float64) - 0.5
    res = 0.02 * F.avg_pool2d(F.pad(res, (3, 3, 3, 3), mode='reflect'), 7, 1)
    vecs = [0.2 * v for v in vecs]          # small motion: the soft arg-max searches a 41 x 41 half-resolution patch
    return image, context, inv_depths[0], res, vecs, synth_ray_template(H, W)


def patch_origin(i, n):
    """First index of the 41-wide patch around i, shifted so that it stays inside [0, n) (camera_generic.py:133-147)."""
    return (i - PATCH).clamp(min=0, max=n - (2 * PATCH + 1))


def raysurf_project(X, rays, progress):
    """`GenericCamera.project(X, progress, downsample=True, frame='c')` (camera_generic.py:86-208): soft arg-max, over a
    41 x 41 patch of half-resolution rays, of the cosine between each point direction and the rays; returns the
    normalised sampling grid [1,H,W,2].  X, rays: [1,3,H,W]."""
    _, _, H, W = X.shape
    h, w = int(H / 2.), int(W / 2.)
    rs = F.interpolate(rays, mode='bilinear', scale_factor=0.5, align_corners=True)[0]        # [3,h,w]
    d = F.interpolate(X, mode='bilinear', scale_factor=0.5, align_corners=True)[0]
    d = d / d.norm(dim=0, keepdim=True)
    temperature = np.maximum(1e-8, 0.0001 / np.exp(0.1 * progress))
    ii, jj = torch.meshgrid(torch.arange(h), torch.arange(w), indexing='ij')
    pi, pj = patch_origin(ii, h), patch_origin(jj, w)                                          # [h,w]
    a = torch.arange(2 * PATCH + 1)
    rows = (pi[..., None, None] + a[None, None, :, None]).expand(h, w, 41, 41)
    cols = (pj[..., None, None] + a[None, None, None, :]).expand(h, w, 41, 41)
    patch = rs[:, rows, cols]                                                                   # [3,h,w,41,41]
    logits = (d[:, :, :, None, None] * patch).sum(0).reshape(h, w, -1)
    p = torch.softmax(logits / temperature, -1)
    row = (p * rows.reshape(h, w, -1).to(X.dtype)).sum(-1)
    col = (p * cols.reshape(h, w, -1).to(X.dtype)).sum(-1)
    xn = 2 * row / (h - 1) - 1.
    yn = 2 * col / (w - 1) - 1.
    up = lambda t: F.interpolate(t[None, None], mode='bilinear', scale_factor=2.0, align_corners=True)[0, 0]
    return torch.stack([up(yn), up(xn)], -1).view(1, 2 * h, 2 * w, 2)


def ray_surface(template, residual, progress):
    """losses/generic_multiview_photometric_loss.py:193-195."""
    coeff = float(np.min([((100.0 * progress) ** (4 / 3.) / 100.), 1.]))
    R = template + coeff * residual
    return R / R.norm(dim=1, keepdim=True)


def view_synthesis_generic(ref_image, depth, R, Tcw, padding_mode, progress):
    """geometry/camera_utils.py:64-96 with two GenericCameras sharing the ray surface R."""
    Xw = R * depth[0].unsqueeze(0)                       # reconstruct, identity pose (camera_generic.py:72-80)
    Xc = O.transform_points(Tcw, Xw)                     # frame 'w': Tcw @ X (camera_generic.py:115-116)
    coords = raysurf_project(Xc, R, progress)
    return F.grid_sample(ref_image, coords, mode='bilinear', padding_mode=padding_mode, align_corners=True)


def generic_photometric_loss(image, context, inv_depth, residual, template, pose_mats, progress, num_scales=1,
                             ssim_loss_weight=0.85, smooth_loss_weight=0.001, C1=1e-4, C2=9e-4,
                             photometric_reduce_op='min', clip_loss=0.0, padding_mode='zeros', automask_loss=True):
    """`GenericMultiViewPhotometricLoss.forward` for one scale (:337-400)."""
    assert num_scales == 1
    R = ray_surface(template, residual, progress)
    depth = O.inv2depth(inv_depth)
    hp = dict(ssim_w=ssim_loss_weight, C1=C1, C2=C2, clip_loss=clip_loss)
    cands, warped = [], []
    for ref_image, T in zip(context, pose_mats):
        wimg = view_synthesis_generic(ref_image, depth, R, T, padding_mode, progress)
        warped.append(wimg)
        cands.append(O.photometric_map(wimg, image, **hp))
        if automask_loss:
            cands.append(O.photometric_map(ref_image, image, **hp))
    photo = O.reduce_candidates(cands, photometric_reduce_op)
    loss = photo
    smooth = torch.zeros((), dtype=image.dtype)
    if smooth_loss_weight > 0.0:
        smooth = O.smoothness_loss([inv_depth], [image], 1, smooth_loss_weight)
        loss = loss + smooth
    return dict(loss=loss.unsqueeze(0), photometric_loss=photo.detach(), smoothness_loss=smooth.detach(), warped=warped)
