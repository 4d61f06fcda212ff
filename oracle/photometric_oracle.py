"""CPU oracle for the multi-view photometric loss path  --  TEST INFRASTRUCTURE ONLY.

This file is a restatement, in plain functional PyTorch (fp32 or fp64, CPU), of the
algorithm the reference implements for its self-supervised hot path.  It is the checker
for the CUDA kernels; it is *never* on the product path.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference` legs may
import it.  The product (`packnet-sfm-resnet-san_b200/`) must not import anything from
`oracle/` and raises when its CUDA library is missing.

Parity pin: the reference ships no tests for this path, so the oracle is
pinned against outputs of the reference's own classes run in the build container:
`oracle/gen_golden.py` imports `/root/reference` (pinhole `warp_ref_image`, SURVEY
Appendix C), and `tests/golden/*.npz` hold what it produced; `tests/test_oracle_golden.py`
checks this restatement against those fixtures on every run.  The fisheye (VADAS) camera functions
 are pinned the same way: `oracle/gen_golden_fisheye.py` runs the reference's own
`FisheyeCamera.reconstruct` / `.project` (frame 'c', which execute as shipped) inside its loss class and
writes `tests/golden/fisheye_*.npz`.

Third-party arithmetic: `F.grid_sample`, `F.avg_pool2d`, reflection `F.pad` and
`F.interpolate` are PyTorch primitives (reference pins torch==1.9.0+cu111 in
docker/Dockerfile:25; this image has torch 2.11.0) -- they are called here exactly as
the reference calls them (geometry/camera_utils.py:58-59,
losses/multiview_photometric_loss.py:33-46, utils/image.py:145).

All `file:line` citations are relative to the reference repository root.
"""
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# geometry
# --------------------------------------------------------------------------------------

def euler_to_rotation(angles):
    """[B,3] (rx, ry, rz) -> [B,3,3], R = Rx @ Ry @ Rz  (geometry/pose_utils.py:8-37)."""
    rx, ry, rz = angles[:, 0], angles[:, 1], angles[:, 2]
    o, z = torch.ones_like(rx), torch.zeros_like(rx)
    cx, sx, cy, sy, cz, sz = rx.cos(), rx.sin(), ry.cos(), ry.sin(), rz.cos(), rz.sin()
    Rx = torch.stack([o, z, z, z, cx, -sx, z, sx, cx], 1).view(-1, 3, 3)
    Ry = torch.stack([cy, z, sy, z, o, z, -sy, z, cy], 1).view(-1, 3, 3)
    Rz = torch.stack([cz, -sz, z, sz, cz, z, z, z, o], 1).view(-1, 3, 3)
    return Rx @ Ry @ Rz


def pose_vec_to_mat(vec):
    """[B,6] = [tx,ty,tz,rx,ry,rz] -> [B,4,4]  (geometry/pose_utils.py:41-51, geometry/pose.py:39-46)."""
    B = vec.shape[0]
    top = torch.cat([euler_to_rotation(vec[:, 3:]), vec[:, :3].unsqueeze(-1)], 2)
    bottom = torch.tensor([0, 0, 0, 1], dtype=vec.dtype, device=vec.device).expand(B, 1, 4)
    return torch.cat([top, bottom], 1)


def invert_rigid(T):
    """[B,4,4] rigid inverse [R^T | -R^T t]  (geometry/pose_utils.py:55-60)."""
    Rt = T[:, :3, :3].transpose(1, 2)
    t = -(Rt @ T[:, :3, 3:])
    out = torch.eye(4, dtype=T.dtype, device=T.device).repeat(T.shape[0], 1, 1)
    out = out.clone()
    out[:, :3, :3] = Rt
    out[:, :3, 3:] = t
    return out


def scale_K(K, sx, sy=None):
    """Intrinsics of the resized image (geometry/camera_utils.py:16-22, camera.py:84-108).

    `Camera.scaled(sf)` uses the x factor for y when called with one argument, and returns
    the camera unchanged when both factors are 1."""
    sy = sx if sy is None else sy
    if sx == 1. and sy == 1.:
        return K
    K = K.clone()
    K[:, 0, 0] = K[:, 0, 0] * sx
    K[:, 1, 1] = K[:, 1, 1] * sy
    K[:, 0, 2] = (K[:, 0, 2] + 0.5) * sx - 0.5
    K[:, 1, 2] = (K[:, 1, 2] + 0.5) * sy - 0.5
    return K


def lifting_matrix(K):
    """The reference's `Kinv`: a copy of K with four entries replaced (camera.py:73-81)."""
    Ki = K.clone()
    Ki[:, 0, 0] = 1. / K[:, 0, 0]
    Ki[:, 1, 1] = 1. / K[:, 1, 1]
    Ki[:, 0, 2] = -1. * K[:, 0, 2] / K[:, 0, 0]
    Ki[:, 1, 2] = -1. * K[:, 1, 2] / K[:, 1, 1]
    return Ki


def pixel_grid(B, H, W, dtype, device):
    """[B,3,H*W] rows (u, v, 1), u = 0..W-1 fastest  (utils/image.py:218-282)."""
    u = torch.linspace(0, W - 1, W, dtype=dtype, device=device)
    v = torch.linspace(0, H - 1, H, dtype=dtype, device=device)
    vv, uu = torch.meshgrid(v, u, indexing='ij')
    return torch.stack([uu, vv, torch.ones_like(uu)], 0).reshape(1, 3, -1).repeat(B, 1, 1)


def inv2depth(inv_depth):
    """1 / clamp(inv_depth, 1e-6)  (utils/depth.py:103-120)."""
    return 1. / inv_depth.clamp(min=1e-6)


def reconstruct(depth, K, Twc=None):
    """Back-project a depth map [B,1,H,W] -> points [B,3,H,W]  (camera.py:111-147).

    Twc: optional [B,4,4] world<-camera... applied as R X + t (pose.py:80-86)."""
    B, _, H, W = depth.shape
    rays = (lifting_matrix(K) @ pixel_grid(B, H, W, depth.dtype, depth.device)).view(B, 3, H, W)
    X = rays * depth
    if Twc is not None:
        X = transform_points(Twc, X)
    return X


def transform_points(T, X):
    """T[:, :3, :3] @ X + T[:, :3, 3]  (pose.py:80-86)."""
    B, _, H, W = X.shape
    return (T[:, :3, :3] @ X.reshape(B, 3, -1) + T[:, :3, 3:]).view(B, 3, H, W)


def project(X, K, Tcw=None):
    """Points [B,3,H,W] -> normalised grid [B,H,W,2]  (camera.py:149-190)."""
    B, _, H, W = X.shape
    if Tcw is not None:
        X = transform_points(Tcw, X)
    p = K @ X.reshape(B, 3, -1)
    Z = p[:, 2].clamp(min=1e-5)
    xn = 2 * (p[:, 0] / Z) / (W - 1) - 1.
    yn = 2 * (p[:, 1] / Z) / (H - 1) - 1.
    return torch.stack([xn, yn], -1).view(B, H, W, 2)


def view_synthesis(ref_image, depth, ref_K, Tcw_ref, K, padding_mode='zeros'):
    """Warp `ref_image` into the target view  (geometry/camera_utils.py:27-59).

    The target camera has identity pose; the reference still multiplies by it
    (camera.py:144) which is exact, so it is skipped here."""
    coords = project(reconstruct(depth, K), ref_K, Tcw_ref)
    return F.grid_sample(ref_image, coords, mode='bilinear', padding_mode=padding_mode, align_corners=True)


# --------------------------------------------------------------------------------------
# fisheye (VADAS) camera  --  SURVEY.md §8(f) N1
# Intrinsics are a dict of per-image tensors: k [B,7], s [B], div [B], ux [B], uy [B]
# (datasets/ncdb_dataset.py:577-584; geometry/camera.py:199-220).
# --------------------------------------------------------------------------------------

FISHEYE_EPS = 2.220446049250313e-16      # sys.float_info.epsilon (geometry/camera.py:287,355,373)


def fisheye_scaled(intr, sw, sh):
    """Per-scale intrinsics as the fork's `warp_ref_image` builds them: only the principal point moves,
    k / s / div are passed through unscaled (losses/multiview_photometric_loss.py:158-183)."""
    out = dict(intr)
    out['ux'] = (intr['ux'] + 0.5) * sw - 0.5
    out['uy'] = (intr['uy'] + 0.5) * sh - 0.5
    return out


def fisheye_reconstruct(depth, intr):
    """`FisheyeCamera.reconstruct(depth, frame='c')`: [B,1,H,W] -> [B,3,H,W]  (geometry/camera.py:242-305).
    The inverse of the distortion polynomial is the reference's placeholder theta = r_d."""
    B, _, H, W = depth.shape
    g = pixel_grid(B, H, W, depth.dtype, depth.device)
    xd = (g[:, 0] - intr['ux'].unsqueeze(1)) / intr['s'].unsqueeze(1)
    yd = (g[:, 1] - intr['uy'].unsqueeze(1)) / intr['div'].unsqueeze(1)
    rd = torch.sqrt(xd ** 2 + yd ** 2)
    r = torch.tan(rd)
    rd_safe = torch.where(rd < FISHEYE_EPS, torch.full_like(rd, FISHEYE_EPS), rd)
    d = depth.reshape(B, -1)
    return torch.stack([(r / rd_safe) * xd * d, (r / rd_safe) * yd * d, d], 1).view(B, 3, H, W)


def fisheye_project(X, intr):
    """`FisheyeCamera.project(X, frame='c')`: [B,3,H,W] -> normalised grid [B,H,W,2]  (geometry/camera.py:311-394)."""
    B, _, H, W = X.shape
    Xc = X.reshape(B, 3, -1)
    Z = Xc[:, 2].clamp(min=FISHEYE_EPS)
    xn, yn = Xc[:, 0] / Z, Xc[:, 1] / Z
    r = torch.sqrt(xn ** 2 + yn ** 2)
    theta = torch.atan(r)
    k = intr['k']
    poly = k[:, 0].unsqueeze(1)
    for i in range(1, 7):
        poly = poly + k[:, i].unsqueeze(1) * torch.pow(theta, i)
    r_safe = torch.where(r < FISHEYE_EPS, torch.full_like(r, FISHEYE_EPS), r)
    u = intr['s'].unsqueeze(1) * ((poly / r_safe) * xn) + intr['ux'].unsqueeze(1)
    v = intr['div'].unsqueeze(1) * ((poly / r_safe) * yn) + intr['uy'].unsqueeze(1)
    return torch.stack([2 * u / (W - 1) - 1., 2 * v / (H - 1) - 1.], -1).view(B, H, W, 2)


def fisheye_view_synthesis(ref_image, depth, ref_intr, Tcw_ref, intr, padding_mode='zeros'):
    """`view_synthesis` with two `FisheyeCamera`s (geometry/camera_utils.py:27-59).  As shipped the reference
    raises inside `reconstruct(frame='w')` / `project(frame='w')` because `Pose.transform_points` takes
    [B,3,H,W] only; this is the composition those lines spell out, with the pose
    applied to the 4-D tensor: project_c(Tcw_ref . reconstruct_c(depth))."""
    coords = fisheye_project(transform_points(Tcw_ref, fisheye_reconstruct(depth, intr)), ref_intr)
    return F.grid_sample(ref_image, coords, mode='bilinear', padding_mode=padding_mode, align_corners=True)


def fisheye_pack(intr):
    """dict -> [B,11] rows (k0..k6, s, div, ux, uy): the layout of the C ABI (`psfm_loss_io::fisheye`)."""
    return torch.cat([intr['k'], intr['s'].unsqueeze(1), intr['div'].unsqueeze(1), intr['ux'].unsqueeze(1),
                      intr['uy'].unsqueeze(1)], 1)


def synth_fisheye(B, H, W, dtype=torch.float64):
    """Deterministic VADAS-like intrinsics for the synthetic recipe (per-image variation, principal point off
    the pixel grid so that no pixel sits exactly on the optical axis)."""
    t = torch.arange(B, dtype=torch.float64)
    k = torch.tensor([0.002, 1.0, -0.01, -0.06, 0.004, 0.012, -0.002], dtype=torch.float64).repeat(B, 1)
    k[:, 1] += 0.01 * t
    k[:, 3] -= 0.003 * t
    intr = dict(k=k, s=0.62 * W + 1.5 * t, div=0.60 * W - 0.8 * t, ux=0.5 * W - 0.37 + 0.9 * t,
                uy=0.5 * H + 0.21 - 0.6 * t)
    return {n: v.to(dtype) for n, v in intr.items()}


# --------------------------------------------------------------------------------------
# image pyramid helpers
# --------------------------------------------------------------------------------------

def resize_like(image, target, mode='bilinear', align_corners=True):
    """`interpolate_image`: identity when the spatial shapes agree (utils/image.py:117-146)."""
    if tuple(image.shape[-2:]) == tuple(target.shape[-2:]):
        return image
    return F.interpolate(image, size=tuple(target.shape[-2:]), mode=mode, align_corners=align_corners)


def match_scales(image, targets, n, mode='bilinear', align_corners=True):
    """utils/image.py:178-214."""
    return [resize_like(image, targets[i], mode, align_corners) for i in range(n)]


# --------------------------------------------------------------------------------------
# photometric terms
# --------------------------------------------------------------------------------------

def ssim_index(x, y, C1=1e-4, C2=9e-4):
    """3x3 SSIM with reflection padding  (losses/multiview_photometric_loss.py:15-54)."""
    xp, yp = F.pad(x, (1, 1, 1, 1), mode='reflect'), F.pad(y, (1, 1, 1, 1), mode='reflect')
    box = lambda t: F.avg_pool2d(t, 3, 1)
    mx, my = box(xp), box(yp)
    sx = box(xp * xp) - mx * mx
    sy = box(yp * yp) - my * my
    sxy = box(xp * yp) - mx * my
    return ((2 * mx * my + C1) * (2 * sxy + C2)) / ((mx * mx + my * my + C1) * (sx + sy + C2))


def ssim_distance(x, y, C1=1e-4, C2=9e-4):
    """clamp((1 - ssim) / 2, 0, 1)  (losses/multiview_photometric_loss.py:199-216)."""
    return torch.clamp((1. - ssim_index(x, y, C1, C2)) / 2., 0., 1.)


def photometric_map(est, image, ssim_w=0.85, C1=1e-4, C2=9e-4, clip_loss=0.0, mask=None):
    """One scale of `calc_photometric_loss`  (losses/multiview_photometric_loss.py:218-267)."""
    l1 = (est - image).abs()
    if ssim_w > 0.0:
        p = ssim_w * ssim_distance(est, image, C1, C2).mean(1, True) + (1 - ssim_w) * l1.mean(1, True)
    else:
        p = l1
    if clip_loss > 0.0:
        p = torch.clamp(p, max=float(p.mean() + clip_loss * p.std()))
    if mask is not None:
        p = p * (mask.expand_as(p) if p.shape[1] == 3 and mask.shape[1] == 1 else mask)
    return p


def reduce_candidates(cands, op='min'):
    """One scale of `reduce_photometric_loss`  (losses/multiview_photometric_loss.py:269-297)."""
    if op == 'mean':
        return sum(c.mean() for c in cands) / len(cands)
    if op == 'min':
        return torch.cat(cands, 1).min(1, True)[0].mean()
    raise NotImplementedError('Unknown photometric_reduce_op: {}'.format(op))


def smoothness_maps(inv_depths, images, n):
    """`calc_smoothness` + `inv_depths_normalize`  (utils/depth.py:146-198, utils/image.py:85-113)."""
    sx, sy = [], []
    for i in range(n):
        d = inv_depths[i]
        dn = d / d.mean(2, True).mean(3, True).clamp(min=1e-6)
        gx, gy = dn[:, :, :, :-1] - dn[:, :, :, 1:], dn[:, :, :-1, :] - dn[:, :, 1:, :]
        ix = images[i][:, :, :, :-1] - images[i][:, :, :, 1:]
        iy = images[i][:, :, :-1, :] - images[i][:, :, 1:, :]
        sx.append(gx * torch.exp(-ix.abs().mean(1, True)))
        sy.append(gy * torch.exp(-iy.abs().mean(1, True)))
    return sx, sy


def smoothness_loss(inv_depths, images, n, weight):
    """`calc_smoothness_loss`  (losses/multiview_photometric_loss.py:301-327)."""
    sx, sy = smoothness_maps(inv_depths, images, n)
    return weight * sum((sx[i].abs().mean() + sy[i].abs().mean()) / 2 ** i for i in range(n)) / n


def fork_sigmoid_premap(sig, min_depth, max_depth):
    """Fork-only sigmoid -> inverse depth pre-map  (losses/multiview_photometric_loss.py:362-369,
    utils/post_process_depth.py:68-108)."""
    lo, hi = 1.0 / max(max_depth, 1e-6), 1.0 / max(min_depth, 1e-6)
    depth = 1.0 / ((lo + (hi - lo) * sig) + 1e-8)
    return 1.0 / (depth + 1e-8)


# --------------------------------------------------------------------------------------
# the whole path
# --------------------------------------------------------------------------------------

def multiview_photometric_loss(image, context, inv_depths, K, ref_K, pose_mats, num_scales=4,
                               ssim_loss_weight=0.85, smooth_loss_weight=0.001, C1=1e-4, C2=9e-4,
                               photometric_reduce_op='min', clip_loss=0.0, padding_mode='zeros',
                               automask_loss=True, mask=None, fork_premap=None, return_parts=False):
    """`MultiViewPhotometricLoss.forward` with the pinhole `warp_ref_image`
    (losses/multiview_photometric_loss.py:331-410; upstream warp body = SURVEY Appendix C).

    pose_mats: list of [B,4,4] target->context transforms (Pose.mat).
    fork_premap: None, or (min_depth, max_depth) to apply the fork's sigmoid pre-map.
    Returns dict(loss [1], photometric_loss, smoothness_loss) (+ per-scale lists when return_parts).
    """
    n = num_scales
    raw = inv_depths
    if fork_premap is not None:
        inv_depths = [fork_sigmoid_premap(s, *fork_premap) for s in inv_depths]
    W = image.shape[-1]
    images = match_scales(image, inv_depths, n)
    masks = None if mask is None else match_scales(mask, inv_depths, n, mode='nearest', align_corners=None)
    cands = [[] for _ in range(n)]
    warped_all = []
    hp = dict(ssim_w=ssim_loss_weight, C1=C1, C2=C2, clip_loss=clip_loss)
    for ref_image, T in zip(context, pose_mats):
        ref_images = match_scales(ref_image, inv_depths, n)
        warped = []
        for i in range(n):
            sf = inv_depths[i].shape[3] / float(W)
            if isinstance(K, dict):      # fisheye intrinsics (the fork's forward, :331-410 with :131-195)
                sh = inv_depths[i].shape[2] / float(image.shape[-2])
                warped.append(fisheye_view_synthesis(ref_images[i], inv2depth(inv_depths[i]),
                                                     fisheye_scaled(ref_K, sf, sh), T, fisheye_scaled(K, sf, sh),
                                                     padding_mode))
                continue
            warped.append(view_synthesis(ref_images[i], inv2depth(inv_depths[i]), scale_K(ref_K, sf), T,
                                         scale_K(K, sf), padding_mode))
        warped_all.append(warped)
        for i in range(n):
            cands[i].append(photometric_map(warped[i], images[i], mask=None if masks is None else masks[i], **hp))
        if automask_loss:
            for i in range(n):
                cands[i].append(photometric_map(ref_images[i], images[i],
                                                mask=None if masks is None else masks[i], **hp))
    photo = sum(reduce_candidates(cands[i], photometric_reduce_op) for i in range(n)) / n
    loss = photo
    smooth = torch.zeros((), dtype=image.dtype, device=image.device)
    if smooth_loss_weight > 0.0:
        smooth = smoothness_loss(raw, images, n, smooth_loss_weight)
        loss = loss + smooth
    out = {'loss': loss.unsqueeze(0), 'photometric_loss': photo.detach(), 'smoothness_loss': smooth.detach()}
    if return_parts:
        out['candidates'] = cands
        out['warped'] = warped_all
    return out


# --------------------------------------------------------------------------------------
# deterministic synthetic inputs (recipe S3)
# --------------------------------------------------------------------------------------

def synth_S3(B, H, W, C, dtype=torch.float32, seed=1234, pyramid=False, k=3, n=4, kd=15):
    """Deterministic synthetic batch.  Returns (image, context[C], inv_depths[n], vecs[C], K).
    k / kd: box-filter sizes of the image / inverse-depth textures (kd must stay below the
    smallest pyramid dimension because of the reflect pad)."""
    g = torch.Generator('cpu').manual_seed(seed)

    def tex(m, ch, h, w, kk):
        x = torch.rand(m, ch, h, w, generator=g, dtype=torch.float64)
        if kk > 1:
            x = F.avg_pool2d(F.pad(x, (kk // 2,) * 4, mode='reflect'), kk, 1)
        lo, hi = x.amin((2, 3), True), x.amax((2, 3), True)
        return (x - lo) / (hi - lo)

    image = tex(B, 3, H, W, k)
    context = [tex(B, 3, H, W, k) for _ in range(C)]
    inv_depths = [1.0 / (2.0 + 38.0 * tex(B, 1, (H >> s) if pyramid else H, (W >> s) if pyramid else W, kd))
                  for s in range(n)]
    base = torch.tensor([0.02, -0.01, 0.8, 0.003, -0.01, 0.002], dtype=torch.float64)
    vecs = []
    for j in range(C):
        v = base.clone()
        v[2] *= (-1.0 if j % 2 == 0 else 1.0)
        vecs.append(v.repeat(B, 1) + 0.01 * torch.randn(B, 6, generator=g, dtype=torch.float64))
    K = torch.tensor([[0.58 * W, 0, 0.5 * W], [0, 1.92 * H, 0.5 * H], [0, 0, 1.]], dtype=torch.float64).repeat(B, 1, 1)
    c = lambda t: t.to(dtype)
    return c(image), [c(x) for x in context], [c(d) for d in inv_depths], [c(v) for v in vecs], c(K)
