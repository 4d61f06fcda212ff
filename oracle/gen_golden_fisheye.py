"""Generate tests/golden/fisheye_*.npz from the REAL reference  --  TEST INFRASTRUCTURE, build container only.

SURVEY.md §8(f) N1: the fisheye (VADAS) camera of the fork's `warp_ref_image`.  The reference's own
`FisheyeCamera.reconstruct(depth, frame='c')` and `FisheyeCamera.project(X, frame='c')` run as shipped; the
`frame='w'` branches raise (they hand a [B,3,HW] tensor to `Pose.transform_points`).
The fixtures therefore come from the reference's classes composed the way its `view_synthesis` spells out,
with the pose applied to the 4-D point tensor:  project_c( pose @ reconstruct_c(depth) ), followed by the
reference's `grid_sample` call and its unchanged SSIM / L1 / automask / min / smoothness tail.  The two
other defects on that path (`.get_device()` on CPU, `masks_scaled = [None] * n`) are avoided the same way
`gen_golden.py` does for the pinhole path.

    python oracle/gen_golden_fisheye.py      # writes tests/golden/fisheye_*.npz
"""
import os

import numpy as np
import torch
import torch.nn.functional as funct

import gen_golden as G          # stubs yacs / matplotlib and puts /root/reference first on sys.path
from gen_golden import PinholeReference, Pose, inv2depth, match_scales, YAML_LOSS, ROOT

from packnet_sfm.geometry.camera import FisheyeCamera  # noqa: E402
from photometric_oracle import synth_S3, synth_fisheye  # noqa: E402  (input recipe only)


class FisheyeReference(PinholeReference):
    """The reference loss with its own fisheye cameras (losses/multiview_photometric_loss.py:131-195)."""

    def warp_ref_image(self, inv_depths, ref_image, intrinsics, ref_intrinsics, pose):
        B, _, H, W = ref_image.shape
        out = []
        ref_images = match_scales(ref_image, inv_depths, self.n)
        for i in range(self.n):
            _, _, DH, DW = inv_depths[i].shape
            sw, sh = DW / float(W), DH / float(H)
            scaled = lambda d: {'k': d['k'].clone(), 's': d['s'].clone(), 'div': d['div'].clone(),
                                'ux': (d['ux'].clone() + 0.5) * sw - 0.5, 'uy': (d['uy'].clone() + 0.5) * sh - 0.5}
            cam = FisheyeCamera(intrinsics=scaled(intrinsics), image_size=(DH, DW))
            ref_cam = FisheyeCamera(intrinsics=scaled(ref_intrinsics), Tcw=pose, image_size=(DH, DW))
            depth = inv2depth(inv_depths[i])
            points = cam.reconstruct(depth, frame='c')                       # [B,3,H,W], camera.py:242-305
            coords = ref_cam.project(ref_cam.Tcw @ points, frame='c')        # pose.py:80-99, camera.py:311-394
            out.append(funct.grid_sample(ref_images[i], coords, mode='bilinear', padding_mode=self.padding_mode,
                                         align_corners=True))               # camera_utils.py:58-59
        return out


def run_case(B, H, W, C, pyramid, kd=15, full=False, with_mask=False, seed=1234, **hp):
    cfg = dict(YAML_LOSS)
    cfg.update(hp)
    image, context, inv_depths, vecs, _ = synth_S3(B, H, W, C, torch.float64, seed=seed, pyramid=pyramid, kd=kd)
    intr = synth_fisheye(B, H, W)
    inv_depths = [d.requires_grad_() for d in inv_depths]
    vecs = [v.requires_grad_() for v in vecs]
    poses = [Pose.from_vec(v, 'euler') for v in vecs]
    for p in poses:
        p.mat.retain_grad()
    mask = None
    if with_mask:
        gm = torch.Generator('cpu').manual_seed(seed + 7)
        mask = (torch.rand(B, 1, H, W, generator=gm, dtype=torch.float64) > 0.3).to(torch.float64)
    L = FisheyeReference(**cfg)
    out = L(image, context, inv_depths, intr, intr, poses, mask=mask)
    out['loss'].backward()
    res = {
        'loss': out['loss'].detach().numpy(),
        'photometric_loss': out['metrics']['photometric_loss'].numpy(),
        'smoothness_loss': out['metrics'].get('smoothness_loss', torch.zeros(())).numpy(),
        'd_vec': torch.stack([v.grad for v in vecs]).numpy(),
        'd_posemat': torch.stack([p.mat.grad for p in poses]).numpy(),
        'd_invd_norm': np.array([d.grad.norm().item() for d in inv_depths]),
        'd_invd_sum': np.array([d.grad.sum().item() for d in inv_depths]),
        'image_probe': image[0, 0, 0, :3].numpy(),
    }
    for i in range(L.n):
        w0 = L.kept['warped'][0][i].detach()
        res['warp0_s%d_stats' % i] = np.array([w0.mean().item(), w0.abs().sum().item(), (w0 == 0).double().mean().item()])
    if full:
        for i in range(L.n):
            res['d_invd_s%d' % i] = inv_depths[i].grad.numpy()
    if full == 2:
        for i in range(L.n):
            for j in range(C):
                res['warp%d_s%d' % (j, i)] = L.kept['warped'][j][i].detach().numpy()
    return res


def unit_vectors():
    """Known answers of the reference's FisheyeCamera methods (fp64)."""
    g = torch.Generator('cpu').manual_seed(321)
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64)
    B, H, W = 2, 7, 11
    intr = synth_fisheye(B, H, W)
    depth = 1.0 + 5.0 * r(B, 1, H, W)
    vec = (r(B, 6) - 0.5) * torch.tensor([0.5, 0.5, 0.5, 0.2, 0.2, 0.2], dtype=torch.float64)
    pose = Pose.from_vec(vec, 'euler')
    img = r(B, 3, H, W)
    cam = FisheyeCamera(intrinsics={k: v.clone() for k, v in intr.items()}, image_size=(H, W))
    X = cam.reconstruct(depth, frame='c')
    Xr = pose @ X
    out = dict(depth=depth.numpy(), vec=vec.numpy(), img=img.numpy(), reconstruct_c=X.numpy(),
               project_c=cam.project(Xr, frame='c').numpy(), pose_mat=pose.mat.numpy())
    # points behind / on the clamp of Z and far off-axis
    Xe = Xr.clone()
    Xe[:, 2, 0, :] = -0.5
    Xe[:, 2, 1, :] = 0.0
    Xe[:, 0, 2, :] *= 50.0
    out['edge_points'] = Xe.numpy()
    out['project_edge'] = cam.project(Xe, frame='c').numpy()
    for pm in ('zeros', 'border'):
        out['view_synthesis_' + pm] = funct.grid_sample(img, cam.project(Xr, frame='c'), mode='bilinear', padding_mode=pm,
                                                        align_corners=True).numpy()
    for name, v in intr.items():
        out['intr_' + name] = v.numpy()
    return out


def main():
    save = lambda name, d: np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', name + '.npz'), **d)
    save('fisheye_units', unit_vectors())
    tiny = dict(B=2, H=24, W=40, C=2, kd=5, full=2)
    save('fisheye_tiny_U_zeros', run_case(pyramid=False, **tiny))
    save('fisheye_tiny_U_border', run_case(pyramid=False, padding_mode='border', **tiny))
    save('fisheye_small_P_zeros', run_case(pyramid=True, **dict(tiny, H=32, W=64, kd=3)))     # smallest scale 4 x 8
    save('fisheye_tiny_U_mask', run_case(pyramid=False, with_mask=True, **dict(tiny, full=1)))
    save('fisheye_tiny_U_odd_C4', run_case(pyramid=False, **dict(tiny, H=23, W=37, C=4, full=1)))
    for pm in ('zeros', 'border'):
        save('fisheye_B2_64x160_C2_U_' + pm, run_case(2, 64, 160, 2, False, padding_mode=pm))
    save('fisheye_B1_192x640_C2_U_zeros', run_case(1, 192, 640, 2, False))
    print('done')


if __name__ == '__main__':
    main()
