"""Generate tests/golden/*.npz from the REAL reference  --  TEST INFRASTRUCTURE, build container only.

Imports the reference's own classes from /root/reference (read-only, never copied) and runs
them in fp64 on the deterministic synthetic recipe `synth_S3` (oracle/photometric_oracle.py).
The composed loss uses the reference's `MultiViewPhotometricLoss` building blocks with the
pinhole `warp_ref_image` body (the fork's fisheye `warp_ref_image` cannot execute, SURVEY.md
§0 fact 1 / Appendix C).  /root/reference does not exist on the GPU box, so only the produced
fixtures travel; this script is committed so the fixtures can be regenerated:

    python oracle/gen_golden.py            # writes tests/golden/*.npz

Nothing here is imported by the product or by the tests.
"""
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get('PSFM_REFERENCE', '/root/reference')

# import-time-only dependencies of the reference that this image lacks
_y, _yc = types.ModuleType('yacs'), types.ModuleType('yacs.config')
_yc.CfgNode = type('CfgNode', (dict,), {})
_y.config = _yc
_m, _mc = types.ModuleType('matplotlib'), types.ModuleType('matplotlib.cm')
_mc.get_cmap = lambda *a, **k: None
_m.cm = _mc
sys.modules.update({'yacs': _y, 'yacs.config': _yc, 'matplotlib': _m, 'matplotlib.cm': _mc})
sys.path.insert(0, REF)
sys.path.insert(1, HERE)

import warnings
warnings.filterwarnings('ignore')

from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss, SSIM  # noqa: E402
from packnet_sfm.geometry.camera import Camera  # noqa: E402
from packnet_sfm.geometry.camera_utils import view_synthesis, scale_intrinsics  # noqa: E402
from packnet_sfm.geometry.pose import Pose  # noqa: E402
from packnet_sfm.geometry.pose_utils import euler2mat, pose_vec2mat, invert_pose  # noqa: E402
from packnet_sfm.utils.depth import inv2depth, depth2inv, calc_smoothness, inv_depths_normalize  # noqa: E402
from packnet_sfm.utils.image import match_scales, gradient_x, gradient_y, interpolate_scales  # noqa: E402
from packnet_sfm.utils.post_process_depth import sigmoid_to_depth_linear  # noqa: E402
import packnet_sfm  # noqa: E402
assert os.path.realpath(packnet_sfm.__file__).startswith(os.path.realpath(REF)), packnet_sfm.__file__

from photometric_oracle import synth_S3  # noqa: E402  (input recipe only)


class PinholeReference(MultiViewPhotometricLoss):
    """The reference loss with the upstream pinhole warp."""

    def __init__(self, *a, fork_premap=False, **k):
        super().__init__(*a, **k)
        self.fork_premap = fork_premap
        self.kept = {}

    def warp_ref_image(self, inv_depths, ref_image, K, ref_K, pose):
        B, _, H, W = ref_image.shape
        cams, ref_cams = [], []
        for i in range(self.n):
            sf = inv_depths[i].shape[3] / float(W)
            cams.append(Camera(K=K, Tcw=Pose.identity(B, device=K.device, dtype=K.dtype)).scaled(sf))
            ref_cams.append(Camera(K=ref_K, Tcw=pose).scaled(sf))
        depths = [inv2depth(inv_depths[i]) for i in range(self.n)]
        ref_images = match_scales(ref_image, inv_depths, self.n)
        return [view_synthesis(ref_images[i], depths[i], ref_cams[i], cams[i], padding_mode=self.padding_mode)
                for i in range(self.n)]

    def forward(self, image, context, inv_depths, K, ref_K, poses, return_logs=False, progress=0.0, mask=None):
        raw = inv_depths
        if self.fork_premap:       # reference :364-369
            depths = [sigmoid_to_depth_linear(s, self.min_depth, self.max_depth) for s in raw]
            inv_depths = [1.0 / (d + 1e-8) for d in depths]
        self.n = self.progressive_scaling(progress)
        losses = [[] for _ in range(self.n)]
        images = match_scales(image, inv_depths, self.n)
        masks = None if mask is None else match_scales(mask, inv_depths, self.n, mode='nearest', align_corners=None)
        self.kept = {'warped': [], 'cands': losses}
        for ref_image, pose in zip(context, poses):
            warped = self.warp_ref_image(inv_depths, ref_image, K, ref_K, pose)
            self.kept['warped'].append(warped)
            pl = self.calc_photometric_loss(warped, images, masks)
            for i in range(self.n):
                losses[i].append(pl[i])
            if self.automask_loss:
                ul = self.calc_photometric_loss(match_scales(ref_image, inv_depths, self.n), images, masks)
                for i in range(self.n):
                    losses[i].append(ul[i])
        loss = self.reduce_photometric_loss(losses)
        if self.smooth_loss_weight > 0.0:
            loss = loss + self.calc_smoothness_loss(raw, images)
        return {'loss': loss.unsqueeze(0), 'metrics': self.metrics}


YAML_LOSS = dict(num_scales=4, ssim_loss_weight=0.85, smooth_loss_weight=0.001, photometric_reduce_op='min',
                 clip_loss=0.0, padding_mode='zeros', automask_loss=True)


def run_case(B, H, W, C, pyramid, kd=15, full=False, with_mask=False, seed=1234, **hp):
    cfg = dict(YAML_LOSS)
    cfg.update(hp)
    fork = cfg.pop('fork_premap', False)
    image, context, inv_depths, vecs, K = synth_S3(B, H, W, C, torch.float64, seed=seed, pyramid=pyramid, kd=kd)
    if fork:   # sigmoid-like raw inputs in (0,1): re-map inverse depths into sigmoids of a 0.5..80 m range
        lo, hi = 1.0 / 80.0, 1.0 / 0.5
        inv_depths = [((d - lo) / (hi - lo)).clamp(0.0, 1.0) for d in inv_depths]
        cfg.update(min_depth=0.5, max_depth=80.0)
    inv_depths = [d.requires_grad_() for d in inv_depths]
    vecs = [v.requires_grad_() for v in vecs]
    poses = [Pose.from_vec(v, 'euler') for v in vecs]
    for p in poses:
        p.mat.retain_grad()
    mask = None
    if with_mask:
        gm = torch.Generator('cpu').manual_seed(seed + 7)
        mask = (torch.rand(B, 1, H, W, generator=gm, dtype=torch.float64) > 0.3).to(torch.float64)
    L = PinholeReference(fork_premap=fork, **cfg)
    out = L(image, context, inv_depths, K, K, poses, mask=mask)
    out['loss'].backward()
    res = {
        'loss': out['loss'].detach().numpy(),
        'photometric_loss': out['metrics']['photometric_loss'].numpy(),
        'smoothness_loss': out['metrics'].get('smoothness_loss', torch.zeros(())).numpy(),
        'd_vec': torch.stack([v.grad for v in vecs]).numpy(),                # [C,B,6]
        'd_posemat': torch.stack([p.mat.grad for p in poses]).numpy(),       # [C,B,4,4]
        'd_invd_norm': np.array([d.grad.norm().item() for d in inv_depths if d.grad is not None]),
        'd_invd_sum': np.array([d.grad.sum().item() for d in inv_depths if d.grad is not None]),
        'image_probe': image[0, 0, 0, :3].numpy(),
    }
    n = L.n
    for i in range(n):
        w0 = L.kept['warped'][0][i].detach()
        res['warp0_s%d_stats' % i] = np.array([w0.mean().item(), w0.abs().sum().item(),
                                               (w0 == 0).double().mean().item()])
    if full:
        for i in range(n):
            res['d_invd_s%d' % i] = inv_depths[i].grad.numpy()
    if full == 2:
        for i in range(n):
            for j in range(C):
                res['warp%d_s%d' % (j, i)] = L.kept['warped'][j][i].detach().numpy()
            res['cands_s%d' % i] = torch.cat(L.kept['cands'][i], 1).detach().numpy()
    return res


def unit_vectors():
    """Small known-answer vectors of the reference's building blocks (fp64)."""
    g = torch.Generator('cpu').manual_seed(99)
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64)
    out = {}
    B, H, W = 2, 6, 9
    K = torch.tensor([[7.3, 0.02, 4.1], [0.01, 8.9, 2.7], [0.003, 0.002, 1.0]], dtype=torch.float64).repeat(B, 1, 1)
    K[1] *= 1.1
    vec = (r(B, 6) - 0.5) * torch.tensor([0.5, 0.5, 0.5, 0.2, 0.2, 0.2], dtype=torch.float64)
    pose = Pose.from_vec(vec, 'euler')
    depth = 1.0 + 5.0 * r(B, 1, H, W)
    img = r(B, 3, H, W)
    out.update(K=K.numpy(), vec=vec.numpy(), depth=depth.numpy(), img=img.numpy())
    out['euler2mat'] = euler2mat(vec[:, 3:]).numpy()
    out['pose_mat'] = pose.mat.numpy()
    out['pose_inv'] = pose.inverse().mat.numpy()
    out['pose_compose'] = (pose @ pose.inverse()).mat.numpy()
    out['invert_pose'] = invert_pose(pose.mat).numpy()
    cam = Camera(K=K.clone(), Tcw=pose)
    out['Kinv'] = cam.Kinv.numpy()
    out['K_scaled_half'] = Camera(K=K.clone()).scaled(0.5).K.numpy()
    out['K_scaled_xy'] = Camera(K=K.clone()).scaled(0.25, 0.5).K.numpy()
    out['scale_intrinsics'] = scale_intrinsics(K.clone(), 0.5, 0.25).numpy()
    out['reconstruct_c'] = Camera(K=K.clone(), Tcw=pose).reconstruct(depth, frame='c').numpy()
    out['reconstruct_w'] = Camera(K=K.clone(), Tcw=pose).reconstruct(depth, frame='w').numpy()
    X = Camera(K=K.clone(), Tcw=Pose.identity(B, dtype=torch.float64)).reconstruct(depth, frame='w')
    out['project_w'] = Camera(K=K.clone(), Tcw=pose).project(X, frame='w').numpy()
    out['project_c'] = Camera(K=K.clone(), Tcw=pose).project(X, frame='c').numpy()
    out['transform_points'] = pose.transform_points(X).numpy()
    for pm in ('zeros', 'border'):
        out['view_synthesis_' + pm] = view_synthesis(
            img, depth, Camera(K=K.clone(), Tcw=pose), Camera(K=K.clone(), Tcw=Pose.identity(B, dtype=torch.float64)),
            padding_mode=pm).numpy()
    img2 = r(B, 3, H, W)
    out['img2'] = img2.numpy()
    out['ssim'] = SSIM(img, img2).numpy()
    L = MultiViewPhotometricLoss()
    out['ssim_loss'] = L.SSIM(img, img2).numpy()
    invd = [0.05 + r(B, 1, H >> i, W >> i) for i in range(2)]
    imgs = match_scales(img, invd, 2)
    sx, sy = calc_smoothness(invd, imgs, 2)
    out.update(invd0=invd[0].numpy(), invd1=invd[1].numpy(), match_scale1=imgs[1].numpy(),
               smooth_x0=sx[0].numpy(), smooth_y0=sy[0].numpy(), smooth_x1=sx[1].numpy(), smooth_y1=sy[1].numpy(),
               invd_norm0=inv_depths_normalize(invd)[0].numpy(),
               gradient_x=gradient_x(img).numpy(), gradient_y=gradient_y(img).numpy(),
               inv2depth=inv2depth(invd[0]).numpy(), depth2inv=depth2inv(depth - 3.0).numpy(),
               interp_scales=interpolate_scales(invd, mode='nearest', align_corners=None)[1].numpy(),
               match_nearest=match_scales((img[:, :1] > 0.5).double(), invd, 2, mode='nearest', align_corners=None)[1].numpy())
    # grid_sample probe
    probe = torch.arange(12, dtype=torch.float64).view(1, 1, 3, 4)
    grid = torch.tensor([[1.5, 0.0], [-1.2, 0.3], [0.999, -0.999], [1.0, 1.0]], dtype=torch.float64).view(1, 1, 4, 2)
    for pm in ('zeros', 'border'):
        gq = grid.clone().requires_grad_()
        o = F.grid_sample(probe, gq, mode='bilinear', padding_mode=pm, align_corners=True)
        o.sum().backward()
        out['gs_out_' + pm] = o.detach().numpy()
        out['gs_dgrid_' + pm] = gq.grad.numpy()
    return out


def main():
    os.makedirs(os.path.join(ROOT, 'tests', 'golden'), exist_ok=True)
    save = lambda name, d: np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', name + '.npz'), **d)

    save('units', unit_vectors())

    # tiny cases with every per-pixel array kept (exact array parity of oracle and kernels)
    tiny = dict(B=2, H=24, W=40, C=2, kd=5, full=1)
    tiny2 = dict(tiny, full=2)
    save('tiny_U_zeros', run_case(pyramid=False, **tiny2))
    save('tiny_U_border', run_case(pyramid=False, padding_mode='border', **tiny2))
    save('tiny_P_zeros', run_case(pyramid=True, **tiny2))
    save('tiny_P_border', run_case(pyramid=True, padding_mode='border', **tiny2))
    save('tiny_U_mean_noautomask', run_case(pyramid=False, photometric_reduce_op='mean', automask_loss=False, **tiny))
    save('tiny_U_min_noautomask', run_case(pyramid=False, automask_loss=False, **tiny))
    save('tiny_U_clip', run_case(pyramid=False, clip_loss=0.5, **tiny))
    save('tiny_U_mask', run_case(pyramid=False, with_mask=True, **tiny))
    save('tiny_P_mask', run_case(pyramid=True, with_mask=True, **tiny))
    save('tiny_U_l1only', run_case(pyramid=False, ssim_loss_weight=0.0, **tiny))
    save('tiny_U_nosmooth', run_case(pyramid=False, smooth_loss_weight=0.0, **tiny))
    save('tiny_U_fork', run_case(pyramid=False, fork_premap=True, **tiny))
    save('tiny_U_classdefaults', run_case(pyramid=False, ssim_loss_weight=0.85, smooth_loss_weight=0.1,
                                           photometric_reduce_op='mean', clip_loss=0.5, automask_loss=False, **tiny))
    t4 = dict(tiny)
    t4['C'] = 4
    save('tiny_U_C4', run_case(pyramid=False, **t4))
    t3 = dict(tiny)
    t3.update(H=23, W=37)       # odd sizes, scalar tails
    save('tiny_U_odd', run_case(pyramid=False, **t3))
    save('tiny_U_scales2', run_case(pyramid=False, num_scales=2, **tiny))

    # SURVEY Appendix B cases: scalars, norms and pose gradients only
    for (B, H, W) in ((2, 64, 160), (1, 192, 640)):
        for pyr in (False, True):
            for pm in ('zeros', 'border'):
                name = 'B%d_%dx%d_C2_%s_%s' % (B, H, W, 'P' if pyr else 'U', pm)
                save(name, run_case(B, H, W, 2, pyr, padding_mode=pm))
                print('wrote', name)
    save('B2_64x160_C2_U_zeros_mean', run_case(2, 64, 160, 2, False, photometric_reduce_op='mean', automask_loss=False))
    save('B2_64x160_C2_U_zeros_clip', run_case(2, 64, 160, 2, False, clip_loss=0.5))


def time_port_vs_reference(B=2, H=192, W=640, C=2, iters=5, threads=8):
    """`--time`: frames/s of the reference's own classes and of the oracle port on the same fp32 inputs (build container
    only; bench.py's reference arm runs the port because /root/reference does not exist on the GPU box)."""
    import time
    import photometric_oracle as O
    torch.set_num_threads(threads)
    image, context, inv_depths, vecs, K = synth_S3(B, H, W, C, torch.float32)
    hp = dict(YAML_LOSS)

    def ref_step():
        d = [x.clone().requires_grad_() for x in inv_depths]
        v = [x.clone().requires_grad_() for x in vecs]
        out = PinholeReference(**hp)(image, context, d, K, K, [Pose.from_vec(x, 'euler') for x in v])
        out['loss'].backward()
        return float(out['loss'])

    def port_step():
        d = [x.clone().requires_grad_() for x in inv_depths]
        v = [x.clone().requires_grad_() for x in vecs]
        out = O.multiview_photometric_loss(image, context, d, K, K, [O.pose_vec_to_mat(x) for x in v], **hp)
        out['loss'].backward()
        return float(out['loss'])

    res = {}
    for name, fn in (('reference classes', ref_step), ('oracle port', port_step)):
        loss = fn()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        res[name] = (B * iters / (time.perf_counter() - t0), loss)
        print('%-18s %.2f frames/s  (loss %.9f, %d threads, B=%d %dx%d C=%d fp32)' % (name, res[name][0], loss, threads, B, H, W, C))
    print('port / reference = %.3f' % (res['oracle port'][0] / res['reference classes'][0]))


if __name__ == '__main__':
    if '--time' in sys.argv:
        time_port_vs_reference()
    else:
        main()
