"""the fisheye (VADAS) camera model of the fork's `warp_ref_image`
(losses/multiview_photometric_loss.py:131-195, geometry/camera.py:194-394) inside the sweep kernels.

Fixtures: tests/golden/fisheye_*.npz, produced by the reference's own FisheyeCamera / loss classes
(oracle/gen_golden_fisheye.py).  fp64 kernels against the fixtures to rounding; fp32 kernels against the fp64
oracle at the tolerances of tests/test_loss_fp32_gpu.py (T2 loss, T3 gradients)."""
import pytest
import torch

import photometric_oracle as O
from golden_cases import CASES, YAML_LOSS, make_inputs
from psfm_testlib import cuda_run, oracle_run, rel_l2
from test_loss_gpu import check_against_golden

pytestmark = pytest.mark.gpu

FISHEYE = [n for n in CASES if n.startswith('fisheye_')]


@pytest.mark.parametrize('name', FISHEYE)
def test_fp64_kernels_match_reference_fisheye_fixtures(name):
    check_against_golden(name, cuda_run(name, torch.float64, 'fused'), 1e-10, 1e-8)


@pytest.mark.parametrize('name', ['fisheye_B2_64x160_C2_U_zeros', 'fisheye_B2_64x160_C2_U_border',
                                  'fisheye_B1_192x640_C2_U_zeros'])
def test_fp32_fisheye_kernels(name):
    inp64 = make_inputs(name, torch.float64)
    ref = oracle_run(name, torch.float64, inputs=inp64)
    to32 = lambda v: ([t.float() for t in v] if isinstance(v, list) else
                      ({k: t.float() for k, t in v.items()} if isinstance(v, dict) and k_is_tensor(v) else
                       (v.float() if torch.is_tensor(v) else v)))
    inp32 = {k: to32(v) for k, v in inp64.items()}
    res = cuda_run(name, torch.float32, 'fused', inputs=inp32)
    ref32 = oracle_run(name, torch.float32, inputs=inp32)
    for k in range(2):
        assert abs(res['loss3'][k].double() - ref['loss3'][k]) <= 1e-5 * abs(ref['loss3'][k]), k
    for s in range(len(ref['d_invd'])):
        g, g64, g32 = res['d_invd'][s].double(), ref['d_invd'][s], ref32['d_invd'][s].double()
        assert rel_l2(g, g64) <= max(3.0 * rel_l2(g32, g64), 1e-2), 'scale %d' % s
    for key in ('d_vec', 'd_mat'):
        a, b, c = torch.stack(res[key]).double(), torch.stack(ref[key]), torch.stack(ref32[key]).double()
        assert rel_l2(a, b) <= max(3.0 * rel_l2(c, b), 1e-2), key


def k_is_tensor(d):
    return all(torch.is_tensor(t) for t in d.values())


def test_fisheye_with_pyramid_upsampling_and_premap():
    """The fisheye instantiation is the run-time generic one: the fork's sigmoid pre-map and in-kernel nearest
    upsampling of the decoder pyramid compose with it (fp64, live oracle)."""
    import torch.nn.functional as F
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    B, H, W, C, n = 2, 32, 64, 2, 3
    hp = dict(YAML_LOSS, num_scales=n)
    image, context, pyr, vecs, _ = O.synth_S3(B, H, W, C, torch.float64, pyramid=True, kd=5, n=n)
    intr = O.synth_fisheye(B, H, W)
    lo, hi = 1.0 / 80.0, 1.0 / 0.5
    pyr = [((d - lo) / (hi - lo)).clamp(0.0, 1.0) for d in pyr]            # sigmoid-like raw inputs
    d64 = [d.clone().requires_grad_() for d in pyr]
    v64 = [v.clone().requires_grad_() for v in vecs]
    up = [d if s == 0 else F.interpolate(d, size=(H, W), mode='nearest') for s, d in enumerate(d64)]
    ref = O.multiview_photometric_loss(image, context, up, intr, intr, [O.pose_vec_to_mat(v) for v in v64],
                                       fork_premap=(0.5, 80.0), **hp)
    ref['loss'].backward()
    dev = 'cuda:0'
    dg = [d.to(dev).requires_grad_() for d in pyr]
    vg = [v.to(dev).requires_grad_() for v in vecs]
    L = MultiViewPhotometricLoss(upsample_depth_maps=True, sigmoid_premap=True, min_depth=0.5, max_depth=80.0, **hp)
    K = {k: v.to(dev) for k, v in intr.items()}
    out = L(image.to(dev), [c.to(dev) for c in context], dg, intrinsics=K, ref_intrinsics=K,
            poses=[Pose.from_vec(v, 'euler') for v in vg])
    out['loss'].backward()
    torch.cuda.synchronize()
    assert abs(out['loss'].item() - ref['loss'].item()) <= 1e-10 * abs(ref['loss'].item())
    for s in range(n):
        assert rel_l2(dg[s].grad.cpu(), d64[s].grad) <= 1e-8, 'scale %d' % s
    for a, b in zip(vg, v64):
        assert rel_l2(a.grad.cpu(), b.grad) <= 1e-8


def test_fisheye_rejections():
    """Settings the sweep kernels do not cover are refused loudly (no fallback)."""
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    inp = make_inputs('fisheye_tiny_U_zeros', torch.float64)
    dev = 'cuda:0'
    K = {k: v.to(dev) for k, v in inp['K'].items()}
    args = (inp['image'].to(dev), [c.to(dev) for c in inp['context']], [d.to(dev) for d in inp['inv_depths']])
    poses = [Pose.from_vec(v.to(dev), 'euler') for v in inp['vecs']]
    hp = dict(inp['hp'], photometric_reduce_op='mean', automask_loss=False)
    with pytest.raises(ValueError, match='fisheye'):
        MultiViewPhotometricLoss(**hp)(*args, K, K, poses)
    with pytest.raises(NotImplementedError):
        MultiViewPhotometricLoss(kernel_mode='split', **inp['hp'])(*args, K, K, poses)
    with pytest.raises(ValueError):
        MultiViewPhotometricLoss(**inp['hp'])(*args, K, torch.eye(3, device=dev, dtype=torch.float64), poses)


# ---- building blocks: FisheyeCamera / grid_sample / view_synthesis (geometry/camera.py:194-394) ----------------

def _units():
    from golden_cases import load_golden
    u = load_golden('fisheye_units')
    t = lambda k: torch.from_numpy(u[k]).to('cuda:0')
    return u, t, {n: t('intr_' + n) for n in ('k', 's', 'div', 'ux', 'uy')}


def test_fisheye_camera_methods_match_reference():
    """FisheyeCamera.reconstruct / project (both frames) and view_synthesis against outputs of the reference's
    own class (fisheye_units.npz), fp64."""
    import numpy as np
    from packnet_sfm.geometry.camera import FisheyeCamera
    from packnet_sfm.geometry.camera_utils import view_synthesis
    from packnet_sfm.geometry.pose import Pose
    u, t, intr = _units()
    eq = lambda a, key: np.testing.assert_allclose(a.cpu().numpy(), u[key], rtol=1e-12, atol=1e-13)
    B, _, H, W = u['depth'].shape
    pose = Pose(t('pose_mat'))
    cam = FisheyeCamera(intrinsics=intr, image_size=(H, W))
    X = cam.reconstruct(t('depth'), frame='c')
    eq(X, 'reconstruct_c')
    eq(cam.project(pose @ X, frame='c'), 'project_c')
    eq(cam.project(t('edge_points'), frame='c'), 'project_edge')              # Z <= 0, far off-axis points
    # world frame: what the reference's frame='w' branches spell out (they raise there)
    ref_cam = FisheyeCamera(intrinsics=intr, Tcw=pose, image_size=(H, W))
    eq(ref_cam.project(cam.reconstruct(t('depth'), frame='w'), frame='w'), 'project_c')
    # [B,3,N] point clouds are normalised with image_size (:325-333)
    eq(cam.project((pose @ X).reshape(B, 3, -1), frame='c').view(B, H, W, 2), 'project_c')
    with pytest.raises(ValueError, match='image_size'):
        FisheyeCamera(intrinsics=intr).project(X.reshape(B, 3, -1), frame='c')
    for pm in ('zeros', 'border'):
        eq(view_synthesis(t('img'), t('depth'), ref_cam, cam, padding_mode=pm), 'view_synthesis_' + pm)
    assert len(cam) == B and cam.to('cuda:0') is cam


@pytest.mark.parametrize('pm', ['zeros', 'border'])
def test_fisheye_view_synthesis_gradients(pm):
    """d/d depth and d/d pose of the kernel chain reconstruct -> transform -> project -> grid_sample against
    autograd through the oracle (fp64)."""
    from packnet_sfm.geometry.camera import FisheyeCamera
    from packnet_sfm.geometry.camera_utils import view_synthesis
    from packnet_sfm.geometry.pose import Pose
    B, H, W = 2, 20, 36
    image, context, invd, vecs, _ = O.synth_S3(B, H, W, 1, torch.float64, kd=5)
    intr = O.synth_fisheye(B, H, W)
    gen = torch.Generator('cpu').manual_seed(5)
    wgt = torch.rand(B, 3, H, W, generator=gen, dtype=torch.float64)
    d64 = O.inv2depth(invd[0]).clone().requires_grad_()
    v64 = vecs[0].clone().requires_grad_()
    ref = O.fisheye_view_synthesis(context[0], d64, intr, O.pose_vec_to_mat(v64), intr, pm)
    (ref * wgt).sum().backward()
    dev = 'cuda:0'
    dg = O.inv2depth(invd[0]).to(dev).requires_grad_()
    vg = vecs[0].to(dev).requires_grad_()
    K = {k: v.to(dev) for k, v in intr.items()}
    out = view_synthesis(context[0].to(dev), dg, FisheyeCamera(K, Tcw=Pose.from_vec(vg, 'euler'), image_size=(H, W)),
                         FisheyeCamera(K, image_size=(H, W)), padding_mode=pm)
    (out * wgt.to(dev)).sum().backward()
    torch.cuda.synchronize()
    assert (out.detach().cpu() - ref.detach()).abs().max().item() <= 1e-12
    assert rel_l2(dg.grad.cpu(), d64.grad) <= 1e-10
    assert rel_l2(vg.grad.cpu(), v64.grad) <= 1e-10


@pytest.mark.parametrize('name', ['fisheye_tiny_U_zeros', 'fisheye_tiny_U_border', 'fisheye_small_P_zeros'])
def test_warp_ref_image_with_fisheye_intrinsics(name):
    """MultiViewPhotometricLoss.warp_ref_image with the fork's dict intrinsics (:131-195) against the warped
    images the reference produced."""
    import numpy as np
    from golden_cases import load_golden
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    gold = load_golden(name)
    inp = make_inputs(name, torch.float64)
    dev = 'cuda:0'
    L = MultiViewPhotometricLoss(**inp['hp'])
    K = {k: v.to(dev) for k, v in inp['K'].items()}
    H, W = inp['image'].shape[2:]
    for j, (c, v) in enumerate(zip(inp['context'], inp['vecs'])):
        warped = L.warp_ref_image([d.to(dev) for d in inp['inv_depths']], c.to(dev), K, K,
                                  Pose.from_vec(v.to(dev), 'euler'), (H, W))
        for i, w in enumerate(warped):
            np.testing.assert_allclose(w.cpu().numpy(), gold['warp%d_s%d' % (j, i)], rtol=1e-10, atol=1e-12)
