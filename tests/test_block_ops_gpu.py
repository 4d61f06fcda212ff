"""GPU parity of the building-block operators (Camera.reconstruct/project, Pose.transform_points,
SSIM, calc_photometric_loss, reduce_photometric_loss, calc_smoothness, view_synthesis, match_scales)
against the CPU oracle and the reference-produced `units` fixture.  fp64 kernels: values 1e-11,
gradients 1e-9 (relative L2)."""
import numpy as np
import pytest
import torch

import photometric_oracle as O
from golden_cases import load_golden
from psfm_testlib import rel_l2

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _units():
    u = load_golden('units')
    t = lambda k: torch.from_numpy(u[k])
    return u, t


def _grad_pair(f_ref, f_gpu, inputs, seed=0):
    """Runs both on copies of `inputs` that require grad, back-propagates one random cotangent
    through every output, returns (outs_ref, grads_ref, outs_gpu, grads_gpu)."""
    g = torch.Generator().manual_seed(seed)
    a = [x.detach().clone().requires_grad_() for x in inputs]
    b = [x.detach().clone().to(DEV).requires_grad_() for x in inputs]
    ra, rb = f_ref(*a), f_gpu(*b)
    ra = ra if isinstance(ra, (list, tuple)) else [ra]
    rb = rb if isinstance(rb, (list, tuple)) else [rb]
    cot = [torch.randn(o.shape, generator=g, dtype=o.dtype) for o in ra]
    sum((o * c).sum() for o, c in zip(ra, cot)).backward()
    sum((o * c.to(DEV)).sum() for o, c in zip(rb, cot)).backward()
    torch.cuda.synchronize()
    return ra, [x.grad for x in a], [o.detach().cpu() for o in rb], [x.grad.cpu() for x in b]


def _check(f_ref, f_gpu, inputs, vtol=1e-11, gtol=1e-9):
    ra, ga, rb, gb = _grad_pair(f_ref, f_gpu, inputs)
    for x, y in zip(ra, rb):
        assert x.shape == y.shape
        assert rel_l2(y, x.detach()) <= vtol
    for x, y in zip(ga, gb):
        assert rel_l2(y, x) <= gtol


def test_reconstruct_project_transform_match_reference_units():
    from packnet_sfm.geometry.camera import Camera
    from packnet_sfm.geometry.pose import Pose
    u, t = _units()
    K, vec, depth = t('K').to(DEV), t('vec').to(DEV), t('depth').to(DEV)
    T = Pose.from_vec(vec, 'euler')
    eq = lambda a, k: np.testing.assert_allclose(a.cpu().numpy(), u[k], rtol=1e-11, atol=1e-13)
    eq(Camera(K).reconstruct(depth, frame='c'), 'reconstruct_c')
    eq(Camera(K, Tcw=T).reconstruct(depth, frame='w'), 'reconstruct_w')
    X = Camera(K).reconstruct(depth, frame='c')
    eq(Camera(K, Tcw=T).project(X, frame='w'), 'project_w')
    eq(Camera(K).project(X, frame='c'), 'project_c')
    eq(T @ X, 'transform_points')
    eq(T.transform_pairs(X), 'transform_points')


def test_geometry_gradients_match_oracle():
    from packnet_sfm import ops
    u, t = _units()
    K, T, depth = t('K'), O.pose_vec_to_mat(t('vec')), t('depth')
    _check(lambda d: O.reconstruct(d, K), lambda d: ops.reconstruct_op(d, K.to(DEV)), [depth])
    _check(lambda d: O.reconstruct(d, K, T), lambda d: ops.reconstruct_op(d, K.to(DEV), T.to(DEV)), [depth])
    X = O.reconstruct(depth, K)
    _check(lambda x, m: O.project(x, K, m), lambda x, m: ops.project_op(x, K.to(DEV), m), [X, T])
    _check(lambda x: O.project(x, K), lambda x: ops.project_op(x, K.to(DEV)), [X])
    _check(lambda m, x: O.transform_points(m, x), lambda m, x: ops.transform_points_op(m, x), [T, X])


def test_ssim_and_photometric_map():
    from packnet_sfm.losses.multiview_photometric_loss import SSIM, MultiViewPhotometricLoss
    u, t = _units()
    img, img2 = t('img'), t('img2')
    np.testing.assert_allclose(SSIM(img.to(DEV), img2.to(DEV)).cpu().numpy(), u['ssim'], rtol=1e-11, atol=1e-13)
    L = MultiViewPhotometricLoss()
    np.testing.assert_allclose(L.SSIM(img.to(DEV), img2.to(DEV)).cpu().numpy(), u['ssim_loss'], rtol=1e-11, atol=1e-13)
    _check(lambda a, b: O.ssim_index(a, b), lambda a, b: SSIM(a, b), [img, img2])
    _check(lambda a, b: O.ssim_distance(a, b), lambda a, b: L.SSIM(a, b), [img, img2])
    for alpha, clip in ((0.85, 0.0), (0.0, 0.0), (0.85, 0.5)):
        L = MultiViewPhotometricLoss(num_scales=1, ssim_loss_weight=alpha, clip_loss=clip)
        _check(lambda a: O.photometric_map(a, img2, alpha, 1e-4, 9e-4, clip),
               lambda a: L.calc_photometric_loss([a], [img2.to(DEV)])[0], [img])


@pytest.mark.parametrize('op', ['min', 'mean'])
def test_reduce_photometric_loss(op):
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    g = torch.Generator().manual_seed(3)
    cands = [torch.rand(2, 1, 9, 13, generator=g, dtype=torch.float64) for _ in range(4)]
    cands[2][0, 0, 3, 4] = cands[0][0, 0, 3, 4]            # a tie: the first index takes the gradient
    L = MultiViewPhotometricLoss(num_scales=1, photometric_reduce_op=op)
    _check(lambda *c: O.reduce_candidates(list(c), op), lambda *c: L.reduce_photometric_loss([list(c)]), cands)


def test_smoothness_matches_reference_units_and_oracle():
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    from packnet_sfm.utils.depth import calc_smoothness, inv_depths_normalize, inv_depths_smoothness, inv2depth
    from packnet_sfm.utils.image import match_scales
    u, t = _units()
    img = t('img').to(DEV)
    invd = [t('invd0').to(DEV), t('invd1').to(DEV)]
    imgs = match_scales(img, invd, 2)
    eq = lambda a, k: np.testing.assert_allclose(a.cpu().numpy(), u[k], rtol=1e-11, atol=1e-13)
    eq(imgs[1], 'match_scale1')
    sx, sy = calc_smoothness(invd, imgs, 2)
    eq(sx[0], 'smooth_x0'); eq(sy[0], 'smooth_y0'); eq(sx[1], 'smooth_x1'); eq(sy[1], 'smooth_y1')
    assert inv_depths_smoothness is calc_smoothness
    eq(inv2depth(invd[0]), 'inv2depth')
    n0 = inv_depths_normalize(invd)[0]
    np.testing.assert_allclose(n0.mean((2, 3)).cpu().numpy(), 1.0, rtol=1e-12)
    cpu_imgs = [i.cpu() for i in imgs]
    L = MultiViewPhotometricLoss(num_scales=2, smooth_loss_weight=0.1)
    _check(lambda a, b: O.smoothness_loss([a, b], cpu_imgs, 2, 0.1),
           lambda a, b: L.calc_smoothness_loss([a, b], imgs), [t('invd0'), t('invd1')])


@pytest.mark.parametrize('pm', ['zeros', 'border'])
def test_view_synthesis_matches_reference_units_and_oracle(pm):
    from packnet_sfm.geometry.camera import Camera
    from packnet_sfm.geometry.camera_utils import view_synthesis
    from packnet_sfm.geometry.pose import Pose
    u, t = _units()
    K, vec, depth, img = t('K'), t('vec'), t('depth'), t('img')
    out = view_synthesis(img.to(DEV), depth.to(DEV), Camera(K.to(DEV), Tcw=Pose.from_vec(vec.to(DEV), 'euler')),
                         Camera(K.to(DEV)), padding_mode=pm)
    np.testing.assert_allclose(out.cpu().numpy(), u['view_synthesis_' + pm], rtol=1e-10, atol=1e-12)
    _check(lambda d, m: O.view_synthesis(img, d, K, m, K, pm),
           lambda d, m: view_synthesis(img.to(DEV), d, Camera(K.to(DEV), Tcw=Pose(m)), Camera(K.to(DEV)), padding_mode=pm),
           [depth, O.pose_vec_to_mat(vec)], vtol=1e-10, gtol=1e-8)


def test_warp_ref_image_building_block_path_equals_fused_loss():
    """The reference-style composition of the building blocks reproduces the fused kernel's loss."""
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.utils.image import match_scales
    from golden_cases import make_inputs
    inp = make_inputs('tiny_U_zeros', torch.float64)
    L = MultiViewPhotometricLoss(**inp['hp'])
    mv = lambda x: x.to(DEV)
    image, context, invd, K = mv(inp['image']), [mv(c) for c in inp['context']], [mv(d) for d in inp['inv_depths']], mv(inp['K'])
    poses = [Pose.from_vec(mv(v), 'euler') for v in inp['vecs']]
    fused = L(image, context, invd, K, K, poses)['loss'].item()
    n = L.n
    images = match_scales(image, invd, n)
    losses = [[] for _ in range(n)]
    for ref, pose in zip(context, poses):
        warped = L.warp_ref_image(invd, ref, K, K, pose)
        pl = L.calc_photometric_loss(warped, images)
        ul = L.calc_photometric_loss(match_scales(ref, invd, n), images)
        for i in range(n):
            losses[i] += [pl[i], ul[i]]
    total = L.reduce_photometric_loss(losses) + L.calc_smoothness_loss(invd, images)
    assert abs(total.item() - fused) <= 1e-11 * abs(fused)


def test_cpu_tensors_are_refused():
    from packnet_sfm import ops
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ops.ssim_op(torch.rand(1, 3, 8, 8), torch.rand(1, 3, 8, 8))


@pytest.mark.parametrize('dtype', [torch.float64, torch.float32])
def test_pose_from_vec_kernel_matches_oracle(dtype):
    """Pose.from_vec(vec, 'euler') (geometry/pose.py:39-46, pose_utils.py:8-51) as one kernel each way:
    matrix and d/d vec against the oracle's torch composition Rx Ry Rz, for batch sizes 1 and 5 (the
    kernel's rounding does not depend on the batch size, torch's batched matmul does)."""
    from packnet_sfm.geometry.pose import Pose
    g = torch.Generator().manual_seed(3)
    vec = torch.randn(5, 6, generator=g, dtype=torch.float64) * torch.tensor([1., 1., 1., 0.5, 0.5, 0.5], dtype=torch.float64)
    w = torch.randn(5, 4, 4, generator=g, dtype=torch.float64)
    v_ref = vec.clone().requires_grad_()
    m_ref = O.pose_vec_to_mat(v_ref)
    (m_ref * w).sum().backward()
    tol = 1e-12 if dtype == torch.float64 else 2e-6
    v = vec.to(DEV).to(dtype).requires_grad_()
    pose = Pose.from_vec(v, 'euler')
    (pose.mat * w.to(DEV).to(dtype)).sum().backward()
    assert (pose.mat.detach().cpu().double() - m_ref.detach()).abs().max().item() <= tol
    assert (v.grad.cpu().double() - v_ref.grad).abs().max().item() <= 10 * tol
    one = Pose.from_vec(vec[2:3].to(DEV).to(dtype), 'euler').mat
    assert torch.equal(one[0], pose.mat.detach()[2])
