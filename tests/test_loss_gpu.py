"""GPU parity of the fused photometric loss against the oracle / golden fixtures.

T1 (math parity): the fp64 kernel instantiation vs the fp64 oracle and the reference-produced
    fixtures: loss rel <= 1e-10, gradients rel-L2 <= 1e-8.
T2 (north-star tolerances): fp32 kernels vs the fp64 oracle: loss rel <= 1e-5; gradients judged
    norm-wise outside the flip set (tests/test_loss_fp32_gpu.py).
"""
import numpy as np
import pytest
import torch

from golden_cases import CASES, load_golden
from psfm_testlib import cuda_run, oracle_run, rel_l2

pytestmark = pytest.mark.gpu

TINY = [n for n in CASES if n.startswith('tiny')]
MID = [n for n in CASES if n.startswith('B2_')]
BIG = [n for n in CASES if n.startswith('B1_')]


def check_against_golden(name, res, loss_tol, grad_tol):
    gold = load_golden(name)
    n = len(gold['d_invd_norm'])
    np.testing.assert_allclose(res['loss3'][0].item(), gold['loss'][0], rtol=loss_tol)
    np.testing.assert_allclose(res['loss3'][1].item(), gold['photometric_loss'], rtol=loss_tol)
    np.testing.assert_allclose(res['loss3'][2].item(), gold['smoothness_loss'], rtol=loss_tol, atol=1e-300)
    d_mat = torch.stack(res['d_mat']).double()
    assert rel_l2(d_mat, torch.from_numpy(gold['d_posemat'])) <= grad_tol
    d_vec = torch.stack(res['d_vec']).double()
    assert rel_l2(d_vec, torch.from_numpy(gold['d_vec'])) <= grad_tol
    for i in range(n):
        if 'd_invd_s%d' % i in gold:
            assert rel_l2(res['d_invd'][i], torch.from_numpy(gold['d_invd_s%d' % i])) <= grad_tol, 'scale %d' % i
        np.testing.assert_allclose(res['d_invd'][i].double().norm().item(), gold['d_invd_norm'][i], rtol=max(grad_tol, 1e-9))


@pytest.mark.parametrize('mode', ['split', 'fused'])
@pytest.mark.parametrize('name', TINY + MID)
def test_fp64_kernels_match_reference_fixtures(name, mode):
    check_against_golden(name, cuda_run(name, torch.float64, mode), 1e-10, 1e-8)


@pytest.mark.parametrize('name', BIG)
def test_fp64_kernels_match_reference_fixtures_kitti(name):
    check_against_golden(name, cuda_run(name, torch.float64, 'fused'), 1e-10, 1e-8)


@pytest.mark.parametrize('name', ['tiny_U_zeros', 'tiny_P_border', 'tiny_U_mask', 'tiny_U_fork'])
def test_fp64_kernels_match_live_oracle(name):
    ref, res = oracle_run(name), cuda_run(name, torch.float64, 'split')
    assert abs(res['loss3'][0] - ref['loss3'][0]) / ref['loss3'][0] <= 1e-10
    for a, b in zip(res['d_invd'], ref['d_invd']):
        assert rel_l2(a, b) <= 1e-8
    for a, b in zip(res['d_mat'], ref['d_mat']):
        assert rel_l2(a, b) <= 1e-8


@pytest.mark.parametrize('case', [
    dict(B=2, H=64, W=160, C=2, pm='zeros', dtype=torch.float64),
    dict(B=2, H=24, W=40, C=2, pm='border', dtype=torch.float64, kd=5, n=3),
    dict(B=1, H=192, W=640, C=2, pm='zeros', dtype=torch.float64),
    dict(B=2, H=64, W=160, C=2, pm='zeros', dtype=torch.float32),
    dict(B=2, H=30, W=50, C=2, pm='zeros', dtype=torch.float64, kd=3, n=2, sizes=[(30, 50), (10, 25)]),   # factors 3 x 2
])
def test_in_kernel_upsampling_of_the_depth_pyramid(case):
    """`upsample_depth_maps=True`: the loss takes the decoder's native pyramid [B,1,H>>s,W>>s] and does what the
    reference's MODEL does before its loss -- nearest upsampling to full resolution, models/SfmModel.py:86-88 ->
    models/model_utils.py:152-176 -> F.interpolate(mode='nearest') -- inside the kernels, adjoint included.
    Oracle: F.interpolate + the U-mode oracle loss, gradients taken w.r.t. the pyramid."""
    import torch.nn.functional as F
    import photometric_oracle as O
    from golden_cases import YAML_LOSS
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    n = case.get('n', 4)
    hp = dict(YAML_LOSS, padding_mode=case['pm'], num_scales=n)
    image, context, pyr, vecs, K = O.synth_S3(case['B'], case['H'], case['W'], case['C'], torch.float64, pyramid=True,
                                              kd=case.get('kd', 15 if case['H'] >= 128 else 7), n=n)
    if 'sizes' in case:     # arbitrary (non power-of-two) map sizes: the general pooling kernel
        pyr = [F.interpolate(pyr[0], size=sz, mode='bilinear', align_corners=False) for sz in case['sizes']]
    d64 = [d.clone().requires_grad_() for d in pyr]
    v64 = [v.clone().requires_grad_() for v in vecs]
    up = [d if s == 0 else F.interpolate(d, size=(case['H'], case['W']), mode='nearest') for s, d in enumerate(d64)]
    ref = O.multiview_photometric_loss(image, context, up, K, K, [O.pose_vec_to_mat(v) for v in v64], **hp)
    ref['loss'].backward()
    dt, dev = case['dtype'], 'cuda:0'
    dg = [d.to(dev).to(dt).requires_grad_() for d in pyr]
    vg = [v.to(dev).to(dt).requires_grad_() for v in vecs]
    L = MultiViewPhotometricLoss(upsample_depth_maps=True, **hp)
    out = L(image.to(dev).to(dt), [c.to(dev).to(dt) for c in context], dg, K.to(dev).to(dt), K.to(dev).to(dt),
            [Pose.from_vec(v, 'euler') for v in vg])
    out['loss'].backward()
    torch.cuda.synchronize()
    f64 = dt == torch.float64
    assert abs(out['loss'].item() - ref['loss'].item()) <= (1e-10 if f64 else 1e-5) * abs(ref['loss'].item())
    for s in range(n):
        assert dg[s].grad.shape == pyr[s].shape
        # fp32: pooled sums average the flip-pixel noise of DESIGN.md T2/T3; bound as in T3
        assert rel_l2(dg[s].grad.cpu(), d64[s].grad) <= (1e-8 if f64 else 1e-2), 'scale %d' % s
    for a, b in zip(vg, v64):
        assert rel_l2(a.grad.cpu(), b.grad) <= (1e-8 if f64 else 1e-2)
