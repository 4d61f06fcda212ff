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


@pytest.mark.parametrize('case', [dict(pm='zeros', mask=False, pyramid=False), dict(pm='border', mask=True, pyramid=False),
                                  dict(pm='zeros', mask=False, pyramid=True)])
def test_uint8_frames_are_bit_identical_to_to_tensor_floats(case):
    """img_dtype PSFM_U8: the kernels read image bytes and apply torchvision's to_tensor (byte / 255) themselves.
    The result must be bit-identical to feeding the float frames that to_tensor would have produced, for the loss
    and every gradient (byte masks hold 0 / 1)."""
    import torch.nn.functional as F
    import photometric_oracle as O
    from golden_cases import YAML_LOSS
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    B, H, W, C = 2, 64, 160, 2
    image, context, invd, vecs, K = O.synth_S3(B, H, W, C, torch.float32, pyramid=case['pyramid'], kd=7)
    dev = 'cuda:0'
    q = lambda t: (t * 255.0).round().clamp(0, 255).to(torch.uint8).to(dev)
    img8, ctx8 = q(image), [q(c) for c in context]
    # torchvision.transforms.functional.to_tensor, on the CPU like the reference's data pipeline: a true division
    # (on a CUDA tensor torch turns div-by-scalar into a multiplication by 1/255, which is off by an ulp for 126 bytes)
    to_tensor = lambda b: b.cpu().float().div(255).to(dev)
    gm = torch.Generator('cpu').manual_seed(11)
    mask = (torch.rand(B, 1, H, W, generator=gm) > 0.3) if case['mask'] else None
    hp = dict(YAML_LOSS, padding_mode=case['pm'])
    res = []
    for as_bytes in (True, False):
        d = [x.to(dev).clone().requires_grad_() for x in invd]
        v = [x.to(dev).clone().requires_grad_() for x in vecs]
        L = MultiViewPhotometricLoss(upsample_depth_maps=case['pyramid'], **hp)
        m = None if mask is None else (mask.to(torch.uint8) if as_bytes else mask.float()).to(dev)
        out = L(img8 if as_bytes else to_tensor(img8), ctx8 if as_bytes else [to_tensor(c) for c in ctx8], d,
                K.to(dev), K.to(dev), [Pose.from_vec(x, 'euler') for x in v], mask=m)
        out['loss'].backward()
        torch.cuda.synchronize()
        res.append((out['loss'].detach().clone(), [x.grad for x in d], [x.grad for x in v]))
    assert torch.equal(res[0][0], res[1][0])
    for a, b in zip(res[0][1] + res[0][2], res[1][1] + res[1][2]):
        assert torch.equal(a, b)
    # settings outside the sweep kernels are refused, not converted behind the caller's back
    with pytest.raises(ValueError, match='u8 images'):
        MultiViewPhotometricLoss(**dict(hp, photometric_reduce_op='mean', automask_loss=False))(
            img8, ctx8, [x.to(dev) for x in invd][:1] * 4 if case['pyramid'] else [x.to(dev) for x in invd], K.to(dev), K.to(dev),
            [Pose.from_vec(x.to(dev), 'euler') for x in vecs])
    with pytest.raises(ValueError, match='uint8 context'):
        MultiViewPhotometricLoss(**hp)(img8, [to_tensor(c) for c in ctx8], [x.to(dev) for x in invd][:1] * 4, K.to(dev),
                                       K.to(dev), [Pose.from_vec(x.to(dev), 'euler') for x in vecs])
