"""Size-independent properties of the CUDA path at BASELINE.json's full sizes,
where the CPU oracle would take minutes: batch-shard additivity (what the multi-GPU path relies on),
agreement of the two independent kernel families, a directional finite-difference check of the
gradients, bit-wise determinism, and the identity-pose round trip of view synthesis.
Inputs are the synthetic frames of bench.py (recipe S3 shapes); everything goes through the public
Python API -> ctypes -> C ABI."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))

pytestmark = pytest.mark.gpu

HP = dict(num_scales=4, ssim_loss_weight=0.85, smooth_loss_weight=0.001, photometric_reduce_op='min', clip_loss=0.0,
          automask_loss=True)
# name: per-GPU shapes of BASELINE.json configs[1..4] (batch trimmed where only the shape matters)
SHAPES = {
    'cfg2_kitti_B8': dict(B=8, H=192, W=640, C=2, n=4, img='fp32', padding='zeros'),
    'cfg3_ncdb_bf16_B4': dict(B=4, H=384, W=640, C=2, n=4, img='bf16', padding='zeros'),
    'cfg4_ddad_border_B2': dict(B=2, H=1216, W=1936, C=2, n=4, img='fp32', padding='border'),
    'cfg5_c4_B4': dict(B=4, H=384, W=640, C=4, n=4, img='fp32', padding='zeros'),
}


def _inputs(cfg, real=torch.float32, seed=1234):
    import bench
    h = bench.synth_batch(cfg, None, seed)
    dev = torch.device('cuda:0')
    img_dt = torch.bfloat16 if cfg['img'] == 'bf16' else real
    if real == torch.float64:
        img_dt = torch.float64
    from packnet_sfm.geometry.pose import Pose
    # pose matrices are formed ONCE for the whole batch (torch's batched matmul may round differently for
    # different batch sizes, which would leak into shard-vs-batch comparisons)
    mats = [Pose.from_vec(v.to(dev).to(real), 'euler').mat.detach() for v in h['vecs']]
    return dict(image=h['image'].to(dev).to(img_dt), context=[c.to(dev).to(img_dt) for c in h['context']],
                inv_depths=[d.to(dev).to(real) for d in h['inv_depths']], mats=mats, K=h['K'].to(dev).to(real))


def _run(cfg, inp, sl=slice(None), mode=None, inv_depths=None):
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    invd = [d[sl].detach().clone().requires_grad_() for d in (inv_depths or inp['inv_depths'])]
    vecs = [m[sl].detach().clone().requires_grad_() for m in inp['mats']]
    L = MultiViewPhotometricLoss(padding_mode=cfg['padding'], kernel_mode=mode, **HP)
    out = L(inp['image'][sl], [c[sl] for c in inp['context']], invd, inp['K'][sl], inp['K'][sl],
            [Pose(m) for m in vecs])
    out['loss'].backward()
    torch.cuda.synchronize()
    return out['loss'].detach().double().item(), [d.grad for d in invd], [v.grad for v in vecs]


def _rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300)).item()


@pytest.mark.parametrize('pin_rows', [True, False])
@pytest.mark.parametrize('name', list(SHAPES))
def test_batch_shards_add_up(name, pin_rows, monkeypatch):
    """loss(batch) = mean of the shard losses, d/d inv_depth of a half-batch shard = 2 x the full-batch one
    (every reduction is a mean over batch x pixels; the factor 2 is exact in binary floating point).
    The library picks the rows per sweep unit from the batch size; a row's position in the 3-step unrolled
    sweep changes fp32 rounding (different FMA contraction per unrolled copy), so the comparison is exact to
    1e-6 with the unit height pinned and at fp32 noise level (a few flipped pixels) otherwise."""
    cfg = SHAPES[name]
    if pin_rows:
        monkeypatch.setenv('PSFM_SWEEP_CHUNK', '48')
    tol_g, tol_v = (1e-6, 1e-5) if pin_rows else (5e-3, 5e-3)
    inp = _inputs(cfg)
    B = cfg['B']
    full, g_full, v_full = _run(cfg, inp)
    parts = [_run(cfg, inp, slice(k * B // 2, (k + 1) * B // 2)) for k in range(2)]
    assert abs(0.5 * (parts[0][0] + parts[1][0]) - full) <= 1e-6 * abs(full)
    for k in range(2):
        sl = slice(k * B // 2, (k + 1) * B // 2)
        for s in range(cfg['n']):
            assert _rel(parts[k][1][s], 2.0 * g_full[s][sl]) <= tol_g
        for j in range(cfg['C']):
            assert _rel(parts[k][2][j], 2.0 * v_full[j][sl]) <= tol_v


@pytest.mark.parametrize('name', ['cfg2_kitti_B8', 'cfg4_ddad_border_B2'])
def test_two_kernel_families_agree_fp64(name):
    """The sweep kernels (fused) and the first-generation tile kernels (split) are independent
    implementations; in double precision they agree to rounding at full size."""
    cfg = dict(SHAPES[name])
    cfg['B'] = 2
    inp = _inputs(cfg, torch.float64)
    lf, gf, vf = _run(cfg, inp, mode='fused')
    ls, gs, vs = _run(cfg, inp, mode='split')
    assert abs(lf - ls) <= 1e-11 * abs(ls)
    for a, b in zip(gf, gs):
        assert _rel(a, b) <= 1e-9
    for a, b in zip(vf, vs):
        assert _rel(a, b) <= 1e-9


def test_directional_derivative_fp64_kitti_size():
    """<grad, delta> against a central difference of the loss along a smooth random direction (fp64, config-2
    shape).  The loss has kinks (min, abs, clamp, bilinear cell edges): measured agreement 8e-4, bound 5e-3."""
    cfg = dict(SHAPES['cfg2_kitti_B8'])
    cfg['B'] = 2
    inp = _inputs(cfg, torch.float64)
    _, g, _ = _run(cfg, inp)
    gen = torch.Generator(device='cuda').manual_seed(7)
    delta = [torch.nn.functional.avg_pool2d(torch.randn(d.shape, generator=gen, device=d.device, dtype=d.dtype), 9, 1, 4)
             for d in inp['inv_depths']]
    eps = 1e-7
    lp = _run(cfg, inp, inv_depths=[d + eps * x for d, x in zip(inp['inv_depths'], delta)])[0]
    lm = _run(cfg, inp, inv_depths=[d - eps * x for d, x in zip(inp['inv_depths'], delta)])[0]
    fd = (lp - lm) / (2 * eps)
    an = sum((a * x).sum().item() for a, x in zip(g, delta))
    assert abs(fd - an) <= 5e-3 * max(abs(an), abs(fd)), (fd, an)


@pytest.mark.parametrize('name', ['cfg2_kitti_B8', 'cfg5_c4_B4'])
def test_bitwise_determinism(name):
    """Fixed-order partial reductions: two runs give identical bits for the loss and every gradient."""
    cfg = SHAPES[name]
    inp = _inputs(cfg)
    a, b = _run(cfg, inp), _run(cfg, inp)
    assert a[0] == b[0]
    for x, y in zip(a[1] + a[2], b[1] + b[2]):
        assert torch.equal(x, y)


@pytest.mark.parametrize('pm', ['zeros', 'border'])
def test_identity_pose_view_synthesis_is_identity_full_size(pm):
    """view_synthesis with the identity pose and the same intrinsics returns the image itself, whatever the depth
    (geometry/camera_utils.py:27-59), at the DDAD frame size."""
    from packnet_sfm.geometry.camera import Camera
    from packnet_sfm.geometry.camera_utils import view_synthesis
    from packnet_sfm.geometry.pose import Pose
    cfg = dict(SHAPES['cfg4_ddad_border_B2'], B=1)
    inp = _inputs(cfg)
    depth = 1.0 / inp['inv_depths'][0]
    cam = Camera(inp['K'])
    ref_cam = Camera(inp['K'], Tcw=Pose.identity(1, device=depth.device))
    out = view_synthesis(inp['context'][0], depth, ref_cam, cam, padding_mode=pm)
    torch.cuda.synchronize()
    assert (out - inp['context'][0]).abs().max().item() <= 2e-3      # fp32 sample position error x image gradient
    inner = (out - inp['context'][0])[..., 2:-2, 2:-2].abs().mean().item()
    assert inner <= 1e-4
