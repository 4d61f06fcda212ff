"""SupervisedLoss with the l1 / mse / silog loss functions
(reference losses/supervised_loss.py:67-92, :241-437) through psfm_supervised_loss[_backward].
fp64 kernels against fixtures produced by the reference's own class; fp32 kernels against the fp64 oracle."""
import numpy as np
import pytest
import torch

import supervised_oracle as S
from golden_cases import load_golden
from supervised_cases import CASES, EXTRA, key_of, make_inputs

pytestmark = pytest.mark.gpu


def _run(case, dtype):
    from packnet_sfm.losses.supervised_loss import SupervisedLoss
    preds, gt, masks = make_inputs(case, dtype)
    dev = 'cuda:0'
    preds = [p.to(dev).requires_grad_() for p in preds]
    L = SupervisedLoss(supervised_method=case[0], supervised_num_scales=case[1], **EXTRA.get(case[0], {}))
    out = L(preds, gt.to(dev), masks=None if masks is None else [m.to(dev) for m in masks])
    out['loss'].backward()
    torch.cuda.synchronize()
    return out, preds


@pytest.mark.parametrize('case', CASES, ids=[key_of(c) for c in CASES])
def test_fp64_kernels_match_reference_fixtures(case):
    gold = load_golden('supervised')
    out, preds = _run(case, torch.float64)
    k = key_of(case)
    assert out['loss'].shape == (1,)
    np.testing.assert_allclose(out['loss'].item(), gold[k + '/loss'][0], rtol=1e-12)
    for i, p in enumerate(preds):
        np.testing.assert_allclose(p.grad.cpu().numpy(), gold[k + '/d_pred%d' % i], rtol=1e-10, atol=1e-16)
    if case[0].startswith('sparse'):
        m = out['metrics']
        np.testing.assert_allclose([m['s%d/valid_px' % i].item() for i in range(case[1])], gold[k + '/valid_px'])
        np.testing.assert_allclose([m['s%d/loss' % i].item() for i in range(case[1])], gold[k + '/scale_loss'], rtol=1e-12)
        assert abs(m['supervised_loss'].item() - out['loss'].item()) == 0
    if case[0].endswith('ssi-silog'):
        for name in ('ssi_component', 'silog_component', 'ssi_mean', 'ssi_var', 'silog1', 'silog2'):
            for i in range(case[1]):
                if gold[k + '/valid_px'][i] >= 100:      # below that the reference returns before logging consistent parts
                    np.testing.assert_allclose(out['metrics']['s%d/%s' % (i, name)].item(), gold[k + '/' + name][i],
                                               rtol=1e-10, atol=1e-15)


@pytest.mark.parametrize('method', ['sparse-l1', 'sparse-mse', 'sparse-silog', 'sparse-ssi-silog'])
def test_fp32_kernels_at_full_size(method):
    """640x192, batch 8, one scale (what the shipped yaml runs): fp32 kernels vs the fp64 oracle."""
    preds, gt, _ = S.synth_supervised(8, 192, 640, 1, torch.float64, seed=5)
    p64 = [p.clone().requires_grad_() for p in preds]
    ref = S.supervised_loss(p64, gt, method, 1)
    ref.backward()
    from packnet_sfm.losses.supervised_loss import SupervisedLoss
    dev = 'cuda:0'
    p32 = [p.float().to(dev).requires_grad_() for p in preds]
    out = SupervisedLoss(supervised_method=method, supervised_num_scales=1)(p32, gt.float().to(dev))
    out['loss'].backward()
    torch.cuda.synchronize()
    assert abs(out['loss'].item() - ref.item()) <= 1e-5 * abs(ref.item())
    g, g64 = p32[0].grad.double().cpu(), p64[0].grad
    assert ((g - g64).norm() / g64.norm()).item() <= 1e-5
    # deterministic: fixed-order partial sums
    out2 = SupervisedLoss(supervised_method=method, supervised_num_scales=1)([p32[0].detach()], gt.float().to(dev))
    assert torch.equal(out2['loss'], out['loss'].detach())


def test_unsupported_methods_are_refused():
    from packnet_sfm.losses.supervised_loss import SupervisedLoss, get_loss_func
    for m in ('sparse-berhu', 'sparse-ssi-trim', 'sparse-ssi'):
        with pytest.raises(NotImplementedError):
            get_loss_func(m)
    with pytest.raises(ValueError):
        SupervisedLoss(supervised_method='sparse-huber')
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        SupervisedLoss(supervised_method='l1', supervised_num_scales=1)([torch.rand(1, 1, 4, 4)], torch.rand(1, 1, 4, 4))


def test_odd_sizes_and_unaligned_views():
    """Vector loads: sizes that are not a multiple of 4 (scalar tail) and a prediction that is an offset view."""
    from packnet_sfm.losses.supervised_loss import SupervisedLoss
    dev = 'cuda:0'
    g = torch.Generator('cpu').manual_seed(9)
    for (B, H, W) in ((1, 7, 9), (3, 5, 11)):
        base = (0.02 + torch.rand(B * H * W + 1, generator=g, dtype=torch.float64))
        gt = (0.02 + torch.rand(B, 1, H, W, generator=g, dtype=torch.float64))
        gt[gt < 0.3] = 0
        for method in ('sparse-silog', 'sparse-l1'):
            p64 = base[1:].view(B, 1, H, W).clone().requires_grad_()
            ref = S.supervised_loss([p64], gt, method, 1)
            ref.backward()
            holder = base.to(dev)
            pg = holder[1:].view(B, 1, H, W).requires_grad_()           # data_ptr offset by 8 bytes
            out = SupervisedLoss(supervised_method=method, supervised_num_scales=1)([pg], gt.to(dev))
            out['loss'].backward()
            np.testing.assert_allclose(out['loss'].item(), ref.item(), rtol=1e-12)
            np.testing.assert_allclose(pg.grad.cpu().numpy(), p64.grad.numpy(), rtol=1e-10, atol=1e-16)


# ---- DualHeadDepthLoss (reference losses/dual_head_depth_loss.py:23-201) -----------------------------------------

@pytest.mark.parametrize('name', ['default', 'weights', 'resized_gt', 'no_valid'])
def test_dual_head_fp64_kernels_match_reference_fixtures(name):
    from packnet_sfm.losses.dual_head_depth_loss import DualHeadDepthLoss
    from supervised_cases import DUAL_CASES
    gold = load_golden('dual_head')
    kw, shape, gt_size = DUAL_CASES[name]
    ip, fp, gt = S.synth_dual_head(*shape, gt_size=gt_size)
    dev = 'cuda:0'
    ip, fp = ip.to(dev).requires_grad_(), fp.to(dev).requires_grad_()
    L = DualHeadDepthLoss(**kw)
    res = L({('integer', 0): ip, ('fractional', 0): fp}, gt.to(dev), return_logs=True)
    res['loss'].backward()
    torch.cuda.synchronize()
    got = [res[k].item() for k in ('loss', 'integer_loss', 'fractional_loss', 'consistency_loss')]
    np.testing.assert_allclose(got, gold[name + '/losses'], rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(ip.grad.cpu().numpy(), gold[name + '/d_integer'], rtol=1e-10, atol=1e-16)
    np.testing.assert_allclose(fp.grad.cpu().numpy(), gold[name + '/d_fractional'], rtol=1e-10, atol=1e-16)
    if name != 'no_valid':
        logs = [L.metrics[k].item() for k in ('mean_depth_error', 'integer_accuracy', 'fractional_rmse')]
        # the reference forms integer_accuracy with .float() (fp32) even for fp64 inputs
        np.testing.assert_allclose(logs, gold[name + '/logs'], rtol=1e-6)


def test_dual_head_fp32_full_size():
    from packnet_sfm.losses.dual_head_depth_loss import DualHeadDepthLoss
    ip, fp, gt = S.synth_dual_head(16, 384, 640, seed=4)
    i64, f64 = ip.clone().requires_grad_(), fp.clone().requires_grad_()
    ref = S.dual_head_loss(i64, f64, gt)
    ref[0].backward()
    dev = 'cuda:0'
    i32, f32 = ip.float().to(dev).requires_grad_(), fp.float().to(dev).requires_grad_()
    res = DualHeadDepthLoss()({('integer', 0): i32, ('fractional', 0): f32}, gt.float().to(dev))
    res['loss'].backward()
    torch.cuda.synchronize()
    assert abs(res['loss'].item() - ref[0].item()) <= 1e-5 * abs(ref[0].item())
    # sign-type gradients: identical except where fp32 rounding of the ground truth flips a sign (|error| ~ 1e-7)
    for a, b in ((i32.grad, i64.grad), (f32.grad, f64.grad)):
        rel = ((a.double().cpu() - b).norm() / b.norm()).item()
        assert rel <= 2e-3, rel
    with pytest.raises(AssertionError):
        DualHeadDepthLoss(max_depth=1.0, min_depth=2.0)


# ---- SSISilogLoss with its multi-scale Sobel gradient term (gradient_weight > 0) ----
from supervised_cases import GRAD_CASES, GRAD_KW, make_grad_inputs  # noqa: E402


@pytest.mark.parametrize('name', sorted(GRAD_CASES))
def test_gradient_term_fp64_matches_reference_fixtures(name):
    """fp64 kernels vs the reference's SupervisedLoss('sparse-ssi-silog', gradient_weight=0.2).  The reference subtracts two
    Sobel responses that both carry the 1e6 m fills of the invalid pixels (cancellation noise ~1e-10 absolute per pixel);
    the kernels take the Sobel of the difference, hence 1e-9 relative on the loss and 1e-12 absolute on the gradient."""
    from packnet_sfm.losses.supervised_loss import SupervisedLoss
    gold = load_golden('ssi_gradient')
    B, H, W, use_mask, gs = GRAD_CASES[name]
    preds, gt, masks = make_grad_inputs(name)
    dev = 'cuda:0'
    preds = [p.to(dev).requires_grad_() for p in preds]
    L = SupervisedLoss(supervised_method='sparse-ssi-silog', supervised_num_scales=1, **dict(GRAD_KW, gradient_scales=gs))
    out = L(preds, gt.to(dev), masks=None if masks is None else [m.to(dev) for m in masks])
    out['loss'].backward()
    np.testing.assert_allclose(out['loss'].item(), gold[name + '/loss'], rtol=1e-9, atol=1e-300)
    np.testing.assert_allclose(out['metrics']['s0/gradient_component'].item() if name != 'few_valid' else 0.0,
                               gold[name + '/gradient_component'], rtol=1e-9)
    g = gold[name + '/d_pred']
    np.testing.assert_allclose(preds[0].grad.cpu().numpy(), g, rtol=1e-9, atol=1e-12 * max(1.0, np.abs(g).max()))


def test_gradient_term_fp32_at_full_size():
    """NCDB 640x384, batch 4: fp32 kernels vs the fp64 oracle (the oracle's own fp32 run loses ~0.1 m per Sobel response
    to the 1e6 m fills; the kernels do not)."""
    preds, gt, _ = S.synth_supervised(4, 384, 640, 1, torch.float64, seed=9)
    p64 = [p.clone().requires_grad_() for p in preds]
    ref = S.supervised_loss(p64, gt, 'sparse-ssi-silog', 1, **GRAD_KW)
    ref.backward()
    from packnet_sfm.losses.supervised_loss import SupervisedLoss
    dev = 'cuda:0'
    p32 = [p.float().to(dev).requires_grad_() for p in preds]
    out = SupervisedLoss(supervised_method='sparse-ssi-silog', supervised_num_scales=1, **GRAD_KW)(p32, gt.float().to(dev))
    out['loss'].backward()
    assert abs(out['loss'].item() - ref.item()) / abs(ref.item()) < 1e-5
    # the gradient of |.| flips sign where a Sobel response of the difference is within rounding of 0: compare in L2
    err = (p32[0].grad.double().cpu() - p64[0].grad).norm() / p64[0].grad.norm()
    assert err < 1e-3, err
    # deterministic
    p32b = [p.float().to(dev).requires_grad_() for p in preds]
    out2 = SupervisedLoss(supervised_method='sparse-ssi-silog', supervised_num_scales=1, **GRAD_KW)(p32b, gt.float().to(dev))
    out2['loss'].backward()
    assert torch.equal(out['loss'], out2['loss']) and torch.equal(p32[0].grad, p32b[0].grad)


@pytest.mark.parametrize('use_mask', [False, True])
def test_gradient_term_block_and_pixel_kernels_agree(use_mask):
    """H, W multiples of 4 with 16-byte aligned maps run the 4 x 4 block pyramid and the quad backward kernels, anything else
    the per-pixel kernels (same expressions): same loss, identical counts and gradient bits."""
    from packnet_sfm.block_ops import sobel_gradient_loss_op
    dev = 'cuda:0'
    preds, gt, _ = S.synth_supervised(2, 48, 72, 1, torch.float32, seed=5)
    mask = (torch.rand(gt.shape, generator=torch.Generator().manual_seed(1)) < 0.8).to(torch.uint8) if use_mask else None
    n = gt.numel()

    def run(shift):
        bp, bg = torch.zeros(n + 8, device=dev), torch.zeros(n + 8, device=dev)
        p, g = bp[shift:shift + n].view(gt.shape), bg[shift:shift + n].view(gt.shape)
        p.copy_(preds[0])
        g.copy_(gt)
        p.requires_grad_()
        assert p.data_ptr() % 16 == 4 * shift
        loss, state = sobel_gradient_loss_op(p, g, mask=None if mask is None else mask.to(dev), scales=4)
        (3.0 * loss).backward()
        return loss.detach().clone(), state.clone(), p.grad.clone()

    a, b = run(0), run(1)
    assert a[0].item() > 0 and a[2].abs().max().item() > 0
    # the response kernel's quad form adds the same |gx|, |gy| in another order: loss and per-scale losses to fp32 summation
    # order, the counts (hence the backward's coefficients) and the gradient bit for bit
    torch.testing.assert_close(a[0], b[0], rtol=2e-6, atol=0)
    torch.testing.assert_close(a[1], b[1], rtol=2e-6, atol=0)
    assert torch.equal(a[1][1:12], b[1][1:12])
    assert torch.equal(a[2], b[2])
