"""compute_depth_metrics (reference utils/depth.py:258-447, scale_depth :450-481) through
psfm_depth_metrics.  fp64 kernels against fixtures produced by the reference's own function; per-image medians (exact
radix selection) against torch.median; fp32 kernels at evaluation sizes against the fp64 oracle."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

import metrics_oracle as M
from golden_cases import load_golden

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _cfg(name):
    B, _, _, crop, scale_output, use_gt_scale, lo, hi, _ = M.CASES[name]
    return SimpleNamespace(min_depth=lo, max_depth=hi, crop=crop, scale_output=scale_output), use_gt_scale


@pytest.mark.parametrize('name', sorted(M.CASES))
def test_fp64_kernels_match_reference_fixtures(name):
    from packnet_sfm.utils.depth import compute_depth_metrics
    gold = load_golden('depth_metrics')[name]
    cfg, ugs = _cfg(name)
    gt, pred = M.synth_metrics(name)
    out, state = compute_depth_metrics(cfg, gt.to(DEV), pred.to(DEV), use_gt_scale=ugs, return_state=True)
    assert out.shape == (7,) and out.dtype == torch.float64 and out.device.type == 'cuda'
    got = out.cpu().numpy()
    # abs_rel, sq_rel, rmse, rmse_log: fp64 arithmetic, summation order differs; a1..a3: the reference's float32 path, exact
    np.testing.assert_allclose(got[:4], gold[:4], rtol=1e-12, atol=1e-300)
    np.testing.assert_array_equal(got[4:], gold[4:])
    # per image: valid count, medians and scale are exact (selection, not approximation)
    _, rows = M.compute_depth_metrics(gt, pred, cfg.min_depth, cfg.max_depth, cfg.crop, cfg.scale_output, ugs,
                                      return_per_image=True)
    st = state.cpu()
    assert torch.equal(st[:, 0], rows[:, 0])
    if cfg.scale_output == 'top-center' or gt.shape[-2:] == pred.shape[-2:]:
        assert torch.equal(st[:, 1:4], rows[:, 1:4])          # no interpolation involved: bit-identical medians
    else:
        torch.testing.assert_close(st[:, 1:4], rows[:, 1:4], rtol=1e-13, atol=0)
    torch.testing.assert_close(st[:, 4:8], rows[:, 4:8], rtol=1e-12, atol=1e-300)
    assert torch.equal(st[:, 8:11], rows[:, 8:11])


@pytest.mark.parametrize('shape', [(8, 192, 640, 192, 640, '', 'resize'), (4, 375, 1242, 192, 640, 'garg', 'resize'),
                                   (2, 1216, 1936, 1216, 1936, '', 'resize'), (3, 375, 1242, 352, 1216, 'garg', 'top-center')])
@pytest.mark.parametrize('use_gt_scale', [True, False])
def test_fp32_kernels_at_evaluation_sizes(shape, use_gt_scale):
    """KITTI evaluation shapes (prediction at network size, ground truth at 375x1242, garg crop) and DDAD: fp32 kernels
    vs the fp64 oracle on the same float32 inputs.  Tolerance 2e-5 relative (float32 per-pixel terms; a threshold
    count may differ by a pixel or two out of >1e4 where a ratio rounds across 1.25^k)."""
    from packnet_sfm.utils.depth import compute_depth_metrics
    B, H, W, h, w, crop, so = shape
    gt, pred = M.synth_metrics_shape(B, H, W, h, w, so, 'inside' if so == 'top-center' else 'sparse', torch.float32, seed=11)
    cfg = SimpleNamespace(min_depth=0.0, max_depth=80.0, crop=crop, scale_output=so)
    out, state = compute_depth_metrics(cfg, gt.to(DEV), pred.to(DEV), use_gt_scale=use_gt_scale, return_state=True)
    ref, rows = M.compute_depth_metrics(gt.double(), pred.double(), 0.0, 80.0, crop, so, use_gt_scale, return_per_image=True)
    assert out.dtype == torch.float32
    np.testing.assert_allclose(out.cpu().numpy()[:4], ref.numpy()[:4], rtol=2e-5)
    np.testing.assert_allclose(out.cpu().numpy()[4:], ref.numpy()[4:], rtol=1e-6, atol=3.0 / rows[:, 0].min().item())
    assert torch.equal(state[:, 0].cpu(), rows[:, 0])
    if use_gt_scale:
        assert torch.equal(state[:, 1].cpu(), rows[:, 1])                    # gt median: same float32 values, exact
        if (h, w) == (H, W) or so == 'top-center':
            assert torch.equal(state[:, 2].cpu(), rows[:, 2])                # no interpolation: exact selection
        else:   # float32 source coordinates (ry * y, as torch's float32 kernel forms them) on a noisy field
            torch.testing.assert_close(state[:, 2].cpu(), rows[:, 2], rtol=1e-4, atol=0)


@pytest.mark.parametrize('n_valid', [1, 2, 3, 4, 255, 256, 2049])
def test_lower_median_and_ties(n_valid):
    """torch.median returns the lower of the two middle elements; duplicates and negative predictions order correctly."""
    from packnet_sfm.utils.depth import compute_depth_metrics
    g = torch.Generator('cpu').manual_seed(n_valid)
    H, W = 48, 64
    gt = torch.zeros(1, 1, H, W, dtype=torch.float32)
    idx = torch.randperm(H * W, generator=g)[:n_valid]
    vals = torch.randint(1, 9, (n_valid,), generator=g).float() * 2.5          # many ties
    gt.view(-1)[idx] = vals
    pred = torch.randn(1, 1, H, W, generator=g)                               # both signs
    cfg = SimpleNamespace(min_depth=0.0, max_depth=80.0, crop='', scale_output='resize')
    _, state = compute_depth_metrics(cfg, gt.to(DEV), pred.to(DEV), return_state=True)
    keep = gt.view(-1) > 0
    assert state[0, 0].item() == n_valid
    assert state[0, 1].item() == gt.view(-1)[keep].median().item()
    assert state[0, 2].item() == pred.view(-1)[keep].median().item()


def test_errors_and_determinism():
    from packnet_sfm.utils.depth import compute_depth_metrics, scale_depth
    gt, pred = M.synth_metrics_shape(2, 64, 96, 32, 48, dtype=torch.float32)
    gt, pred = gt.to(DEV), pred.to(DEV)
    cfg = SimpleNamespace(min_depth=0.0, max_depth=80.0, crop='garg', scale_output='resize')
    a = compute_depth_metrics(cfg, gt, pred)
    b = compute_depth_metrics(cfg, gt, pred)
    assert torch.equal(a, b)
    # the per-pixel evaluation inside the kernels equals the stand-alone resize kernel's output
    c = compute_depth_metrics(cfg, gt, scale_depth(pred, gt, 'resize'))
    assert torch.equal(a, c)
    cfg.scale_output = 'bottom-left'
    with pytest.raises(NotImplementedError, match='not implemented'):
        compute_depth_metrics(cfg, gt, pred)
    cfg.scale_output = 'top-center'
    with pytest.raises(ValueError, match='top-center'):
        compute_depth_metrics(cfg, pred, gt)          # prediction larger than the ground truth
    d = compute_depth_metrics(cfg, gt, pred)
    e = compute_depth_metrics(SimpleNamespace(min_depth=0.0, max_depth=80.0, crop='garg', scale_output='resize'), gt,
                              scale_depth(pred, gt, 'top-center'))
    assert torch.equal(torch.nan_to_num(d, 7., 8., 9.), torch.nan_to_num(e, 7., 8., 9.))
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        compute_depth_metrics(cfg, gt.cpu(), pred.cpu())
