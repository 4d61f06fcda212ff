"""CPU oracle for SupervisedLoss (l1 / mse / silog)  --  TEST INFRASTRUCTURE ONLY.

Restatement in plain functional PyTorch of reference losses/supervised_loss.py:67-82 (SilogLoss), :84-92
(get_loss_func: nn.L1Loss / nn.MSELoss / SilogLoss), :292-417 (calculate_loss) and :420-437 (forward).
Pinned against outputs of the reference's own SupervisedLoss run in the build container
(oracle/gen_golden_supervised.py -> tests/golden/supervised.npz; tests/test_oracle_golden.py).
Never imported by the product."""
import torch
import torch.nn.functional as F


def loss_function(pred, gt, method, ratio2=0.85):
    if method == 'l1':
        return (pred - gt).abs().mean()
    if method == 'mse':
        return ((pred - gt) ** 2).mean()
    if method == 'silog':                       # :73-82
        l = torch.log(pred) - torch.log(gt)
        return torch.sqrt((l ** 2).mean() - ratio2 * l.mean() ** 2)
    raise NotImplementedError(method)


def ssi_silog(pred, gt, mask, alpha=0.85, ratio2=0.85, ssi_weight=0.7, silog_weight=0.3, min_depth=None, max_depth=None):
    """`SSISilogLoss.forward` with gradient_weight = 0 (losses/ssi_silog_loss.py:176-197, :218-262, :366-429)."""
    if mask.sum() < 100:
        return torch.zeros((), dtype=pred.dtype)
    d = pred[mask] - gt[mask]
    mean = d.mean()
    ssi = ((d ** 2).mean() - mean ** 2) + alpha * mean ** 2
    cmin = 1e-3 if min_depth is None else float(min_depth)
    cmax = 100.0 if max_depth is None else float(max_depth)
    P = (1. / pred.clamp(min=1e-6)).clamp(min=cmin, max=cmax)[mask]
    G = (1. / gt.clamp(min=1e-6)).clamp(min=cmin, max=cmax)[mask]
    l = torch.log(P) - torch.log(G)
    silog = torch.sqrt(((l ** 2).mean() - ratio2 * l.mean() ** 2).abs() + 1e-8)
    return ssi_weight * ssi + silog_weight * silog


def supervised_loss(inv_depths, gt_inv_depth, supervised_method, num_scales, masks=None, **kw):
    """`SupervisedLoss.forward` (:420-437): ground truth matched to every scale with nearest interpolation
    (utils/image.py:178-214), sparse methods fill invalid pixels of both maps with 1e-6 (:323-345)."""
    method = 'ssi-silog' if supervised_method.endswith('ssi-silog') else supervised_method.split('-')[-1]
    sparse = supervised_method.startswith('sparse')
    total = 0.0
    for i in range(num_scales):
        p = inv_depths[i]
        g = gt_inv_depth if p.shape[-2:] == gt_inv_depth.shape[-2:] else F.interpolate(gt_inv_depth, size=p.shape[-2:],
                                                                                       mode='nearest')
        if sparse:
            valid = (g > 0.).detach()
            if masks is not None and masks[i] is not None:
                valid = valid & masks[i]
            p, g = p.masked_fill(~valid, 1e-6), g.masked_fill(~valid, 1e-6)
        if method == 'ssi-silog':
            total = total + ssi_silog(p, g, valid if sparse else (1. / g.clamp(min=1e-6)) > 0, **kw)
            continue
        total = total + loss_function(p, g, method)
    return total / float(num_scales)


def synth_supervised(B, H, W, n, dtype=torch.float64, seed=77, pyramid=False):
    """Deterministic inputs: predictions per scale, a sparse ground truth (about 40 % valid, LiDAR-like), byte masks."""
    g = torch.Generator('cpu').manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64)
    sizes = [((H >> s), (W >> s)) if pyramid else (H, W) for s in range(n)]
    preds = [1.0 / (2.0 + 38.0 * r(B, 1, h, w)) for h, w in sizes]
    gt = 1.0 / (2.0 + 38.0 * r(B, 1, H, W))
    gt = torch.where(r(B, 1, H, W) < 0.4, gt, torch.zeros_like(gt))
    masks = [r(B, 1, h, w) > 0.25 for h, w in sizes]
    return [p.to(dtype) for p in preds], gt.to(dtype), masks
