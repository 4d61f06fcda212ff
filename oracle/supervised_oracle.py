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


def sobel_gradient_term(pred, gt, mask, scales=4):
    """`SSISilogLoss.compute_gradient_loss` (losses/ssi_silog_loss.py:116-174) with `Gradient2D` (:12-50) on the depth maps
    of the (filled) inverse depths.  pred, gt: [B,1,H,W] inverse depths; mask: [B,1,H,W] bool."""
    kx = torch.tensor([[-1., 0., 1.], [-2., 0., 2.], [-1., 0., 1.]], dtype=pred.dtype).view(1, 1, 3, 3)
    ky = kx.transpose(2, 3).contiguous()
    P, G = 1. / pred.clamp(min=1e-6), 1. / gt.clamp(min=1e-6)
    total, used = 0.0, 0
    for s in range(scales):
        if s == 0:
            p, g, m = P, G, mask
        else:
            f = 1.0 / 2 ** s
            p = F.interpolate(P, scale_factor=f, mode='bilinear', align_corners=False)
            g = F.interpolate(G, scale_factor=f, mode='bilinear', align_corners=False)
            m = F.interpolate(mask.to(pred.dtype), scale_factor=f, mode='nearest') > 0.5
        if p.shape[2] < 3 or p.shape[3] < 3:
            continue
        m = m[:, :, 1:-1, 1:-1]
        if m.sum() > 0:
            total = total + ((F.conv2d(p, kx) - F.conv2d(g, kx)).abs()[m].mean() +
                             (F.conv2d(p, ky) - F.conv2d(g, ky)).abs()[m].mean())
            used += 1
    return total / used if used else torch.zeros((), dtype=pred.dtype)


def supervised_loss(inv_depths, gt_inv_depth, supervised_method, num_scales, masks=None, gradient_weight=0.0,
                    gradient_scales=4, **kw):
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
            m = valid if sparse else (1. / g.clamp(min=1e-6)) > 0
            total = total + ssi_silog(p, g, m, **kw)
            if gradient_weight > 0 and m.sum() >= 100:
                total = total + gradient_weight * sobel_gradient_term(p, g, m, gradient_scales)
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


def dual_head_loss(integer_pred, fractional_pred, depth_gt, max_depth=15.0, min_depth=0.5, integer_weight=1.0,
                   fractional_weight=10.0, consistency_weight=0.5):
    """`DualHeadDepthLoss.forward` (losses/dual_head_depth_loss.py:83-201; decompose_depth / dual_head_to_depth,
    networks/layers/resnet/layers.py:76-149).  Returns (loss, integer, fractional, consistency)."""
    if depth_gt.shape[-2:] != integer_pred.shape[-2:]:
        depth_gt = F.interpolate(depth_gt, size=integer_pred.shape[-2:], mode='nearest')
    mask = (depth_gt > min_depth) & (depth_gt < max_depth)
    if mask.sum() == 0:
        z = torch.zeros((), dtype=depth_gt.dtype)
        return z, z, z, z
    fl = torch.floor(depth_gt)
    li = (integer_pred[mask] - (fl / max_depth)[mask]).abs().mean()
    lf = (fractional_pred[mask] - (depth_gt - fl)[mask]).abs().mean()
    lc = ((integer_pred * max_depth + fractional_pred)[mask] - depth_gt[mask]).abs().mean()
    return integer_weight * li + fractional_weight * lf + consistency_weight * lc, li, lf, lc


def synth_dual_head(B, H, W, dtype=torch.float64, seed=91, gt_size=None):
    """Deterministic dual-head inputs: sigmoid-like predictions near the decomposition of a 0.2..18 m ground truth."""
    g = torch.Generator('cpu').manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64)
    gh, gw = gt_size or (H, W)
    gt = 0.2 + 17.8 * r(B, 1, gh, gw)
    gt = torch.where(r(B, 1, gh, gw) < 0.6, gt, torch.zeros_like(gt))
    ip = (r(B, 1, H, W)).clamp(0.01, 0.99)
    fp = (r(B, 1, H, W)).clamp(0.01, 0.99)
    return ip.to(dtype), fp.to(dtype), gt.to(dtype)
