"""CPU oracle for compute_depth_metrics  --  TEST INFRASTRUCTURE ONLY.

Restatement in plain functional PyTorch of reference utils/depth.py:258-447 (compute_depth_metrics, without its
diagnostic printing) and :450-481 (scale_depth).  Pinned against outputs of the reference's own function run in the
build container (oracle/gen_golden_metrics.py -> tests/golden/depth_metrics.npz; tests/test_oracle_golden.py).
Never imported by the product."""
import torch
import torch.nn.functional as F

NAMES = ('abs_rel', 'sq_rel', 'rmse', 'rmse_log', 'a1', 'a2', 'a3')


def scale_depth(pred, gt, scale_fn):
    """:450-481 -- 'resize': bilinear, align_corners=True (the input itself when the sizes agree, utils/image.py:138-142);
    'top-center': prediction placed at the bottom, centred horizontally, zeros elsewhere."""
    H, W = gt.shape[-2:]
    h, w = pred.shape[-2:]
    if scale_fn == 'resize':
        return pred if (h, w) == (H, W) else F.interpolate(pred, size=(H, W), mode='bilinear', align_corners=True)
    if scale_fn != 'top-center':
        raise NotImplementedError(scale_fn)
    top, left = H - h, (W - w) // 2
    return F.pad(pred, (left, W - w - left, top, 0))


def garg_window(H, W):
    """:335-339"""
    return int(0.40810811 * H), int(0.99189189 * H), int(0.03594771 * W), int(0.96405229 * W)


def compute_depth_metrics(gt, pred, min_depth, max_depth, crop='', scale_output='resize', use_gt_scale=True,
                          return_per_image=False):
    """Returns [7] (abs_rel, sq_rel, rmse, rmse_log, a1, a2, a3): per-image values summed over the images that have a
    valid pixel, divided by the batch size (:342-447)."""
    B, _, H, W = gt.shape
    pred = scale_depth(pred, gt, scale_output)
    window = torch.ones(H, W, dtype=torch.bool)
    if crop == 'garg':
        y1, y2, x1, x2 = garg_window(H, W)
        window = torch.zeros(H, W, dtype=torch.bool)
        window[y1:y2, x1:x2] = True
    total = torch.zeros(4, dtype=gt.dtype)
    total_a = torch.zeros(3, dtype=torch.float32)    # (thresh < t).float().mean() stays float32 for any map type (:417-419)
    rows = []
    for b in range(B):
        g, p = gt[b, 0], pred[b, 0]
        keep = (g > min_depth) & (g < max_depth) & window
        row = torch.zeros(11, dtype=torch.float64)
        if keep.any():
            g, p = g[keep], p[keep]
            scale = torch.ones((), dtype=gt.dtype)
            if use_gt_scale:
                gm, pm = g.median(), p.median()          # lower median (:376-379)
                scale = gm / pm
                p = p * scale
                row[1], row[2] = gm, pm
            ratio = torch.maximum(g / p, p / g)
            d = g - p
            m = torch.stack([(d.abs() / g).mean(), (d * d / g).mean(), (d * d).mean().sqrt(),
                             ((g.log() - p.log()) ** 2).mean().sqrt()])
            a = torch.stack([(ratio < 1.25 ** k).float().mean() for k in (1, 2, 3)])
            total, total_a = total + m, total_a + a
            row[0], row[3] = float(keep.sum()), scale
            row[4:8], row[8:] = m, a
        else:
            row[3] = 1.0
        rows.append(row)
    out = torch.cat([total / B, (total_a / B).to(gt.dtype)])
    return (out, torch.stack(rows)) if return_per_image else out


CASES = {
    # name: (B, (H, W), (h, w), crop, scale_output, use_gt_scale, min_depth, max_depth, flavour)
    'same_scaled':       (2, (24, 40), (24, 40), '',     'resize',     True,  0.0, 80.0, 'sparse'),
    'same_unscaled':     (2, (24, 40), (24, 40), '',     'resize',     False, 0.0, 80.0, 'sparse'),
    'garg_scaled':       (2, (37, 61), (37, 61), 'garg', 'resize',     True,  0.0, 80.0, 'sparse'),
    'resize_half':       (2, (24, 40), (12, 20), '',     'resize',     True,  0.0, 80.0, 'sparse'),
    'resize_odd_garg':   (3, (37, 61), (13, 21), 'garg', 'resize',     False, 0.0, 80.0, 'sparse'),
    'resize_down':       (1, (24, 40), (48, 80), '',     'resize',     True,  0.0, 80.0, 'dense'),
    'top_center':        (2, (24, 40), (20, 36), '',     'top-center', True,  0.0, 80.0, 'inside'),
    'top_center_garg':   (2, (37, 61), (30, 55), 'garg', 'top-center', False, 0.0, 80.0, 'inside'),
    'depth_range':       (2, (24, 40), (24, 40), '',     'resize',     True,  5.0, 30.0, 'dense'),
    'one_image_empty':   (3, (24, 40), (24, 40), '',     'resize',     True,  0.0, 80.0, 'empty1'),
    'all_empty':         (2, (16, 16), (16, 16), '',     'resize',     True,  50.0, 80.0, 'dense'),
    'single_pixel':      (1, (8, 8),   (8, 8),   '',     'resize',     True,  0.0, 80.0, 'one'),
}


def synth_metrics(name, dtype=torch.float64, seed=303):
    """Deterministic inputs of a catalogue case: a LiDAR-like sparse (or dense) ground truth of 2..40 m and a prediction
    that is a noisy multiple of it (same size) or an independent 2.6..52 m field (other sizes)."""
    B, (H, W), (h, w), crop, scale_output, use_gt_scale, lo, hi, flavour = CASES[name]
    return synth_metrics_shape(B, H, W, h, w, scale_output, flavour, dtype, seed + sorted(CASES).index(name))


def synth_metrics_shape(B, H, W, h, w, scale_output='resize', flavour='sparse', dtype=torch.float64, seed=303):
    g = torch.Generator('cpu').manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64)
    dense = 2.0 + 38.0 * r(B, 1, H, W)
    keep = r(B, 1, H, W) < 0.4
    if (h, w) == (H, W):
        pred = 1.3 * dense * (0.8 + 0.4 * r(B, 1, H, W))
    else:
        pred = 1.3 * (2.0 + 38.0 * r(B, 1, h, w))
    gt = dense.clone()
    if flavour in ('sparse', 'inside', 'empty1'):
        gt = torch.where(keep, dense, torch.zeros_like(dense))
    if flavour == 'inside':          # ground truth only where the placed prediction is (elsewhere the metrics are infinite)
        top, left = H - h, (W - w) // 2
        m = torch.zeros(1, 1, H, W, dtype=torch.bool)
        m[..., top:, left:left + w] = True
        gt = torch.where(m, gt, torch.zeros_like(gt))
    if flavour == 'empty1':
        gt[1] = 0.0
    if flavour == 'one':
        gt = torch.zeros_like(gt)
        gt[0, 0, 3, 5] = 7.5
    return gt.to(dtype), pred.to(dtype)
