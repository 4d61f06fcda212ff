"""Depth helpers on the photometric-loss path (reference: utils/depth.py:103-198) and the evaluation metrics
(reference: utils/depth.py:258-481)."""
from packnet_sfm.utils.types import is_seq


def inv2depth(inv_depth):
    """1 / clamp(inv_depth, 1e-6), list-aware (reference :103-120)."""
    if is_seq(inv_depth):
        return [inv2depth(item) for item in inv_depth]
    return 1. / inv_depth.clamp(min=1e-6)


def depth2inv(depth):
    """1 / clamp(depth, 1e-6) with zeros where depth <= 0, list-aware (reference :123-143)."""
    if is_seq(depth):
        return [depth2inv(item) for item in depth]
    inv_depth = 1. / depth.clamp(min=1e-6)
    inv_depth[depth <= 0.] = 0.
    return inv_depth


def inv_depths_normalize(inv_depths):
    """Divide every map by its per-image mean (clamped at 1e-6) (reference :146-162)."""
    from packnet_sfm.ops import normalize_op
    return [normalize_op(d) for d in inv_depths]


def calc_smoothness(inv_depths, images, num_scales):
    """Edge-aware first-order smoothness maps (reference :165-198): returns
    (smoothness_x [B,1,H,W-1], smoothness_y [B,1,H-1,W]) per scale, CUDA kernels through the C ABI."""
    from packnet_sfm.ops import smoothness_op
    out = [smoothness_op(inv_depths[i], images[i]) for i in range(num_scales)]
    return [o[0] for o in out], [o[1] for o in out]


# north-star name (BASELINE.json) for the same function; the reference only has calc_smoothness
inv_depths_smoothness = calc_smoothness


def compute_depth_metrics(config, gt, pred, use_gt_scale=True, return_state=False):
    """Depth metrics of a batch (reference :258-447): `config` carries `min_depth`, `max_depth`, `crop` ('garg' or
    anything else) and `scale_output` ('resize' | 'top-center', scale_depth :450-481); gt [B,1,H,W], pred [B,1,h,w].
    Returns a [7] tensor (abs_rel, sq_rel, rmse, rmse_log, a1, a2, a3) of gt's dtype on gt's device.  One C-ABI call
    (`psfm_depth_metrics`): exact per-image medians by radix selection, no gather / sort / host read-back.
    `return_state=True` (extension) also returns the per-image block [B,12] float64 = {valid count, gt median, pred
    median, scale, the image's seven metrics, 0}."""
    from packnet_sfm.block_ops import depth_metrics_op
    if config.scale_output == 'resize':
        mode = 0
    elif config.scale_output == 'top-center':
        mode = 1
    else:
        raise NotImplementedError('Depth scale function {} not implemented.'.format(config.scale_output))
    out, state = depth_metrics_op(gt, pred, mode, config.crop == 'garg', bool(use_gt_scale), float(config.min_depth),
                                  float(config.max_depth))
    return (out, state) if return_state else out


def scale_depth(pred, gt, scale_fn):
    """Match depth maps to the ground-truth resolution (reference :450-481): 'resize' = bilinear, align_corners=True
    (the resize kernel behind `interpolate_image`); 'top-center' = prediction placed at the bottom, centred, zeros
    elsewhere.  `compute_depth_metrics` does not call this: it evaluates the same values per pixel inside its kernels."""
    if scale_fn == 'resize':
        from packnet_sfm.utils.image import interpolate_image
        return interpolate_image(pred, gt.shape, mode='bilinear', align_corners=True)
    if scale_fn == 'top-center':
        import torch
        pred_uncropped = torch.zeros(gt.shape, dtype=pred.dtype, device=pred.device)
        top, left = gt.shape[2] - pred.shape[2], (gt.shape[3] - pred.shape[3]) // 2
        pred_uncropped[:, :, top:(top + pred.shape[2]), left:(left + pred.shape[3])] = pred
        return pred_uncropped
    raise NotImplementedError('Depth scale function {} not implemented.'.format(scale_fn))
