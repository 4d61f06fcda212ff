"""Depth helpers on the photometric-loss path (reference: utils/depth.py:103-198)."""
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
