"""Catalogue of tests/golden/supervised.npz (mirrors oracle/gen_golden_supervised.py)."""
import torch

CASES = [(m, n, pyr, use_mask) for m in ('sparse-l1', 'sparse-mse', 'sparse-silog', 'sparse-ssi-silog', 'l1', 'mse', 'silog')
         for (n, pyr, use_mask) in ((1, False, False), (4, True, False), (2, False, True))
         if not (use_mask and not m.startswith('sparse'))]

EXTRA = {'sparse-ssi-silog': dict(min_depth=0.5, max_depth=30.0)}


def key_of(case):
    method, n, pyr, use_mask = case
    return '%s_n%d_%s%s' % (method, n, 'P' if pyr else 'U', '_mask' if use_mask else '')


def make_inputs(case, dtype=torch.float64):
    from supervised_oracle import synth_supervised
    method, n, pyr, use_mask = case
    preds, gt, masks = synth_supervised(2, 24, 40, n, dtype, pyramid=pyr)
    if not method.startswith('sparse'):
        gt = gt.clamp(min=0.02)
    return preds, gt, (masks if use_mask else None)

DUAL_CASES = {
    'default': ({}, (2, 24, 40), None),
    'weights': (dict(max_depth=12.0, min_depth=1.0, integer_weight=0.5, fractional_weight=4.0, consistency_weight=2.0),
                (2, 24, 40), None),
    'resized_gt': ({}, (2, 12, 20), (24, 40)),
    'no_valid': (dict(max_depth=0.15, min_depth=0.1), (1, 8, 8), None),
}

# SSISilogLoss with its multi-scale Sobel gradient term (tests/golden/ssi_gradient.npz; oracle/gen_golden_supervised.py)
GRAD_KW = dict(gradient_weight=0.2, gradient_scales=4, min_depth=0.5, max_depth=30.0)
GRAD_CASES = {
    # name: (B, H, W, use_mask, gradient_scales)
    'even': (2, 24, 40, False, 4),
    'odd': (2, 37, 61, False, 4),
    'mask': (2, 24, 40, True, 4),
    'small_top_scale_skipped': (2, 16, 16, False, 4),      # 2 x 2 at scale 3: no Sobel output
    'two_scales': (1, 32, 48, False, 2),
    'few_valid': (1, 12, 16, False, 4),                    # < 100 valid pixels: the constant 0
}


def make_grad_inputs(name, dtype=torch.float64):
    from supervised_oracle import synth_supervised
    B, H, W, use_mask, _ = GRAD_CASES[name]
    preds, gt, masks = synth_supervised(B, H, W, 1, dtype, seed=131)
    return preds, gt, (masks if use_mask else None)
