"""Catalogue of tests/golden/*.npz: how each fixture's inputs are re-created (recipe S3) and
which loss settings it was produced with (mirrors oracle/gen_golden.py)."""
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')

YAML_LOSS = dict(num_scales=4, ssim_loss_weight=0.85, smooth_loss_weight=0.001, photometric_reduce_op='min',
                 clip_loss=0.0, padding_mode='zeros', automask_loss=True)

_T = dict(B=2, H=24, W=40, C=2, kd=5)
CASES = {
    # name: (synth kwargs, loss overrides, extras)
    'tiny_U_zeros': (dict(_T, pyramid=False), {}, {}),
    'tiny_U_border': (dict(_T, pyramid=False), dict(padding_mode='border'), {}),
    'tiny_P_zeros': (dict(_T, pyramid=True), {}, {}),
    'tiny_P_border': (dict(_T, pyramid=True), dict(padding_mode='border'), {}),
    'tiny_U_mean_noautomask': (dict(_T, pyramid=False), dict(photometric_reduce_op='mean', automask_loss=False), {}),
    'tiny_U_min_noautomask': (dict(_T, pyramid=False), dict(automask_loss=False), {}),
    'tiny_U_clip': (dict(_T, pyramid=False), dict(clip_loss=0.5), {}),
    'tiny_U_mask': (dict(_T, pyramid=False), {}, dict(with_mask=True)),
    'tiny_P_mask': (dict(_T, pyramid=True), {}, dict(with_mask=True)),
    'tiny_U_l1only': (dict(_T, pyramid=False), dict(ssim_loss_weight=0.0), {}),
    'tiny_U_nosmooth': (dict(_T, pyramid=False), dict(smooth_loss_weight=0.0), {}),
    'tiny_U_fork': (dict(_T, pyramid=False), {}, dict(fork=True)),
    'tiny_U_classdefaults': (dict(_T, pyramid=False),
                             dict(ssim_loss_weight=0.85, smooth_loss_weight=0.1, photometric_reduce_op='mean',
                                  clip_loss=0.5, automask_loss=False), {}),
    'tiny_U_C4': (dict(_T, C=4, pyramid=False), {}, {}),
    'tiny_U_odd': (dict(_T, H=23, W=37, pyramid=False), {}, {}),
    'tiny_U_scales2': (dict(_T, pyramid=False), dict(num_scales=2), {}),
    'B2_64x160_C2_U_zeros_mean': (dict(B=2, H=64, W=160, C=2, pyramid=False),
                                  dict(photometric_reduce_op='mean', automask_loss=False), {}),
    'B2_64x160_C2_U_zeros_clip': (dict(B=2, H=64, W=160, C=2, pyramid=False), dict(clip_loss=0.5), {}),
}
for _B, _H, _W in ((2, 64, 160), (1, 192, 640)):
    for _pyr in (False, True):
        for _pm in ('zeros', 'border'):
            CASES['B%d_%dx%d_C2_%s_%s' % (_B, _H, _W, 'P' if _pyr else 'U', _pm)] = (
                dict(B=_B, H=_H, W=_W, C=2, pyramid=_pyr), dict(padding_mode=_pm), {})

# the same recipe seen through fisheye (VADAS) cameras (oracle/gen_golden_fisheye.py)
CASES.update({
    'fisheye_tiny_U_zeros': (dict(_T, pyramid=False), {}, dict(fisheye=True)),
    'fisheye_tiny_U_border': (dict(_T, pyramid=False), dict(padding_mode='border'), dict(fisheye=True)),
    'fisheye_small_P_zeros': (dict(_T, H=32, W=64, kd=3, pyramid=True), {}, dict(fisheye=True)),
    'fisheye_tiny_U_mask': (dict(_T, pyramid=False), {}, dict(fisheye=True, with_mask=True)),
    'fisheye_tiny_U_odd_C4': (dict(_T, H=23, W=37, C=4, pyramid=False), {}, dict(fisheye=True)),
    'fisheye_B2_64x160_C2_U_zeros': (dict(B=2, H=64, W=160, C=2, pyramid=False), {}, dict(fisheye=True)),
    'fisheye_B2_64x160_C2_U_border': (dict(B=2, H=64, W=160, C=2, pyramid=False), dict(padding_mode='border'),
                                      dict(fisheye=True)),
    'fisheye_B1_192x640_C2_U_zeros': (dict(B=1, H=192, W=640, C=2, pyramid=False), {}, dict(fisheye=True)),
})

# Shapes of the benchmarked configurations (BASELINE.json configs[1..4]) for the live-oracle parity tests
# (tests/test_loss_shapes_gpu.py); no fixture files: the CPU oracle runs in the test.
CASES.update({
    'cfg2_B2_192x640_C2': (dict(B=2, H=192, W=640, C=2, pyramid=False), {}, {}),
    'cfg3_B1_384x640_C2_bf16': (dict(B=1, H=384, W=640, C=2, pyramid=False), {}, dict(bf16=True)),
    'cfg4_B1_1216x1936_C2_border': (dict(B=1, H=1216, W=1936, C=2, pyramid=False), dict(padding_mode='border'), {}),
    'cfg5_B1_384x640_C4': (dict(B=1, H=384, W=640, C=4, pyramid=False), {}, {}),
})

FORK_RANGE = (0.5, 80.0)


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN_DIR, name + '.npz')))


def make_inputs(name, dtype=torch.float64, device='cpu', seed=1234):
    """Re-create the inputs of a golden case.  Returns dict(image, context, inv_depths, vecs, K, mask, hp, fork)."""
    from photometric_oracle import synth_S3
    skw, over, extra = CASES[name]
    hp = dict(YAML_LOSS)
    hp.update(over)
    image, context, inv_depths, vecs, K = synth_S3(dtype=torch.float64, seed=seed, **skw)
    fork = None
    if extra.get('fork'):
        lo, hi = 1.0 / 80.0, 1.0 / 0.5
        inv_depths = [((d - lo) / (hi - lo)).clamp(0.0, 1.0) for d in inv_depths]
        fork = FORK_RANGE
    mask = None
    if extra.get('with_mask'):
        gm = torch.Generator('cpu').manual_seed(seed + 7)
        mask = (torch.rand(skw['B'], 1, skw['H'], skw['W'], generator=gm, dtype=torch.float64) > 0.3).to(torch.float64)
    mv = lambda t: t.to(device=device, dtype=dtype)
    if extra.get('fisheye'):      # K becomes the intrinsics dict (k, s, div, ux, uy) of the fork's forward
        from photometric_oracle import synth_fisheye
        K = {n: mv(v) for n, v in synth_fisheye(skw['B'], skw['H'], skw['W']).items()}
        return dict(image=mv(image), context=[mv(c) for c in context], inv_depths=[mv(d) for d in inv_depths],
                    vecs=[mv(v) for v in vecs], K=K, mask=None if mask is None else mv(mask), hp=hp, fork=fork)
    return dict(image=mv(image), context=[mv(c) for c in context], inv_depths=[mv(d) for d in inv_depths],
                vecs=[mv(v) for v in vecs], K=mv(K), mask=None if mask is None else mv(mask), hp=hp, fork=fork)
