"""Generate tests/golden/supervised.npz from the REAL reference  --  TEST INFRASTRUCTURE, build container only.

Runs the reference's own `SupervisedLoss` (losses/supervised_loss.py) in fp64 on the deterministic recipe
`synth_supervised` (oracle/supervised_oracle.py) for the methods this library provides.

    python oracle/gen_golden_supervised.py
"""
import contextlib
import io
import os

import numpy as np
import torch

import gen_golden as G          # stubs yacs / matplotlib and puts /root/reference first on sys.path
from packnet_sfm.losses.supervised_loss import SupervisedLoss  # noqa: E402
from supervised_oracle import synth_supervised  # noqa: E402

CASES = [(m, n, pyr, use_mask) for m in ('sparse-l1', 'sparse-mse', 'sparse-silog', 'sparse-ssi-silog', 'l1', 'mse', 'silog')
         for (n, pyr, use_mask) in ((1, False, False), (4, True, False), (2, False, True))
         if not (use_mask and not m.startswith('sparse'))]


EXTRA = {'sparse-ssi-silog': dict(min_depth=0.5, max_depth=30.0)}     # clamps some of the 2..40 m depths


def main():
    out = {}
    for method, n, pyr, use_mask in CASES:
        preds, gt, masks = synth_supervised(2, 24, 40, n, pyramid=pyr)
        if not method.startswith('sparse'):
            gt = gt.clamp(min=0.02)                       # dense methods take the log of every pixel
        preds = [p.requires_grad_() for p in preds]
        with contextlib.redirect_stdout(io.StringIO()):  # the reference prints banners
            L = SupervisedLoss(supervised_method=method, supervised_num_scales=n, **EXTRA.get(method, {}))
            res = L(preds, gt, masks=[m.clone() for m in masks] if use_mask else None)
        res['loss'].backward()
        key = '%s_n%d_%s%s' % (method, n, 'P' if pyr else 'U', '_mask' if use_mask else '')
        out[key + '/loss'] = res['loss'].detach().numpy()
        for i, p in enumerate(preds):
            # a scale with fewer than 100 valid pixels returns a constant in SSISilogLoss: no gradient
            out[key + '/d_pred%d' % i] = (torch.zeros_like(p) if p.grad is None else p.grad).numpy()
        if method.startswith('sparse'):
            out[key + '/valid_px'] = np.array([float(res['metrics']['s%d/valid_px' % i]) for i in range(n)])
            out[key + '/scale_loss'] = np.array([float(res['metrics']['s%d/loss' % i]) for i in range(n)])
        if method.endswith('ssi-silog'):
            for name in ('ssi_component', 'silog_component', 'ssi_mean', 'ssi_var', 'silog1', 'silog2'):
                out[key + '/' + name] = np.array([float(res['metrics']['s%d/%s' % (i, name)]) for i in range(n)])
    np.savez_compressed(os.path.join(G.ROOT, 'tests', 'golden', 'supervised.npz'), **out)
    print('wrote supervised.npz with', len(out), 'arrays')


if __name__ == '__main__':
    main()


def dual_head():
    """DualHeadDepthLoss of the reference (fp64) -> tests/golden/dual_head.npz."""
    from packnet_sfm.losses.dual_head_depth_loss import DualHeadDepthLoss
    from supervised_oracle import synth_dual_head
    out = {}
    for name, kw, shape, gt_size in (('default', {}, (2, 24, 40), None),
                                     ('weights', dict(max_depth=12.0, min_depth=1.0, integer_weight=0.5, fractional_weight=4.0,
                                                      consistency_weight=2.0), (2, 24, 40), None),
                                     ('resized_gt', {}, (2, 12, 20), (24, 40)),
                                     ('no_valid', dict(max_depth=0.15, min_depth=0.1), (1, 8, 8), None)):
        ip, fp, gt = synth_dual_head(*shape, gt_size=gt_size)
        ip, fp = ip.requires_grad_(), fp.requires_grad_()
        with contextlib.redirect_stdout(io.StringIO()):
            L = DualHeadDepthLoss(**kw)
            res = L({('integer', 0): ip, ('fractional', 0): fp}, gt, return_logs=True)
        if res['loss'].requires_grad and ip.requires_grad and name != 'no_valid':
            res['loss'].backward()
        out[name + '/losses'] = np.array([float(res[k]) for k in ('loss', 'integer_loss', 'fractional_loss', 'consistency_loss')])
        out[name + '/d_integer'] = (torch.zeros_like(ip) if ip.grad is None else ip.grad).numpy()
        out[name + '/d_fractional'] = (torch.zeros_like(fp) if fp.grad is None else fp.grad).numpy()
        if name != 'no_valid':
            out[name + '/logs'] = np.array([float(L.metrics[k]) for k in ('mean_depth_error', 'integer_accuracy', 'fractional_rmse')])
    np.savez_compressed(os.path.join(G.ROOT, 'tests', 'golden', 'dual_head.npz'), **out)
    print('wrote dual_head.npz with', len(out), 'arrays')


if __name__ == '__main__':
    dual_head()


def gradient_term():
    """SSISilogLoss with gradient_weight = 0.2 through the reference's SupervisedLoss (fp64) -> tests/golden/ssi_gradient.npz."""
    import sys
    sys.path.insert(0, os.path.join(G.ROOT, 'tests'))
    from supervised_cases import GRAD_CASES, GRAD_KW, make_grad_inputs
    out = {}
    for name, (B, H, W, use_mask, gs) in GRAD_CASES.items():
        preds, gt, masks = make_grad_inputs(name)
        preds = [p.requires_grad_() for p in preds]
        with contextlib.redirect_stdout(io.StringIO()):
            L = SupervisedLoss(supervised_method='sparse-ssi-silog', supervised_num_scales=1, **dict(GRAD_KW, gradient_scales=gs))
            L = L.double()          # the Sobel weights are float32 buffers of the module (:25-38)
            res = L(preds, gt, masks=[m.clone() for m in masks] if masks is not None else None)
        if res['loss'].requires_grad:
            res['loss'].backward()
        out[name + '/loss'] = res['loss'].detach().numpy()
        out[name + '/d_pred'] = (torch.zeros_like(preds[0]) if preds[0].grad is None else preds[0].grad).numpy()
        out[name + '/gradient_component'] = np.array(float(res['metrics'].get('s0/gradient_component', 0.0)))
        print(name, float(res['loss']), out[name + '/gradient_component'])
    np.savez_compressed(os.path.join(G.ROOT, 'tests', 'golden', 'ssi_gradient.npz'), **out)
    print('wrote ssi_gradient.npz with', len(out), 'arrays')


if __name__ == '__main__':
    gradient_term()
