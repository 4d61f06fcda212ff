"""Device-timed SupervisedLoss forward + backward at the shipped shapes; prints one line per method.
Algorithmic bytes per pixel: forward reads pred + gt (8 B), backward reads pred + gt and writes d pred (12 B)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'packnet-sfm-resnet-san_b200'))
from packnet_sfm.losses.supervised_loss import SupervisedLoss  # noqa: E402


def main():
    dev = torch.device('cuda:0')
    peak = 6548.8
    try:
        peak = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
    except Exception:
        pass
    shapes = ((8, 192, 640), (16, 384, 640), (4, 1216, 1936))
    if os.environ.get('PSFM_BENCH_SHAPES'):      # e.g. 4x1216x1936,8x192x640 (for a short ncu launch list)
        shapes = tuple(tuple(int(v) for v in t.split('x')) for t in os.environ['PSFM_BENCH_SHAPES'].split(','))
    only = os.environ.get('PSFM_BENCH_METHODS')
    for (B, H, W) in shapes:
        g = torch.Generator('cpu').manual_seed(3)
        # rotate > 3x L2 worth of inputs
        R = max(2, int(3 * 126e6 / (B * H * W * 12)) + 1)
        sets = []
        for r in range(R):
            pred = (1.0 / (2.0 + 38.0 * torch.rand(B, 1, H, W, generator=g))).to(dev).requires_grad_()
            gt = 1.0 / (2.0 + 38.0 * torch.rand(B, 1, H, W, generator=g))
            gt = torch.where(torch.rand(B, 1, H, W, generator=g) < 0.4, gt, torch.zeros_like(gt)).to(dev)
            sets.append((pred, gt))
        for method in ('sparse-l1', 'sparse-silog', 'sparse-ssi-silog', 'sparse-ssi-silog+sobel'):
            if only and method not in only.split(','):
                continue
            if method.endswith('+sobel'):   # the dual-head NCDB yaml: gradient_weight 0.2 over 4 scales
                L = SupervisedLoss(supervised_method='sparse-ssi-silog', supervised_num_scales=1, gradient_weight=0.2,
                                   gradient_scales=4, min_depth=0.5, max_depth=80.0)
            else:
                L = SupervisedLoss(supervised_method=method, supervised_num_scales=1)

            def step(i):
                pred, gt = sets[i % R]
                pred.grad = None
                L([pred], gt)['loss'].backward()
            for i in range(10):
                step(i)
            torch.cuda.synchronize()
            K = 200

            def timed(fn):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(K):
                    fn(i)
                e1.record()
                torch.cuda.synchronize()
                return e0.elapsed_time(e1) / K
            ms_eager = timed(step)
            # device time: one CUDA graph per resident input set (the eager loop is bound by Python + launch overhead)
            graphs = []
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for r in range(R):
                    step(r)
            torch.cuda.current_stream().wait_stream(side)
            for r in range(R):
                gr = torch.cuda.CUDAGraph()
                sets[r][0].grad = None
                with torch.cuda.graph(gr):
                    step(r)
                graphs.append(gr)
            ms = timed(lambda i: graphs[i % R].replay())
            gbs = B * H * W * (40 if method.endswith('+sobel') else 20) / (ms * 1e-3) / 1e9
            print(json.dumps({'method': method, 'shape': [B, H, W], 'ms_per_step_eager': ms_eager, 'ms_per_step_graph': ms, 'frames_per_s': B / (ms * 1e-3),
                              'algorithmic_GBs': gbs, 'frac_of_hbm_peak': gbs / peak}))


if __name__ == '__main__':
    main()
