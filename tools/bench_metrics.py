"""Device-timed compute_depth_metrics at evaluation shapes next to the reference's arithmetic in
PyTorch eager on the same GPU (the CPU oracle moved to the device: boolean gathers, two sorts, ~20 temporaries and a host
read-back per image).  Algorithmic bytes per ground-truth pixel: (selection passes + 1) x (gt 4 B + prediction 4 B);
prints one JSON line per shape."""
import json
import os
import sys
from types import SimpleNamespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'packnet-sfm-resnet-san_b200'))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
from packnet_sfm.utils.depth import compute_depth_metrics  # noqa: E402


def main():
    import metrics_oracle as M            # the eager comparison line only; never on the product path
    dev = torch.device('cuda:0')
    peak = 6548.8
    try:
        peak = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
    except Exception:
        pass
    for (B, H, W, h, w, crop, ugs) in ((8, 192, 640, 192, 640, '', True), (8, 375, 1242, 192, 640, 'garg', True),
                                       (4, 1216, 1936, 1216, 1936, '', True), (4, 1216, 1936, 1216, 1936, '', False)):
        cfg = SimpleNamespace(min_depth=0.0, max_depth=80.0, crop=crop, scale_output='resize')
        R = max(2, int(3 * 126e6 / (B * (H * W + h * w) * 4)) + 1)      # rotate > 3x L2 worth of inputs
        sets = []
        for r in range(R):
            gt, pred = M.synth_metrics_shape(B, H, W, h, w, 'resize', 'sparse', torch.float32, seed=r)
            sets.append((gt.to(dev), pred.to(dev)))
        outs = [None] * R

        def step(i):
            gt, pred = sets[i % R]
            outs[i % R] = compute_depth_metrics(cfg, gt, pred, use_gt_scale=ugs)
        for i in range(2 * R):
            step(i)
        torch.cuda.synchronize()
        K = 200

        def timed(fn, k=K):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(k):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / k
        ms_eager = timed(step)
        graphs = []
        for r in range(R):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                step(r)
            graphs.append(gr)
        ms = timed(lambda i: graphs[i % R].replay())

        def torch_eager(i):
            gt, pred = sets[i % R]
            # the reference's statement sequence on the device
            p = pred if (h, w) == (H, W) else torch.nn.functional.interpolate(pred, size=(H, W), mode='bilinear', align_corners=True)
            acc = torch.zeros(7, device=dev)
            for b in range(B):
                g_i, p_i = gt[b, 0], p[b, 0]
                valid = (g_i > 0.0) & (g_i < 80.0)
                if crop:
                    y1, y2, x1, x2 = M.garg_window(H, W)
                    cm = torch.zeros_like(valid)
                    cm[y1:y2, x1:x2] = True
                    valid = valid & cm
                if valid.sum() == 0:
                    continue
                g_i, p_i = g_i[valid], p_i[valid]
                if ugs:
                    p_i = p_i * (g_i.median() / p_i.median())
                th = torch.max(g_i / p_i, p_i / g_i)
                d = g_i - p_i
                acc += torch.stack([(d.abs() / g_i).mean(), (d * d / g_i).mean(), (d * d).mean().sqrt(),
                                    ((g_i.log() - p_i.log()) ** 2).mean().sqrt(), (th < 1.25).float().mean(),
                                    (th < 1.25 ** 2).float().mean(), (th < 1.25 ** 3).float().mean()])
            return acc / B
        for i in range(3):
            torch_eager(i)
        torch.cuda.synchronize()
        ms_torch = timed(torch_eager, 20)
        passes = 4 if ugs else 1
        gbs = B * (H * W) * 8 * passes / (ms * 1e-3) / 1e9
        print(json.dumps({'op': 'compute_depth_metrics', 'shape': [B, H, W], 'pred': [h, w], 'crop': crop, 'use_gt_scale': ugs,
                          'ms_eager_api': ms_eager, 'ms_graph': ms, 'ms_torch_eager_same_gpu': ms_torch,
                          'images_per_s': B / (ms * 1e-3), 'algorithmic_GBs': gbs, 'frac_of_hbm_peak': gbs / peak,
                          'speedup_vs_torch_eager': ms_torch / ms}))


if __name__ == '__main__':
    main()
