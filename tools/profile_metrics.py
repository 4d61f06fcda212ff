"""A few compute_depth_metrics calls at one shape, for `ncu` (launch list / --set full); prints nothing that is a bench value."""
import os
import sys
from types import SimpleNamespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'packnet-sfm-resnet-san_b200'))
from packnet_sfm.utils.depth import compute_depth_metrics  # noqa: E402

B, H, W, h, w = [int(v) for v in (sys.argv[1:6] if len(sys.argv) > 5 else (4, 1216, 1936, 1216, 1936))]
g = torch.Generator('cpu').manual_seed(0)
dense = 2.0 + 38.0 * torch.rand(B, 1, H, W, generator=g)
gt = torch.where(torch.rand(B, 1, H, W, generator=g) < 0.4, dense, torch.zeros_like(dense)).cuda()
pred = (1.3 * (2.0 + 38.0 * torch.rand(B, 1, h, w, generator=g))).cuda()
cfg = SimpleNamespace(min_depth=0.0, max_depth=80.0, crop='garg' if (h, w) != (H, W) else '', scale_output='resize')
for _ in range(3):
    out = compute_depth_metrics(cfg, gt, pred)
torch.cuda.synchronize()
print('done', out.tolist())
