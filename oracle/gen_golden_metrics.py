"""Generate tests/golden/depth_metrics.npz from the REAL reference  --  TEST INFRASTRUCTURE, build container only.

Runs the reference's own `compute_depth_metrics` (utils/depth.py:258-447) in fp64 on the deterministic catalogue
`metrics_oracle.CASES`.

    python oracle/gen_golden_metrics.py
"""
import contextlib
import io
import os
from types import SimpleNamespace

import numpy as np

import gen_golden as G          # stubs yacs / matplotlib and puts /root/reference first on sys.path
from packnet_sfm.utils.depth import compute_depth_metrics  # noqa: E402
from metrics_oracle import CASES, synth_metrics  # noqa: E402


def main():
    out = {}
    for name, (B, _, _, crop, scale_output, use_gt_scale, lo, hi, _) in CASES.items():
        gt, pred = synth_metrics(name)
        cfg = SimpleNamespace(min_depth=lo, max_depth=hi, crop=crop, scale_output=scale_output)
        with contextlib.redirect_stdout(io.StringIO()):  # the reference prints a report per configuration
            m = compute_depth_metrics(cfg, gt, pred, use_gt_scale=use_gt_scale)
        out[name] = m.numpy()
        print(name, m.numpy())
    np.savez_compressed(os.path.join(G.ROOT, 'tests', 'golden', 'depth_metrics.npz'), **out)
    print('wrote depth_metrics.npz with', len(out), 'arrays')


if __name__ == '__main__':
    main()
