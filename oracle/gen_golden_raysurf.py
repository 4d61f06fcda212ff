"""Generate tests/golden/raysurf_*.npz from the REAL reference  --  TEST INFRASTRUCTURE, build container only.

SURVEY.md 8(f) N3: the ray-surface (generic camera) photometric loss -- `GenericMultiViewPhotometricLoss`
(losses/generic_multiview_photometric_loss.py:92-400) with `GenericCamera` (geometry/camera_generic.py:17-208) and
`view_synthesis_generic` (geometry/camera_utils.py:64-96).  The reference's classes run as shipped except for what
cannot run in this container:
  * the hard-coded `.cuda()` calls (camera_generic.py:131,168-171) and `ref_image.get_device()` (:190 of the loss) are
    made no-ops / 'cpu' (no GPU here; they do not touch the arithmetic);
  * fp64: the default dtype is double and `.float()` of the integer patch coordinates (:193) returns double;
  * `np.load("omnicam_ray_template.npy")` (:146-148): the file is not shipped.  It is only the initial ray table, so a
    deterministic synthetic template (`raysurf_oracle.synth_ray_template`: unit rays of a wide-angle camera with a mild
    radial distortion) takes its place, for the reference run, the oracle and the kernels alike.
The class is batch-1 by construction (`.squeeze()`, `depth[0]`), so are the fixtures.

    python oracle/gen_golden_raysurf.py      # writes tests/golden/raysurf_*.npz
"""
import os

import numpy as np
import torch

import gen_golden as G          # stubs yacs / matplotlib and puts /root/reference first on sys.path
from gen_golden import Pose, ROOT

import raysurf_oracle as RO     # input recipe + restatement

# ---- make the reference's GPU-only statements runnable on the CPU (no arithmetic involved) --------------------------
torch.Tensor.cuda = lambda self, *a, **k: self
torch.Tensor.get_device = lambda self: torch.device('cpu')
# fp64 fixtures: the class builds its identity pose with the default dtype and casts the (integer) patch coordinates
# with .float(); both become double here (integers below 2^24: no value changes)
torch.set_default_dtype(torch.float64)
torch.Tensor.float = lambda self, *a, **k: self.double()
_TEMPLATE = {}
_np_load = np.load
np.load = lambda f, *a, **k: _TEMPLATE['t'] if str(f).endswith('omnicam_ray_template.npy') else _np_load(f, *a, **k)

from packnet_sfm.losses.generic_multiview_photometric_loss import GenericMultiViewPhotometricLoss  # noqa: E402

_identity = Pose.identity.__func__
Pose.identity = classmethod(lambda cls, N=1, device=None, dtype=torch.float64: _identity(cls, N, device, torch.float64))


def run_case(name, H, W, C, progress, seed=1234, **hp):
    cfg = dict(RO.RAYSURF_LOSS)
    cfg.update(hp)
    image, context, inv_depth, residual, vecs, template = RO.synth_raysurf(H, W, C, seed=seed)
    _TEMPLATE['t'] = template.numpy()
    inv_depths = [inv_depth.clone().requires_grad_()]
    residual = residual.clone().requires_grad_()
    vecs = [v.clone().requires_grad_() for v in vecs]
    poses = [Pose.from_vec(v, 'euler') for v in vecs]
    for p in poses:
        p.mat.retain_grad()
    L = GenericMultiViewPhotometricLoss(**cfg).double()
    L.canonical_ray_surface = L.canonical_ray_surface.double()
    out = L(image, context, inv_depths, {('raysurf', 0): residual}, None, None, poses, False, progress)
    out['loss'].backward()
    warped = L.warp_ref_image(inv_depths, context[0], residual, poses[0], progress)[0].detach()
    rec = dict(loss=out['loss'].detach().numpy(), photometric_loss=np.array(float(L.metrics['photometric_loss'])),
               smoothness_loss=np.array(float(L.metrics.get('smoothness_loss', 0.0))),
               d_inv_depth=inv_depths[0].grad.numpy(), d_residual=residual.grad.numpy(),
               d_vec=np.stack([v.grad.numpy() for v in vecs]), d_mat=np.stack([p.mat.grad.numpy() for p in poses]),
               warped0=warped.numpy(), progress=np.array(progress))
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', name + '.npz'), **rec)
    print(name, 'loss %.12f' % float(out['loss']), 'd_res norm %.4g' % float(residual.grad.norm()))


if __name__ == '__main__':
    for name, (skw, hp) in RO.RAYSURF_CASES.items():
        run_case(name, **skw, **hp)
