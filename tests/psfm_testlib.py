"""Shared helpers of the GPU parity tests."""
import torch

import photometric_oracle as O
from golden_cases import make_inputs


def oracle_run(name, dtype=torch.float64, inputs=None):
    """CPU oracle on a golden case.  Returns dict(loss3, d_invd, d_vec, d_mat, warped, cands)."""
    inp = inputs if inputs is not None else make_inputs(name, dtype)
    invd = [d.detach().clone().requires_grad_() for d in inp['inv_depths']]
    vecs = [v.detach().clone().requires_grad_() for v in inp['vecs']]
    mats = [O.pose_vec_to_mat(v) for v in vecs]
    for m in mats:
        m.retain_grad()
    out = O.multiview_photometric_loss(inp['image'], inp['context'], invd, inp['K'], inp['K'], mats, mask=inp['mask'],
                                       fork_premap=inp['fork'], return_parts=True, **inp['hp'])
    out['loss'].backward()
    n = inp['hp']['num_scales']
    return dict(loss3=torch.stack([out['loss'][0].detach(), out['photometric_loss'], out['smoothness_loss']]),
                d_invd=[invd[i].grad for i in range(n)], d_vec=[v.grad for v in vecs], d_mat=[m.grad for m in mats],
                warped=out['warped'], cands=out['candidates'])


def cuda_run(name, dtype=torch.float64, mode='split', img_dtype=None, inputs=None, device='cuda:0'):
    """The CUDA path through the public Python API on a golden case."""
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    from packnet_sfm.geometry.pose import Pose
    inp = inputs if inputs is not None else make_inputs(name, dtype)
    dev = torch.device(device)
    idt = img_dtype or dtype
    invd = [d.to(dev).detach().clone().requires_grad_() for d in inp['inv_depths']]
    vecs = [v.to(dev).detach().clone().requires_grad_() for v in inp['vecs']]
    poses = [Pose.from_vec(v, 'euler') for v in vecs]
    for p in poses:
        p.mat.retain_grad()
    hp = dict(inp['hp'])
    if inp['fork'] is not None:
        hp.update(sigmoid_premap=True, min_depth=inp['fork'][0], max_depth=inp['fork'][1])
    L = MultiViewPhotometricLoss(kernel_mode=mode, **hp)
    mask = None if inp['mask'] is None else inp['mask'].to(dev).to(idt)
    K = {k: v.to(dev) for k, v in inp['K'].items()} if isinstance(inp['K'], dict) else inp['K'].to(dev)
    out = L(inp['image'].to(dev).to(idt), [c.to(dev).to(idt) for c in inp['context']], invd, K, K, poses, mask=mask)
    out['loss'].backward()
    torch.cuda.synchronize()
    n = L.n
    z = torch.zeros((), dtype=dtype, device=dev)
    return dict(loss3=torch.stack([out['loss'][0].detach(), out['metrics']['photometric_loss'],
                                   out['metrics'].get('smoothness_loss', z)]).cpu(),
                d_invd=[invd[i].grad.cpu() for i in range(n)], d_vec=[v.grad.cpu() for v in vecs],
                d_mat=[p.mat.grad.cpu() for p in poses])


def rel_l2(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300)).item()
