"""T2 / T3 of the parity protocol: the fp32 production kernels against the fp64
oracle at the north star's tolerances.

  warped image   max-abs <= 1e-4
  loss           relative <= 1e-5
  gradients      norm-wise relative <= 1e-4 on pixels outside the *fragile set*: pixels (1-px dilated)
                 where a discrete decision of the loss sits within fp32 noise of flipping in the fp64
                 oracle itself -- a sampling coordinate within 3e-4 px of an integer (bilinear tap
                 floor; one fp32 ulp of a coordinate near 640 is 6e-5 px), the two best candidates within
                 2e-5 (arg-min) or |warped - target| below 5e-5 (L1 sign): the coordinate noise times the
                 image slope moves a warped value by up to ~3e-5.  One flipped pixel changes its gradient by O(100 %),
                 which is a property of the function, not of the arithmetic: the reference's own fp32
                 run shows the same (T3).
  T3             unmasked error of the kernels <= max(3x the error of the fp32 *reference arithmetic*
                 (oracle run in fp32 on the CPU) against the same fp64 oracle, 1e-2): a handful of
                 flipped pixels decides this number for either implementation (measured floor of the
                 reference 6e-4 ... 1e-2).
"""
import pytest
import torch
import torch.nn.functional as F

import photometric_oracle as O
from golden_cases import make_inputs
from psfm_testlib import cuda_run, oracle_run, rel_l2

pytestmark = pytest.mark.gpu


def fragile_mask(inp, ref, scale):
    """[B,1,H,W] bool: pixels whose gradient hinges on a decision within fp32 noise (dilated by 1 px)."""
    hp = inp['hp']
    d = inp['inv_depths'][scale]
    B, _, H, W = d.shape
    sf = W / inp['image'].shape[3]
    K = O.scale_K(inp['K'], sf)
    frag = torch.zeros(B, 1, H, W, dtype=torch.bool)
    depth = O.inv2depth(d)
    for j, v in enumerate(inp['vecs']):
        T = O.pose_vec_to_mat(v)
        grid = O.project(O.reconstruct(depth, K), K, T)                  # [B,H,W,2] normalised
        ix = (grid[..., 0] + 1) / 2 * (W - 1)
        iy = (grid[..., 1] + 1) / 2 * (H - 1)
        for c, lim in ((ix, W - 1), (iy, H - 1)):
            frag |= ((c - c.round()).abs() < 3e-4).unsqueeze(1)
            if hp['padding_mode'] == 'border':          # the clip points of the border mode
                frag |= ((c.abs() < 3e-4) | ((c - lim).abs() < 3e-4)).unsqueeze(1)
    cands = torch.cat(ref['cands'][scale], 1).detach()                   # [B,K,H,W]
    top2 = cands.topk(2, dim=1, largest=False).values
    frag |= ((top2[:, 1:2] - top2[:, 0:1]) < 2e-5)
    image = inp['image'] if (H, W) == tuple(inp['image'].shape[2:]) else O.match_scales(inp['image'], inp['inv_depths'], scale + 1)[scale]
    for j in range(len(inp['vecs'])):
        w = ref['warped'][j][scale].detach()
        frag |= ((w - image).abs() < 5e-5).any(1, keepdim=True)
    return F.max_pool2d(frag.double(), 3, 1, 1) > 0


@pytest.mark.parametrize('name', ['B2_64x160_C2_U_zeros', 'B2_64x160_C2_U_border', 'B2_64x160_C2_P_zeros',
                                  'B1_192x640_C2_U_zeros'])
def test_fp32_kernels_meet_north_star_tolerances(name):
    inp64 = make_inputs(name, torch.float64)
    ref = oracle_run(name, torch.float64, inputs=inp64)
    inp32 = {k: ([t.float() for t in v] if isinstance(v, list) else (v.float() if torch.is_tensor(v) else v))
             for k, v in inp64.items()}
    res = cuda_run(name, torch.float32, 'fused', inputs=inp32)
    ref32 = oracle_run(name, torch.float32, inputs=inp32)               # the reference arithmetic in fp32 (CPU)

    # loss: 1e-5 relative
    for k in range(2):
        assert abs(res['loss3'][k].double() - ref['loss3'][k]) <= 1e-5 * abs(ref['loss3'][k]), k
    n = len(ref['d_invd'])
    tot_bad = 0
    for s in range(n):
        g, g64, g32 = res['d_invd'][s].double(), ref['d_invd'][s], ref32['d_invd'][s].double()
        frag = fragile_mask(inp64, ref, s)
        keep = ~frag
        tot_bad += frag.double().mean().item() / n
        masked = ((g - g64)[keep].norm() / g64[keep].norm()).item()
        assert masked <= 1e-4, 'scale %d: masked rel-L2 %.3g' % (s, masked)
        # T3: unmasked, against the reference's own fp32 error
        assert rel_l2(g, g64) <= max(3.0 * rel_l2(g32, g64), 1e-2), 'scale %d unmasked' % s
    assert tot_bad < 0.10, 'fragile set covers %.1f %% of the pixels' % (100 * tot_bad)
    dv, dv64, dv32 = torch.stack(res['d_vec']).double(), torch.stack(ref['d_vec']), torch.stack(ref32['d_vec']).double()
    assert rel_l2(dv, dv64) <= max(3.0 * rel_l2(dv32, dv64), 1e-2)
    dm, dm64, dm32 = torch.stack(res['d_mat']).double(), torch.stack(ref['d_mat']), torch.stack(ref32['d_mat']).double()
    assert rel_l2(dm, dm64) <= max(3.0 * rel_l2(dm32, dm64), 1e-2)


@pytest.mark.parametrize('pm', ['zeros', 'border'])
def test_fp32_warped_image_max_abs(pm):
    from packnet_sfm.geometry.camera import Camera
    from packnet_sfm.geometry.camera_utils import view_synthesis
    from packnet_sfm.geometry.pose import Pose
    inp = make_inputs('B1_192x640_C2_U_zeros', torch.float64)
    d, K, img = inp['inv_depths'][0], inp['K'], inp['context'][0]
    want = O.view_synthesis(img, O.inv2depth(d), K, O.pose_vec_to_mat(inp['vecs'][0]), K, pm)
    dev = 'cuda:0'
    f = lambda t: t.float().to(dev)
    got = view_synthesis(f(img), f(O.inv2depth(d)), Camera(f(K), Tcw=Pose.from_vec(f(inp['vecs'][0]), 'euler')),
                         Camera(f(K)), padding_mode=pm)
    assert (got.double().cpu() - want).abs().max().item() <= 1e-4


def test_bf16_images_fp32_accumulation():
    """Config 3: bf16 image storage, fp32 arithmetic; the oracle sees the same bf16-rounded images."""
    name = 'B2_64x160_C2_U_zeros'
    inp = make_inputs(name, torch.float64)
    rnd = lambda t: t.to(torch.bfloat16).double()
    inp['image'], inp['context'] = rnd(inp['image']), [rnd(c) for c in inp['context']]
    ref = oracle_run(name, torch.float64, inputs=inp)
    inp32 = {k: ([t.float() for t in v] if isinstance(v, list) else (v.float() if torch.is_tensor(v) else v))
             for k, v in inp.items()}
    res = cuda_run(name, torch.float32, 'fused', img_dtype=torch.bfloat16, inputs=inp32)
    assert abs(res['loss3'][0].double() - ref['loss3'][0]) <= 1e-5 * abs(ref['loss3'][0])
    for s in range(len(ref['d_invd'])):
        keep = ~fragile_mask(inp, ref, s)
        g, g64 = res['d_invd'][s].double(), ref['d_invd'][s]
        assert ((g - g64)[keep].norm() / g64[keep].norm()).item() <= 1e-4
