"""Where do the fp32 kernels differ from the fp64 oracle?  Prints the worst pixels of d/d inv_depth
with the distances of the oracle's discrete decisions to their flipping points."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, 'packnet-sfm-resnet-san_b200'), os.path.join(ROOT, 'oracle'), os.path.join(ROOT, 'tests')):
    sys.path.insert(0, p)
import torch
import photometric_oracle as O
from golden_cases import make_inputs
from psfm_testlib import cuda_run, oracle_run

name = sys.argv[1] if len(sys.argv) > 1 else 'B1_192x640_C2_U_zeros'
bf16 = len(sys.argv) > 2 and sys.argv[2] == 'bf16'
inp = make_inputs(name, torch.float64)
if bf16:
    rnd = lambda t: t.to(torch.bfloat16).double()
    inp['image'], inp['context'] = rnd(inp['image']), [rnd(c) for c in inp['context']]
ref = oracle_run(name, torch.float64, inputs=inp)
inp32 = {k: ([t.float() for t in v] if isinstance(v, list) else (v.float() if torch.is_tensor(v) else v)) for k, v in inp.items()}
res = cuda_run(name, torch.float32, 'fused', img_dtype=torch.bfloat16 if bf16 else None, inputs=inp32)
ref32 = oracle_run(name, torch.float32, inputs=inp32)
for s in range(len(ref['d_invd'])):
    g, g64, g32 = res['d_invd'][s].double(), ref['d_invd'][s], ref32['d_invd'][s].double()
    e = (g - g64).abs()
    B, _, H, W = g.shape
    print('scale', s, 'rel-L2 kernel %.3g  fp32-oracle %.3g  |g64| %.3g' % (((g - g64).norm() / g64.norm()).item(), ((g32 - g64).norm() / g64.norm()).item(), g64.norm().item()))
    K = O.scale_K(inp['K'], W / inp['image'].shape[3])
    depth = O.inv2depth(inp['inv_depths'][s])
    coords = []
    for v in inp['vecs']:
        grid = O.project(O.reconstruct(depth, K), K, O.pose_vec_to_mat(v))
        coords.append(((grid[..., 0] + 1) / 2 * (W - 1), (grid[..., 1] + 1) / 2 * (H - 1)))
    cands = torch.cat(ref['cands'][s], 1).detach()
    top = e.flatten().topk(6)
    for val, idx in zip(top.values.tolist(), top.indices.tolist()):
        b, r = divmod(idx, H * W); v, u = divmod(r, W)
        near = []
        for dv in (-1, 0, 1):
            for du in (-1, 0, 1):
                vv, uu = min(max(v + dv, 0), H - 1), min(max(u + du, 0), W - 1)
                c = cands[b, :, vv, uu]
                t2 = c.topk(2, largest=False).values
                fr = min(min(abs(cx[b, vv, uu].item() - round(cx[b, vv, uu].item())), abs(cy[b, vv, uu].item() - round(cy[b, vv, uu].item()))) for cx, cy in coords)
                near.append((round((t2[1] - t2[0]).item(), 8), round(fr, 6)))
        print('  err %.3g at (b%d,v%d,u%d) g64 %.3g g %.3g g32 %.3g  min cand gap %.2g  min frac %.2g' % (
            val, b, v, u, g64[b, 0, v, u].item(), g[b, 0, v, u].item(), g32[b, 0, v, u].item(), min(n[0] for n in near), min(n[1] for n in near)))
