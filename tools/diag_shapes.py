"""Where does the fp32 gradient error outside the flip set sit?  (development diagnostic for tests/test_loss_shapes_gpu.py)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'packnet-sfm-resnet-san_b200'), os.path.join(ROOT, 'oracle'), os.path.join(ROOT, 'tests')]
import torch
import test_loss_shapes_gpu as T
name = sys.argv[1] if len(sys.argv) > 1 else 'cfg2_B2_192x640_C2'
inp, ref, ref32, dec = T._oracle(name)
res = T._run_kernels(name, inp, False)
for s in range(len(ref['d_invd'])):
    flip = T._flip_set(dec[s], res['taps'][s], res['sel'][s])
    keep = ~flip
    g, g64, g32 = res['d_invd'][s].double(), ref['d_invd'][s], ref32['d_invd'][s].double()
    den = g64[keep].norm()
    e = ((g - g64) * keep).flatten().abs()
    e32 = ((g32 - g64) * keep).flatten().abs()
    es, _ = e.sort(descending=True)
    tot = (es ** 2).sum()
    print('scale', s, 'excluded %.4f%%' % (100 * flip.double().mean().item()), 'kernel masked %.3g' % (e.norm() / den).item(),
          'ref32 masked %.3g' % (e32.norm() / den).item(),
          'share of squared error in top 1/10/100/1000 px: %.2f %.2f %.2f %.2f' % tuple(((es[:k] ** 2).sum() / tot).item() for k in (1, 10, 100, 1000)),
          'median |err| %.3g, |g| rms %.3g' % (e.median().item(), (g64[keep] ** 2).mean().sqrt().item()))
    # the same with ref32's own error: top pixels
    es32, _ = e32.sort(descending=True)
    print('    ref32: top 1/10/100/1000 share %.2f %.2f %.2f %.2f' % tuple(((es32[:k] ** 2).sum() / (es32 ** 2).sum()).item() for k in (1, 10, 100, 1000)))

# the worst pixels of the scale with the most concentrated error
s = int(sys.argv[2]) if len(sys.argv) > 2 else 0
flip = T._flip_set(dec[s], res['taps'][s], res['sel'][s])
g, g64 = res['d_invd'][s].double(), ref['d_invd'][s]
e = ((g - g64) * (~flip)).abs()
B, _, H, W = e.shape
top = e.flatten().topk(3)
w = res['taps'][s].long()
for v, i in zip(top.values.tolist(), top.indices.tolist()):
    b, r = divmod(i, H * W); y, x = divmod(r, W)
    print('pixel b=%d y=%d x=%d err %.3g g %.6g g64 %.6g g32(ref arithmetic) %.6g' % (b, y, x, v, g[b, 0, y, x].item(), g64[b, 0, y, x].item(), ref32['d_invd'][s][b, 0, y, x].item()))
    print('  3x3 g  ', g[b, 0, max(y-1,0):y+2, max(x-1,0):x+2].numpy().tolist())
    print('  3x3 g64', g64[b, 0, max(y-1,0):y+2, max(x-1,0):x+2].numpy().tolist())
    print('  3x3 g32', ref32['d_invd'][s][b, 0, max(y-1,0):y+2, max(x-1,0):x+2].numpy().tolist())
    ys, xs = slice(max(y - 2, 0), y + 3), slice(max(x - 2, 0), x + 3)
    print('  sel kernel\n', res['sel'][s][b, ys, xs].numpy(), '\n  arg oracle\n', dec[s]['arg'][b, ys, xs].numpy())
    for j in range(w.shape[0]):
        print('  ctx', j, 'kernel x0', ((w[j, b, ys, xs] & 0xfff) - 8).numpy().tolist(), 'oracle x0', dec[s]['x0'][j, b, ys, xs].numpy().tolist())
        print('  ctx', j, 'kernel y0', (((w[j, b, ys, xs] >> 12) & 0x7ff) - 8).numpy().tolist(), 'oracle y0', dec[s]['y0'][j, b, ys, xs].numpy().tolist())
        print('  ctx', j, 'bits kernel', ((w[j, b, ys, xs] >> 23) & 0x1ff).numpy().tolist(), 'oracle', dec[s]['bits'][j, b, ys, xs].numpy().tolist())
    d = inp['inv_depths'][s][b, 0, ys, xs]
    print('  inv depth', d.numpy().tolist())
