"""torch.autograd.Function wrappers over the C ABI (include/psfm.h).

`photometric_loss` is the fused hot path behind `MultiViewPhotometricLoss.forward`
(reference: losses/multiview_photometric_loss.py:331-410); `view_synthesis_op` backs
`geometry.camera_utils.view_synthesis` (reference: geometry/camera_utils.py:27-59);
`resize_op` backs `utils.image.interpolate_image` (reference: utils/image.py:117-146).
PyTorch is used for device memory, streams and autograd plumbing only.
"""
import ctypes as C
import os

import torch

from . import _C
from .block_ops import (abs_mean_op, normalize_op, photometric_map_op, project_op, reconstruct_op,  # noqa: F401
                        reduce_candidates_op, smoothness_op, ssim_op, transform_points_op)

# 'fused': loss and unit-upstream gradients in one sweep (psfm_loss_forward_backward), the
#          backward only applies the upstream scalar.
# 'split': psfm_loss_forward, then psfm_loss_backward recomputes the warp (north-star layout).
DEFAULT_MODE = os.environ.get('PSFM_LOSS_MODE', 'fused')


# set by a call with hp['debug_decisions']: {'taps': [C,B,Hs,Ws] int32 decision words per scale, 'sel': [B,Hs,Ws] uint8}
LAST_DECISIONS = None

FISHEYE_KEYS = ('k', 's', 'div', 'ux', 'uy')


def pack_fisheye(intr, B, dtype, device=None):
    """The fork's fisheye intrinsics dict (k [B,7], s, div, ux, uy [B]; datasets/ncdb_dataset.py:577-584) ->
    [B,11] rows (k0..k6, s, div, ux, uy), the layout of PSFM_CAM_FISHEYE in include/psfm.h."""
    missing = [n for n in FISHEYE_KEYS if n not in intr]
    if missing:
        raise KeyError('fisheye intrinsics need the keys %s (missing %s)' % (FISHEYE_KEYS, missing))
    k = torch.as_tensor(intr['k'])
    device = k.device if device is None else device
    k = k.to(device=device, dtype=dtype).reshape(-1, 7)
    cols = [k.expand(B, 7)] + [torch.as_tensor(intr[n]).to(device=device, dtype=dtype).reshape(-1, 1).expand(B, 1)
                                for n in FISHEYE_KEYS[1:]]
    return torch.cat(cols, 1).contiguous()


def _is_fisheye(K):
    """Fisheye intrinsics: the fork's dict, or its packed [B,11] form (pack_fisheye)."""
    return isinstance(K, dict) or (torch.is_tensor(K) and K.dim() == 2 and K.shape[1] == 11)


def _as_batched_K(K, B, dtype):
    if isinstance(K, dict):
        return pack_fisheye(K, B, dtype)
    if _is_fisheye(K):
        if K.shape[0] != B:
            raise ValueError('packed fisheye intrinsics must be [B,11], got %s' % (tuple(K.shape),))
        return K.to(dtype).contiguous()
    if K.dim() == 2:
        K = K.unsqueeze(0).expand(B, 3, 3)
    if tuple(K.shape) != (B, 3, 3):
        raise ValueError('intrinsics must be [B,3,3], got %s' % (tuple(K.shape),))
    return K.to(dtype).contiguous()


RESIZE_MODES = {'bilinear': 0, 'nearest': 1, 'bilinear_ac_false': 2}


class _ResizeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, Ho, Wo, mode):
        x = x.contiguous()
        Hi, Wi = x.shape[-2:]
        out = torch.empty(x.shape[:-2] + (Ho, Wo), dtype=x.dtype, device=x.device)
        N = x.numel() // (Hi * Wi)
        with torch.cuda.device(x.device):
            _C.check(_C.lib.psfm_resize(N, Hi, Wi, Ho, Wo, _C.dtype_code(x), mode, _C.ptr(x), _C.ptr(out),
                                        _C.stream_ptr(x.device)))
        ctx.geom = (N, Hi, Wi, Ho, Wo, mode, x.shape, x.dtype)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        N, Hi, Wi, Ho, Wo, mode, shape, dtype = ctx.geom
        real = torch.float64 if dtype == torch.float64 else torch.float32
        g = grad_out.to(real).contiguous()
        gin = torch.empty(shape, dtype=real, device=g.device)
        with torch.cuda.device(g.device):
            _C.check(_C.lib.psfm_resize_backward(N, Hi, Wi, Ho, Wo, _C.dtype_code(g), mode, _C.ptr(g), _C.ptr(gin),
                                                 _C.stream_ptr(g.device)))
        return gin.to(dtype), None, None, None


def resize_op(x, size, mode='bilinear'):
    """[..,Hi,Wi] -> [..,Ho,Wo] with F.interpolate's arithmetic: 'bilinear' (align_corners=True), 'nearest',
    'bilinear_ac_false' (align_corners=False).  Differentiable (reference utils/image.py:117-146 runs F.interpolate
    under autograd; the model's upsample_output, models/model_utils.py:175-193, needs the gradient)."""
    _C.require_cuda(x)
    if mode not in RESIZE_MODES:
        raise NotImplementedError('resize mode %s' % mode)
    if x.dtype == torch.uint8:
        raise ValueError('resize takes float images')
    return _ResizeFn.apply(x, int(size[0]), int(size[1]), RESIZE_MODES[mode])


_WS_CACHE = {}


def _workspace(dev, nbytes, shared):
    """Scratch buffer of a loss call.  `shared`: one persistent buffer per (device, stream), grown on demand -- valid
    for the fused mode, whose workspace is dead once the forward kernels have run (stream order protects the next
    call); the split mode keeps per-call scratch because its backward reads the forward's state."""
    if not shared or torch.cuda.is_current_stream_capturing():
        return torch.empty(nbytes, dtype=torch.uint8, device=dev)
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    buf = _WS_CACHE.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _WS_CACHE[key] = buf
    return buf


class _LossPlan:
    """cfg / io structs and buffers for one call."""

    def __init__(self, hp, image, context, inv_depths, K, ref_K, pose_mats, mask):
        n, Cn = len(inv_depths), len(context)
        B, ch, H, W = image.shape
        if ch != 3:
            raise ValueError('image must be [B,3,H,W]')
        if n < 1 or n > _C.MAX_SCALES:
            raise ValueError('between 1 and %d scales are supported, got %d' % (_C.MAX_SCALES, n))
        if Cn < 1 or Cn > _C.MAX_CTX:
            raise ValueError('between 1 and %d context frames are supported, got %d' % (_C.MAX_CTX, Cn))
        if len(pose_mats) != Cn:
            raise ValueError('need one pose per context frame')
        _C.require_cuda(image, mask, *context, *inv_depths, *pose_mats)
        for Kx in (K, ref_K):
            _C.require_cuda(*(list(Kx.values()) if isinstance(Kx, dict) else [Kx]))
        real = inv_depths[0].dtype
        if real not in (torch.float32, torch.float64):
            raise ValueError('inverse depths must be float32 (or float64 for the math-parity build)')
        img_dtype = image.dtype
        dev = image.device
        self.real, self.dev, self.n, self.C, self.B = real, dev, n, Cn, B
        self.upsampled = False
        cfg = _C.LossCfg()
        cfg.B, cfg.C, cfg.n, cfg.H, cfg.W = B, Cn, n, H, W
        cfg.img_dtype = _C.dtype_code(image)
        cfg.real_dtype = _C.F64 if real == torch.float64 else _C.F32
        cfg.padding = _C.PAD[hp['padding_mode']] if hp['padding_mode'] in _C.PAD else -1
        if hp['photometric_reduce_op'] not in _C.REDUCE:
            raise NotImplementedError('Unknown photometric_reduce_op: {}'.format(hp['photometric_reduce_op']))
        cfg.reduce_op = _C.REDUCE[hp['photometric_reduce_op']]
        cfg.automask = int(bool(hp['automask_loss']))
        cfg.premap = int(hp.get('premap', 0))
        cfg.ssim_w, cfg.smooth_w = float(hp['ssim_loss_weight']), float(hp['smooth_loss_weight'])
        cfg.C1, cfg.C2, cfg.clip_loss = float(hp['C1']), float(hp['C2']), float(hp['clip_loss'])
        cfg.min_depth, cfg.max_depth = float(hp.get('min_depth', 0.05)), float(hp.get('max_depth', 80.0))
        io = _C.LossIO()
        keep = []
        image = image.contiguous()
        self.u8 = img_dtype == torch.uint8
        if self.u8 and any(c.dtype != torch.uint8 for c in context):
            raise ValueError('uint8 images (bytes, value = byte / 255) need uint8 context frames')
        context = [c.to(img_dtype).contiguous() for c in context]
        for c in context:
            if tuple(c.shape) != (B, 3, H, W):
                raise ValueError('context frames must match the image shape')
        if mask is not None:
            mask = mask.to(img_dtype).contiguous()
        self.inv_depths = [d.contiguous() for d in inv_depths]
        for s, d in enumerate(self.inv_depths):
            if d.dim() != 4 or d.shape[0] != B or d.shape[1] != 1 or d.dtype != real:
                raise ValueError('inv_depths[%d] must be [B,1,Hs,Ws] %s' % (s, real))
            Hs, Ws = int(d.shape[2]), int(d.shape[3])
            if hp.get('upsample_depth_maps') and (Hs, Ws) != (H, W):
                # the map is nearest-upsampled to the image size inside the kernels (models/model_utils.py:152-176)
                if Hs > H or Ws > W:
                    raise ValueError('upsample_depth_maps: inv_depths[%d] is larger than the image' % s)
                cfg.Hd[s], cfg.Wd[s] = Hs, Ws
                Hs, Ws = H, W
                self.upsampled = True
            cfg.Hs[s], cfg.Ws[s] = Hs, Ws
            if (Hs, Ws) == (H, W):          # interpolate_image returns the image itself (utils/image.py:138-142)
                img_s, ctx_s, mask_s = image, context, mask
            else:                           # match_scales: bilinear align_corners=True; nearest for the mask
                img_s = resize_op(image, (Hs, Ws))
                ctx_s = [resize_op(c, (Hs, Ws)) for c in context]
                mask_s = None if mask is None else resize_op(mask, (Hs, Ws), mode='nearest')
            keep += [img_s, mask_s] + list(ctx_s)
            io.image[s] = img_s.data_ptr()
            io.inv_depth[s] = d.data_ptr()
            io.mask[s] = 0 if mask_s is None else mask_s.data_ptr()
            for j in range(Cn):
                io.ctx[s][j] = ctx_s[j].data_ptr()
        if _is_fisheye(K) != _is_fisheye(ref_K):
            raise ValueError('K and ref_K must both be pinhole matrices or both be fisheye intrinsics')
        cfg.camera_model = _C.CAM_FISHEYE if _is_fisheye(K) else _C.CAM_PINHOLE
        self.K = _as_batched_K(K, B, real)
        self.ref_K = _as_batched_K(ref_K, B, real)
        io.K, io.ref_K = self.K.data_ptr(), self.ref_K.data_ptr()
        self.fisheye = _is_fisheye(K)
        self.pose = []
        for j, m in enumerate(pose_mats):
            if tuple(m.shape) != (B, 4, 4):
                raise ValueError('pose matrices must be [B,4,4]')
            m = m.to(real).contiguous()
            self.pose.append(m)
            io.pose[j] = m.data_ptr()
        self.keep = keep
        self.cfg, self.io = cfg, io
        # scratch for the image sets that exist (one set when every scale is at image resolution); the fused mode
        # needs it only while the forward call's kernels run, so calls on the same stream share one buffer
        self.ws_bytes = int(_C.lib.psfm_loss_workspace_bytes_io(C.byref(cfg), C.byref(io)))
        if self.ws_bytes == 0:
            _C.check(-1)
        self.ws = _workspace(dev, self.ws_bytes, shared=hp.get('_mode') == 'fused')
        self.sel = [torch.empty((B, cfg.Hs[s], cfg.Ws[s]), dtype=torch.uint8, device=dev) for s in range(n)]
        for s in range(n):
            io.sel[s] = self.sel[s].data_ptr()
        # test hook (include/psfm.h, io->dbg): the discrete decisions of the fused kernels, for the parity protocol
        self.dbg = None
        if hp.get('debug_decisions'):
            self.dbg = [torch.zeros((Cn, B, cfg.Hs[s], cfg.Ws[s]), dtype=torch.int32, device=dev) for s in range(n)]
            for s in range(n):
                self.sel[s].fill_(255)
                io.dbg[s] = self.dbg[s].data_ptr()
        self.out = torch.empty(3, dtype=real, device=dev)
        self.grads = None

    def alloc_grads(self):
        sizes = [d.numel() for d in self.inv_depths] + [self.B * 16] * self.C
        flat = torch.empty(sum(sizes), dtype=self.real, device=self.dev)
        views, off = [], 0
        for sz in sizes:
            views.append(flat[off:off + sz])
            off += sz
        self.grads = flat
        self.d_invd = [v.view_as(d) for v, d in zip(views[:self.n], self.inv_depths)]
        self.d_pose = [v.view(self.B, 4, 4) for v in views[self.n:]]
        for s in range(self.n):
            self.io.d_inv_depth[s] = self.d_invd[s].data_ptr()
        for j in range(self.C):
            self.io.d_pose[j] = self.d_pose[j].data_ptr()

    def args(self):
        return C.byref(self.cfg), C.byref(self.io), _C.ptr(self.ws), self.ws_bytes


class _PhotometricLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hp, mode, image, K, ref_K, mask, n, Cn, *tensors):
        context, inv_depths, pose_mats = tensors[:Cn], tensors[Cn:Cn + n], tensors[Cn + n:]
        plan = _LossPlan(dict(hp, _mode=mode), image, context, inv_depths, K, ref_K, pose_mats, mask)
        need_grad = any(t.requires_grad for t in tensors[Cn:])
        st = _C.stream_ptr(plan.dev)
        with torch.cuda.device(plan.dev):
            if (plan.upsampled or plan.fisheye or plan.u8) and mode != 'fused':
                raise NotImplementedError("upsample_depth_maps / fisheye cameras / uint8 images need kernel_mode='fused'")
            if (need_grad or plan.upsampled or plan.fisheye or plan.u8) and mode == 'fused':
                plan.alloc_grads()
                _C.check(_C.lib.psfm_loss_forward_backward(*plan.args(), _C.ptr(plan.out), st))
            else:
                _C.check(_C.lib.psfm_loss_forward(*plan.args(), _C.ptr(plan.out), st))
        if plan.dbg is not None:
            global LAST_DECISIONS
            LAST_DECISIONS = dict(taps=plan.dbg, sel=plan.sel)
        ctx.plan, ctx.mode, ctx.consumed = plan, mode, False
        ctx.counts = (n, Cn)
        out, plan.out = plan.out, None      # the node must not reference its own output (reference cycle)
        # loss [1] (differentiable) and the two logged parts as separate outputs: autograd hands backward the
        # gradient of the loss itself, with no zeros + copy for a slice of a 3-vector
        loss, photo, smooth = out[0:1], out[1], out[2]
        ctx.mark_non_differentiable(photo, smooth)
        return loss, photo, smooth

    @staticmethod
    def backward(ctx, grad_out, _gp=None, _gs=None):
        plan, (n, Cn) = ctx.plan, ctx.counts
        g = grad_out.to(plan.real).contiguous()
        st = _C.stream_ptr(plan.dev)
        with torch.cuda.device(plan.dev):
            if ctx.mode == 'fused' and not ctx.consumed:
                ctx.consumed = True
                _C.check(_C.lib.psfm_scale_inplace(_C.ptr(plan.grads), _C.ptr(g), plan.grads.numel(),
                                                   plan.cfg.real_dtype, st))
            elif ctx.mode == 'fused':
                # a second backward through the same graph (retain_graph=True): the fused pass handed its gradient
                # buffers to autograd the first time, so this one recomputes them with the forward / backward kernel
                # pair from the inputs the node still holds
                if plan.upsampled or plan.fisheye or plan.u8:
                    raise RuntimeError('a second backward through the fused photometric loss is recomputed by the '
                                       'split kernels, which do not take upsample_depth_maps / fisheye cameras / '
                                       'uint8 images; call the loss again instead')
                plan.alloc_grads()
                ws = torch.empty(plan.ws_bytes, dtype=torch.uint8, device=plan.dev)
                out = torch.empty(3, dtype=plan.real, device=plan.dev)
                args = (C.byref(plan.cfg), C.byref(plan.io), _C.ptr(ws), plan.ws_bytes)
                _C.check(_C.lib.psfm_loss_forward(*args, _C.ptr(out), st))
                _C.check(_C.lib.psfm_loss_backward(*args, _C.ptr(g), st))
            else:
                plan.alloc_grads()
                _C.check(_C.lib.psfm_loss_backward(*plan.args(), _C.ptr(g), st))
        return (None,) * 8 + (None,) * Cn + tuple(plan.d_invd) + tuple(plan.d_pose)


def photometric_loss(hp, image, context, inv_depths, K, ref_K, pose_mats, mask=None, mode=None):
    """Returns (loss [1], photometric_loss [], smoothness_loss []); the loss carries the gradient w.r.t.
    `inv_depths` and `pose_mats`, the two parts are logged values."""
    mode = DEFAULT_MODE if mode is None else mode
    if mode not in ('fused', 'split'):
        raise ValueError('mode must be "fused" or "split"')
    n, Cn = len(inv_depths), len(context)
    return _PhotometricLossFn.apply(hp, mode, image, K, ref_K, mask, n, Cn, *context, *inv_depths, *pose_mats)


class _ViewSynthesisFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ref_image, depth, K, ref_K, pose, padding):
        _C.require_cuda(ref_image, depth, K, ref_K, pose)
        B, ch, H, W = ref_image.shape
        if ch != 3:
            raise ValueError('ref_image must be [B,3,H,W]')
        real = depth.dtype
        ref_image = ref_image.contiguous()
        depth = depth.contiguous()
        K, ref_K = _as_batched_K(K, B, real), _as_batched_K(ref_K, B, real)
        pose = pose.to(real).contiguous()
        out = torch.empty((B, 3, H, W), dtype=real, device=depth.device)
        rd = _C.F64 if real == torch.float64 else _C.F32
        with torch.cuda.device(depth.device):
            _C.check(_C.lib.psfm_warp_forward(B, H, W, _C.dtype_code(ref_image), rd, padding, _C.ptr(ref_image),
                                              _C.ptr(depth), _C.ptr(K), _C.ptr(ref_K), _C.ptr(pose), _C.ptr(out),
                                              _C.stream_ptr(depth.device)))
        ctx.save_for_backward(ref_image, depth, K, ref_K, pose)
        ctx.padding = padding
        return out

    @staticmethod
    def backward(ctx, grad_out):
        ref_image, depth, K, ref_K, pose = ctx.saved_tensors
        B, _, H, W = ref_image.shape
        real = depth.dtype
        rd = _C.F64 if real == torch.float64 else _C.F32
        grad_out = grad_out.to(real).contiguous()
        d_depth = torch.empty_like(depth)
        d_pose = torch.empty((B, 4, 4), dtype=real, device=depth.device)
        d_ref = torch.zeros(ref_image.shape, dtype=real, device=depth.device) if ctx.needs_input_grad[0] else None
        nbytes = int(_C.lib.psfm_warp_workspace_bytes(B, H, W, rd))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=depth.device)
        with torch.cuda.device(depth.device):
            _C.check(_C.lib.psfm_warp_backward(B, H, W, _C.dtype_code(ref_image), rd, ctx.padding, _C.ptr(ref_image),
                                               _C.ptr(depth), _C.ptr(K), _C.ptr(ref_K), _C.ptr(pose),
                                               _C.ptr(grad_out), _C.ptr(d_depth), _C.ptr(d_pose), _C.ptr(d_ref),
                                               _C.ptr(ws), nbytes, _C.stream_ptr(depth.device)))
        if d_ref is not None:
            d_ref = d_ref.to(ref_image.dtype)
        return d_ref, d_depth, None, None, d_pose, None


def view_synthesis_op(ref_image, depth, K, ref_K, pose_mat, padding_mode='zeros'):
    if padding_mode not in _C.PAD:
        raise ValueError('padding_mode must be zeros or border (got %s)' % padding_mode)
    if depth.dtype not in (torch.float32, torch.float64):
        raise ValueError('depth must be float32 or float64')
    return _ViewSynthesisFn.apply(ref_image, depth, K, ref_K, pose_mat, _C.PAD[padding_mode])
