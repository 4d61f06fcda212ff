"""Building-block operators of the photometric-loss path over the C ABI (include/psfm.h,
csrc/block_kernels.cu): what `Camera.reconstruct/project`, `Pose.transform_points`, `SSIM`,
`calc_photometric_loss`, `reduce_photometric_loss`, `inv_depths_normalize` and `calc_smoothness`
of the reference compute, one kernel each, with hand-written backward kernels.  PyTorch only
allocates the buffers and records the autograd edges.
"""
import ctypes as C

import torch

from . import _C


def _real(t):
    if t.dtype not in (torch.float32, torch.float64):
        raise ValueError('float32 or float64 tensors expected, got %s' % t.dtype)
    return _C.F64 if t.dtype == torch.float64 else _C.F32


def _ws(B, N, device):
    n = int(_C.lib.psfm_block_workspace_bytes(int(B), int(N)))
    return torch.empty(n, dtype=torch.uint8, device=device), n


def _batched_K(K, B, like):
    if K.dim() == 2:
        K = K.unsqueeze(0).expand(B, 3, 3)
    if tuple(K.shape) != (B, 3, 3):
        raise ValueError('intrinsics must be [B,3,3], got %s' % (tuple(K.shape),))
    return K.to(like.dtype).contiguous()


def _no_grad_supported(t, what):
    if t is not None and t.requires_grad:
        raise NotImplementedError('gradient w.r.t. %s is outside the accelerated path' % what)


class _Reconstruct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depth, K, Twc):
        _C.require_cuda(depth, K, Twc)
        B, _, H, W = depth.shape
        depth = depth.contiguous()
        K = _batched_K(K, B, depth)
        Twc = None if Twc is None else Twc.to(depth.dtype).contiguous()
        out = torch.empty((B, 3, H, W), dtype=depth.dtype, device=depth.device)
        with torch.cuda.device(depth.device):
            _C.check(_C.lib.psfm_reconstruct(B, H, W, _real(depth), _C.ptr(depth), _C.ptr(K), _C.ptr(Twc), _C.ptr(out),
                                             _C.stream_ptr(depth.device)))
        ctx.save_for_backward(K, Twc)
        return out

    @staticmethod
    def backward(ctx, g):
        K, Twc = ctx.saved_tensors
        g = g.contiguous()
        B, _, H, W = g.shape
        d_depth = torch.empty((B, 1, H, W), dtype=g.dtype, device=g.device)
        with torch.cuda.device(g.device):
            _C.check(_C.lib.psfm_reconstruct_backward(B, H, W, _real(g), _C.ptr(K), _C.ptr(Twc), _C.ptr(g),
                                                      _C.ptr(d_depth), _C.stream_ptr(g.device)))
        return d_depth, None, None


def reconstruct_op(depth, K, Twc=None):
    """Camera.reconstruct: depth [B,1,H,W] -> points [B,3,H,W]; differentiable w.r.t. depth."""
    _real(depth)
    _no_grad_supported(K, 'K')
    _no_grad_supported(Twc, 'the camera pose of reconstruct')
    return _Reconstruct.apply(depth, K, Twc)


class _Project(torch.autograd.Function):
    @staticmethod
    def forward(ctx, X, K, Tcw):
        _C.require_cuda(X, K, Tcw)
        B, _, H, W = X.shape
        X = X.contiguous()
        K = _batched_K(K, B, X)
        Tcw = None if Tcw is None else Tcw.to(X.dtype).contiguous()
        out = torch.empty((B, H, W, 2), dtype=X.dtype, device=X.device)
        with torch.cuda.device(X.device):
            _C.check(_C.lib.psfm_project(B, H, W, _real(X), _C.ptr(X), _C.ptr(K), _C.ptr(Tcw), _C.ptr(out),
                                         _C.stream_ptr(X.device)))
        ctx.save_for_backward(X, K, Tcw)
        return out

    @staticmethod
    def backward(ctx, g):
        X, K, Tcw = ctx.saved_tensors
        g = g.contiguous()
        B, _, H, W = X.shape
        dX = torch.empty_like(X)
        d_T = None if Tcw is None else torch.empty((B, 4, 4), dtype=X.dtype, device=X.device)
        ws, n = _ws(B, H * W, X.device)
        with torch.cuda.device(X.device):
            _C.check(_C.lib.psfm_project_backward(B, H, W, _real(X), _C.ptr(X), _C.ptr(K), _C.ptr(Tcw), _C.ptr(g),
                                                  _C.ptr(dX), _C.ptr(d_T), _C.ptr(ws), n, _C.stream_ptr(X.device)))
        return dX, None, d_T


def project_op(X, K, Tcw=None):
    """Camera.project: points [B,3,H,W] -> normalised coordinates [B,H,W,2]; differentiable w.r.t.
    the points and the [B,4,4] pose."""
    _real(X)
    _no_grad_supported(K, 'K')
    return _Project.apply(X, K, Tcw)


class _FisheyeReconstruct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depth, intr):
        _C.require_cuda(depth, intr)
        B, _, H, W = depth.shape
        depth = depth.contiguous()
        out = torch.empty((B, 3, H, W), dtype=depth.dtype, device=depth.device)
        with torch.cuda.device(depth.device):
            _C.check(_C.lib.psfm_fisheye_reconstruct(B, H, W, _real(depth), _C.ptr(depth), _C.ptr(intr), _C.ptr(out),
                                                     _C.stream_ptr(depth.device)))
        ctx.save_for_backward(intr)
        return out

    @staticmethod
    def backward(ctx, g):
        intr, = ctx.saved_tensors
        g = g.contiguous()
        B, _, H, W = g.shape
        d_depth = torch.empty((B, 1, H, W), dtype=g.dtype, device=g.device)
        with torch.cuda.device(g.device):
            _C.check(_C.lib.psfm_fisheye_reconstruct_backward(B, H, W, _real(g), _C.ptr(intr), _C.ptr(g), _C.ptr(d_depth),
                                                              _C.stream_ptr(g.device)))
        return d_depth, None


def fisheye_reconstruct_op(depth, intr):
    """FisheyeCamera.reconstruct(frame='c'): depth [B,1,H,W], packed intrinsics [B,11] -> points [B,3,H,W]."""
    _real(depth)
    _no_grad_supported(intr, 'the fisheye intrinsics')
    return _FisheyeReconstruct.apply(depth, intr.to(depth.dtype).contiguous())


class _FisheyeProject(torch.autograd.Function):
    @staticmethod
    def forward(ctx, X, intr, Hn, Wn):
        _C.require_cuda(X, intr)
        B = X.shape[0]
        X = X.contiguous()
        N = X.numel() // (B * 3)
        out = torch.empty(((B,) + tuple(X.shape[2:]) + (2,)), dtype=X.dtype, device=X.device)
        with torch.cuda.device(X.device):
            _C.check(_C.lib.psfm_fisheye_project(B, N, Hn, Wn, _real(X), _C.ptr(X), _C.ptr(intr), _C.ptr(out),
                                                 _C.stream_ptr(X.device)))
        ctx.save_for_backward(X, intr)
        ctx.size = (Hn, Wn)
        return out

    @staticmethod
    def backward(ctx, g):
        X, intr = ctx.saved_tensors
        g = g.contiguous()
        B = X.shape[0]
        N = X.numel() // (B * 3)
        dX = torch.empty_like(X)
        with torch.cuda.device(X.device):
            _C.check(_C.lib.psfm_fisheye_project_backward(B, N, ctx.size[0], ctx.size[1], _real(X), _C.ptr(X), _C.ptr(intr),
                                                          _C.ptr(g), _C.ptr(dX), _C.stream_ptr(X.device)))
        return dX, None, None, None


def fisheye_project_op(X, intr, image_size=None):
    """FisheyeCamera.project(frame='c'): points [B,3,H,W] (or [B,3,N] with image_size=(H, W)) and packed
    intrinsics [B,11] -> normalised coordinates [B,H,W,2] ([B,N,2])."""
    _real(X)
    _no_grad_supported(intr, 'the fisheye intrinsics')
    if X.dim() == 4:
        Hn, Wn = int(X.shape[2]), int(X.shape[3])
    elif X.dim() == 3:
        if image_size is None or image_size[0] is None or image_size[1] is None:
            raise ValueError('image_size must be provided for 3D point cloud projection.')
        Hn, Wn = int(image_size[0]), int(image_size[1])
    else:
        raise ValueError('Input X must be of shape [B,3,H,W] or [B,3,N]')
    assert X.shape[1] == 3
    return _FisheyeProject.apply(X, intr.to(X.dtype).contiguous(), Hn, Wn)


class _GridSample(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, grid, padding):
        _C.require_cuda(image, grid)
        B, ch, Hi, Wi = image.shape
        if ch != 3:
            raise ValueError('image must be [B,3,H,W]')
        if grid.dim() != 4 or grid.shape[0] != B or grid.shape[3] != 2:
            raise ValueError('grid must be [B,Ho,Wo,2]')
        grid = grid.contiguous()
        image = image.to(grid.dtype).contiguous()
        Ho, Wo = int(grid.shape[1]), int(grid.shape[2])
        out = torch.empty((B, 3, Ho, Wo), dtype=grid.dtype, device=grid.device)
        with torch.cuda.device(grid.device):
            _C.check(_C.lib.psfm_grid_sample(B, Hi, Wi, Ho, Wo, _real(grid), padding, _C.ptr(image), _C.ptr(grid),
                                             _C.ptr(out), _C.stream_ptr(grid.device)))
        ctx.save_for_backward(image, grid)
        ctx.padding = padding
        return out

    @staticmethod
    def backward(ctx, g):
        image, grid = ctx.saved_tensors
        g = g.to(grid.dtype).contiguous()
        B, _, Hi, Wi = image.shape
        Ho, Wo = int(grid.shape[1]), int(grid.shape[2])
        d_grid = torch.empty_like(grid)
        with torch.cuda.device(grid.device):
            _C.check(_C.lib.psfm_grid_sample_backward(B, Hi, Wi, Ho, Wo, _real(grid), ctx.padding, _C.ptr(image),
                                                      _C.ptr(grid), _C.ptr(g), _C.ptr(d_grid), _C.stream_ptr(grid.device)))
        return None, d_grid, None


def grid_sample_op(image, grid, padding_mode='zeros'):
    """funct.grid_sample(image, grid, mode='bilinear', padding_mode, align_corners=True) as view_synthesis calls
    it (reference geometry/camera_utils.py:58-59); differentiable w.r.t. the grid."""
    if padding_mode not in _C.PAD:
        raise ValueError('padding_mode must be zeros or border (got %s)' % padding_mode)
    _real(grid)
    _no_grad_supported(image, 'the sampled image')
    return _GridSample.apply(image, grid, _C.PAD[padding_mode])


class _RaysurfProject(torch.autograd.Function):
    @staticmethod
    def forward(ctx, direction, rays, temperature):
        _C.require_cuda(direction, rays)
        dt = _real(direction)
        if direction.dim() != 4 or direction.shape[1] != 3 or rays.shape != direction.shape:
            raise ValueError('direction and rays must both be [B,3,h,w]')
        direction, rays = direction.contiguous(), rays.to(direction.dtype).contiguous()
        B, _, h, w = direction.shape
        coords = torch.empty((B, h, w, 2), dtype=direction.dtype, device=direction.device)
        dhat = torch.empty_like(direction)
        stats = torch.empty_like(coords)
        with torch.cuda.device(direction.device):
            _C.check(_C.lib.psfm_raysurf_project(B, h, w, dt, float(temperature), _C.ptr(direction), _C.ptr(rays),
                                                 _C.ptr(coords), _C.ptr(dhat), _C.ptr(stats),
                                                 _C.stream_ptr(direction.device)))
        ctx.save_for_backward(direction, rays, coords, dhat, stats)
        ctx.temperature = float(temperature)
        return coords

    @staticmethod
    def backward(ctx, g):
        direction, rays, coords, dhat, stats = ctx.saved_tensors
        B, _, h, w = direction.shape
        g = g.to(direction.dtype).contiguous()
        d_dir, d_rays = torch.empty_like(direction), torch.empty_like(rays)
        with torch.cuda.device(direction.device):
            _C.check(_C.lib.psfm_raysurf_project_backward(B, h, w, _real(direction), ctx.temperature, _C.ptr(direction),
                                                          _C.ptr(rays), _C.ptr(coords), _C.ptr(dhat), _C.ptr(stats),
                                                          _C.ptr(g), _C.ptr(d_dir), _C.ptr(d_rays),
                                                          _C.stream_ptr(direction.device)))
        return d_dir, d_rays, None


def raysurf_project_op(direction, rays, temperature):
    """Soft arg-max projection of the ray-surface camera (reference geometry/camera_generic.py:151-199): point
    directions [B,3,h,w] and viewing rays [B,3,h,w] -> (row, col) coordinates [B,h,w,2]; differentiable w.r.t. both."""
    return _RaysurfProject.apply(direction, rays, temperature)


class _TransformPoints(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mat, X):
        _C.require_cuda(mat, X)
        B = X.shape[0]
        X = X.contiguous()
        mat = mat.to(X.dtype).contiguous()
        if tuple(mat.shape) != (B, 4, 4):
            raise ValueError('pose matrices must be [B,4,4] with the batch of the points')
        N = X.numel() // (B * 3)
        out = torch.empty_like(X)
        with torch.cuda.device(X.device):
            _C.check(_C.lib.psfm_transform_points(B, N, _real(X), _C.ptr(mat), _C.ptr(X), _C.ptr(out),
                                                  _C.stream_ptr(X.device)))
        ctx.save_for_backward(mat, X)
        return out

    @staticmethod
    def backward(ctx, g):
        mat, X = ctx.saved_tensors
        g = g.contiguous()
        B = X.shape[0]
        N = X.numel() // (B * 3)
        dX, d_mat = torch.empty_like(X), torch.empty_like(mat)
        ws, n = _ws(B, N, X.device)
        with torch.cuda.device(X.device):
            _C.check(_C.lib.psfm_transform_points_backward(B, N, _real(X), _C.ptr(mat), _C.ptr(X), _C.ptr(g), _C.ptr(dX),
                                                           _C.ptr(d_mat), _C.ptr(ws), n, _C.stream_ptr(X.device)))
        return d_mat, dX


class _PoseFromVec(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vec):
        _C.require_cuda(vec)
        vec = vec.contiguous()
        B = vec.shape[0]
        mat = torch.empty(B, 4, 4, dtype=vec.dtype, device=vec.device)
        with torch.cuda.device(vec.device):
            _C.check(_C.lib.psfm_pose_from_vec(B, _real(vec), _C.ptr(vec), _C.ptr(mat), _C.stream_ptr(vec.device)))
        ctx.save_for_backward(vec)
        return mat

    @staticmethod
    def backward(ctx, g):
        vec, = ctx.saved_tensors
        g = g.contiguous()
        d_vec = torch.empty_like(vec)
        with torch.cuda.device(vec.device):
            _C.check(_C.lib.psfm_pose_from_vec_backward(vec.shape[0], _real(vec), _C.ptr(vec), _C.ptr(g), _C.ptr(d_vec),
                                                        _C.stream_ptr(vec.device)))
        return d_vec


def pose_from_vec_op(vec):
    """Pose.from_vec(vec, 'euler'): [B,6] -> [B,4,4], one kernel each way (no batch-size dependent rounding)."""
    _real(vec)
    if vec.dim() != 2 or vec.shape[1] != 6:
        raise ValueError('pose vectors must be [B,6]')
    return _PoseFromVec.apply(vec)


def transform_points_op(mat, points):
    """Pose.transform_points: R X + t on [B,3,H,W] (or [B,3,N]) points."""
    _real(points)
    return _TransformPoints.apply(mat, points)


MODE_SSIM, MODE_SSIM_DIST, MODE_PHOTO, MODE_L1 = 0, 1, 2, 3


class _PhotoMap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, mode, ssim_w, C1, C2):
        _C.require_cuda(x, y)
        if x.shape != y.shape or x.dim() != 4 or x.shape[1] != 3:
            raise ValueError('SSIM / photometric maps take two [B,3,H,W] images')
        x = x.contiguous()
        y = y.to(x.dtype).contiguous()
        B, _, H, W = x.shape
        out = torch.empty((B, 1 if mode == MODE_PHOTO else 3, H, W), dtype=x.dtype, device=x.device)
        with torch.cuda.device(x.device):
            _C.check(_C.lib.psfm_photometric_map(B, H, W, _real(x), mode, ssim_w, C1, C2, _C.ptr(x), _C.ptr(y),
                                                 _C.ptr(out), _C.stream_ptr(x.device)))
        ctx.save_for_backward(x, y)
        ctx.args = (mode, ssim_w, C1, C2)
        return out

    @staticmethod
    def backward(ctx, g):
        x, y = ctx.saved_tensors
        mode, ssim_w, C1, C2 = ctx.args
        g = g.contiguous()
        B, _, H, W = x.shape
        dx = torch.empty_like(x)
        dy = torch.empty_like(y) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(x.device):
            _C.check(_C.lib.psfm_photometric_map_backward(B, H, W, _real(x), mode, ssim_w, C1, C2, _C.ptr(x), _C.ptr(y),
                                                          _C.ptr(g), _C.ptr(dx), _C.ptr(dy), _C.stream_ptr(x.device)))
        return dx, dy, None, None, None, None


def ssim_op(x, y, C1=1e-4, C2=9e-4, distance=False):
    """SSIM index of two [B,3,H,W] images (3x3 windows, reflection padding), or with
    `distance` the loss map clamp((1 - SSIM) / 2, 0, 1)."""
    _real(x)
    return _PhotoMap.apply(x, y, MODE_SSIM_DIST if distance else MODE_SSIM, 0.0, float(C1), float(C2))


def photometric_map_op(t_est, image, ssim_loss_weight, C1, C2, clip_loss=0.0, mask=None):
    """calc_photometric_loss for one scale (reference :218-267): [B,1,H,W] map
    alpha * SSIM-distance.mean(1) + (1 - alpha) * L1.mean(1), or the 3-channel L1 when alpha == 0.
    The rarely used clip / mask options are applied on top of the kernel's map: the clip threshold is
    a detached scalar statistic exactly as in the reference (:249-253)."""
    _real(t_est)
    alpha = float(ssim_loss_weight)
    p = _PhotoMap.apply(t_est, image, MODE_PHOTO if alpha > 0.0 else MODE_L1, alpha, float(C1), float(C2))
    if clip_loss > 0.0:
        d = p.detach()
        p = torch.clamp(p, max=float(d.mean() + clip_loss * d.std()))
    if mask is not None:
        p = p * mask
    return p


class _ReduceCandidates(torch.autograd.Function):
    @staticmethod
    def forward(ctx, op, *cands):
        _C.require_cuda(*cands)
        K = len(cands)
        cands = [c.contiguous() for c in cands]
        B, Cg, H, W = cands[0].shape
        for c in cands:
            if c.shape != cands[0].shape or c.dtype != cands[0].dtype:
                raise ValueError('photometric candidates must share one shape and dtype')
        dev, dt = cands[0].device, cands[0].dtype
        out = torch.empty(1, dtype=dt, device=dev)
        argmin = torch.empty((B, H, W), dtype=torch.int32, device=dev) if op == _C.REDUCE['min'] else None
        ptrs = (C.c_void_p * K)(*[c.data_ptr() for c in cands])
        ws, n = _ws(B, H * W, dev)
        with torch.cuda.device(dev):
            _C.check(_C.lib.psfm_reduce_candidates(K, B, Cg, H, W, _real(cands[0]), op, ptrs, _C.ptr(out),
                                                   _C.ptr(argmin), _C.ptr(ws), n, _C.stream_ptr(dev)))
        ctx.argmin, ctx.meta = argmin, (K, B, Cg, H, W, dt, dev, op)
        return out.view(())

    @staticmethod
    def backward(ctx, g):
        K, B, Cg, H, W, dt, dev, op = ctx.meta
        g = g.to(dt).contiguous().view(1)
        grads = [torch.empty((B, Cg, H, W), dtype=dt, device=dev) for _ in range(K)]
        ptrs = (C.c_void_p * K)(*[t.data_ptr() for t in grads])
        with torch.cuda.device(dev):
            _C.check(_C.lib.psfm_reduce_candidates_backward(K, B, Cg, H, W, _C.F64 if dt == torch.float64 else _C.F32,
                                                            op, _C.ptr(ctx.argmin), _C.ptr(g), ptrs, _C.stream_ptr(dev)))
        return (None,) + tuple(grads)


def reduce_candidates_op(candidates, op):
    """reduce_photometric_loss for one scale: 'min' -> mean of the per-pixel min over all candidate
    channels (first index wins ties); 'mean' -> average of the candidates' means."""
    if op not in _C.REDUCE:
        raise NotImplementedError('Unknown photometric_reduce_op: {}'.format(op))
    if len(candidates) > 16:
        raise ValueError('at most 16 photometric candidates per scale')
    _real(candidates[0])
    return _ReduceCandidates.apply(_C.REDUCE[op], *candidates)


class _AbsMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _C.require_cuda(x)
        x = x.contiguous()
        out = torch.empty(1, dtype=x.dtype, device=x.device)
        ws, n = _ws(1, x.numel(), x.device)
        with torch.cuda.device(x.device):
            _C.check(_C.lib.psfm_abs_mean(x.numel(), _real(x), _C.ptr(x), _C.ptr(out), _C.ptr(ws), n,
                                          _C.stream_ptr(x.device)))
        ctx.save_for_backward(x)
        return out.view(())

    @staticmethod
    def backward(ctx, g):
        x, = ctx.saved_tensors
        g = g.to(x.dtype).contiguous().view(1)
        dx = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _C.check(_C.lib.psfm_abs_mean_backward(x.numel(), _real(x), _C.ptr(x), _C.ptr(g), _C.ptr(dx),
                                                   _C.stream_ptr(x.device)))
        return dx


def abs_mean_op(x):
    """x.abs().mean() with a deterministic reduction."""
    _real(x)
    return _AbsMean.apply(x)


class _Normalize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, d):
        _C.require_cuda(d)
        d = d.contiguous()
        B = d.shape[0]
        N = d.numel() // B
        out = torch.empty_like(d)
        mean = torch.empty(B, dtype=d.dtype, device=d.device)
        ws, n = _ws(B, N, d.device)
        with torch.cuda.device(d.device):
            _C.check(_C.lib.psfm_normalize(B, N, _real(d), _C.ptr(d), _C.ptr(out), _C.ptr(mean), _C.ptr(ws), n,
                                           _C.stream_ptr(d.device)))
        ctx.save_for_backward(d, mean)
        return out

    @staticmethod
    def backward(ctx, g):
        d, mean = ctx.saved_tensors
        g = g.contiguous()
        B = d.shape[0]
        N = d.numel() // B
        dd = torch.empty_like(d)
        ws, n = _ws(B, N, d.device)
        with torch.cuda.device(d.device):
            _C.check(_C.lib.psfm_normalize_backward(B, N, _real(d), _C.ptr(d), _C.ptr(mean), _C.ptr(g), _C.ptr(dd),
                                                    _C.ptr(ws), n, _C.stream_ptr(d.device)))
        return dd


def normalize_op(inv_depth):
    """inv_depths_normalize for one map: divide by the per-image mean clamped at 1e-6."""
    _real(inv_depth)
    if inv_depth.dim() != 4 or inv_depth.shape[1] != 1:
        raise ValueError('inverse depth maps must be [B,1,H,W]')
    return _Normalize.apply(inv_depth)


class _SmoothMaps(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dn, image):
        _C.require_cuda(dn, image)
        dn = dn.contiguous()
        image = image.to(dn.dtype).contiguous()
        B, _, H, W = dn.shape
        if tuple(image.shape) != (B, 3, H, W):
            raise ValueError('image must be [B,3,H,W] matching the inverse depth map')
        sx = torch.empty((B, 1, H, W - 1), dtype=dn.dtype, device=dn.device)
        sy = torch.empty((B, 1, H - 1, W), dtype=dn.dtype, device=dn.device)
        with torch.cuda.device(dn.device):
            _C.check(_C.lib.psfm_smoothness_maps(B, H, W, _real(dn), _C.ptr(dn), _C.ptr(image), _C.ptr(sx), _C.ptr(sy),
                                                 _C.stream_ptr(dn.device)))
        ctx.save_for_backward(image)
        return sx, sy

    @staticmethod
    def backward(ctx, gsx, gsy):
        image, = ctx.saved_tensors
        B, _, H, W = image.shape
        gsx, gsy = gsx.contiguous(), gsy.contiguous()
        ddn = torch.empty((B, 1, H, W), dtype=image.dtype, device=image.device)
        with torch.cuda.device(image.device):
            _C.check(_C.lib.psfm_smoothness_maps_backward(B, H, W, _real(image), _C.ptr(image), _C.ptr(gsx), _C.ptr(gsy),
                                                          _C.ptr(ddn), _C.stream_ptr(image.device)))
        return ddn, None


def smoothness_op(inv_depth, image):
    """calc_smoothness for one scale: (smoothness_x [B,1,H,W-1], smoothness_y [B,1,H-1,W]) of the
    mean-normalised inverse depth, weighted by exp(-mean_c |image gradient|)."""
    _no_grad_supported(image, 'the image of calc_smoothness')
    return _SmoothMaps.apply(normalize_op(inv_depth), image)


SUP_METHODS = {'l1': 0, 'mse': 1, 'silog': 2, 'ssi-silog': 3}
SUP_STATE = ('sum_abs', 'sum_sq', 'sum_log', 'sum_log_sq', 'valid', 'sum_diff', 'count', 'loss', 'ssi', 'silog', 'ssi_mean',
             'ssi_var', 'silog1', 'silog2', 'silog_sign', 'constant')


class _SupervisedLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, gt, mask, method, sparse, hp):
        import ctypes
        _C.require_cuda(pred, gt, mask)
        if pred.shape != gt.shape:
            raise ValueError('prediction and ground truth must have the same shape')
        pred = pred.contiguous()
        gt = gt.to(pred.dtype).contiguous()
        if pred.data_ptr() % 16:                 # the kernels use 16-byte vector loads
            pred = pred.clone()
        if gt.data_ptr() % 16:
            gt = gt.clone()
        if mask is not None:
            if mask.numel() != pred.numel():
                raise ValueError('the mask must have one element per pixel')
            mask = (mask if mask.dtype == torch.uint8 else (mask > 0).to(torch.uint8)).contiguous()
        N = pred.numel()
        out = torch.empty(1, dtype=pred.dtype, device=pred.device)
        state = torch.empty(16, dtype=torch.float64, device=pred.device)
        ws, n = _ws(1, N, pred.device)
        hp_c = (ctypes.c_double * 6)(*[float(x) for x in hp])
        with torch.cuda.device(pred.device):
            _C.check(_C.lib.psfm_supervised_loss(method, int(sparse), N, _real(pred), _C.ptr(pred), _C.ptr(gt), _C.ptr(mask),
                                                 hp_c, _C.ptr(out), _C.ptr(state), _C.ptr(ws), n, _C.stream_ptr(pred.device)))
        ctx.save_for_backward(pred, gt, mask, state)
        ctx.cfg = (method, int(sparse), hp_c)
        ctx.mark_non_differentiable(state)
        return out, state

    @staticmethod
    def backward(ctx, g, _gs=None):
        pred, gt, mask, state = ctx.saved_tensors
        method, sparse, hp_c = ctx.cfg
        g = g.to(pred.dtype).contiguous()
        d_pred = torch.empty_like(pred)
        with torch.cuda.device(pred.device):
            _C.check(_C.lib.psfm_supervised_loss_backward(method, sparse, pred.numel(), _real(pred), _C.ptr(pred), _C.ptr(gt),
                                                          _C.ptr(mask), hp_c, _C.ptr(state), _C.ptr(g), _C.ptr(d_pred),
                                                          _C.stream_ptr(pred.device)))
        return d_pred, None, None, None, None, None


def supervised_loss_op(pred, gt, method, sparse=False, mask=None, ratio2=0.85, alpha=0.85, ssi_weight=0.7,
                       silog_weight=0.3, clamp_min=1e-3, clamp_max=100.0):
    """One scale of SupervisedLoss.calculate_loss (reference losses/supervised_loss.py:292-417) with the l1 / mse /
    silog / ssi-silog loss functions.  Returns (loss [], state [16] float64, fields named by SUP_STATE);
    differentiable w.r.t. `pred`."""
    if method not in SUP_METHODS:
        raise NotImplementedError('supervised loss function %r is not provided (%s)' % (method, ', '.join(SUP_METHODS)))
    _real(pred)
    _no_grad_supported(gt, 'the ground truth')
    hp = (ratio2, alpha, ssi_weight, silog_weight, clamp_min, clamp_max)
    out, state = _SupervisedLoss.apply(pred, gt, mask, SUP_METHODS[method], bool(sparse), hp)
    return out[0], state


class _DualHeadLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ip, fp, gt, hp):
        import ctypes
        _C.require_cuda(ip, fp, gt)
        if ip.shape != fp.shape or ip.shape != gt.shape:
            raise ValueError('integer / fractional predictions and the ground truth must have the same shape')
        ip = ip.contiguous()
        fp = fp.to(ip.dtype).contiguous()
        gt = gt.to(ip.dtype).contiguous()
        N = ip.numel()
        out = torch.empty(4, dtype=ip.dtype, device=ip.device)
        state = torch.empty(8, dtype=torch.float64, device=ip.device)
        ws, n = _ws(1, N, ip.device)
        hp_c = (ctypes.c_double * 5)(*[float(x) for x in hp])
        with torch.cuda.device(ip.device):
            _C.check(_C.lib.psfm_dual_head_loss(N, _real(ip), _C.ptr(ip), _C.ptr(fp), _C.ptr(gt), hp_c, _C.ptr(out), _C.ptr(state),
                                                _C.ptr(ws), n, _C.stream_ptr(ip.device)))
        ctx.save_for_backward(ip, fp, gt, state)
        ctx.hp = hp_c
        loss, parts = out[0], out[1:]
        ctx.mark_non_differentiable(parts, state)
        return loss, parts, state

    @staticmethod
    def backward(ctx, g, _gp=None, _gs=None):
        ip, fp, gt, state = ctx.saved_tensors
        g = g.to(ip.dtype).reshape(1).contiguous()
        d_ip, d_fp = torch.empty_like(ip), torch.empty_like(fp)
        with torch.cuda.device(ip.device):
            _C.check(_C.lib.psfm_dual_head_loss_backward(ip.numel(), _real(ip), _C.ptr(ip), _C.ptr(fp), _C.ptr(gt), ctx.hp,
                                                         _C.ptr(state), _C.ptr(g), _C.ptr(d_ip), _C.ptr(d_fp),
                                                         _C.stream_ptr(ip.device)))
        return d_ip, d_fp, None, None


def dual_head_loss_op(integer_pred, fractional_pred, depth_gt, max_depth, min_depth, integer_weight, fractional_weight,
                      consistency_weight):
    """DualHeadDepthLoss.forward (reference losses/dual_head_depth_loss.py:83-201).  Returns (loss [], parts [3] =
    integer / fractional / consistency loss, state [8] float64); differentiable w.r.t. both predictions."""
    _real(integer_pred)
    _no_grad_supported(depth_gt, 'the ground truth')
    return _DualHeadLoss.apply(integer_pred, fractional_pred, depth_gt,
                               (max_depth, min_depth, integer_weight, fractional_weight, consistency_weight))


def depth_metrics_op(gt, pred, scale_mode, garg_crop, use_gt_scale, min_depth, max_depth):
    """compute_depth_metrics (reference utils/depth.py:258-447).  Returns (metrics [7] of gt's dtype, state [B,12] float64)."""
    _C.require_cuda(gt, pred)
    _real(gt)
    if gt.dim() != 4 or pred.dim() != 4 or gt.shape[1] != 1 or pred.shape[1] != 1 or gt.shape[0] != pred.shape[0]:
        raise ValueError('compute_depth_metrics: gt [B,1,H,W] and pred [B,1,h,w] expected, got %s and %s'
                         % (tuple(gt.shape), tuple(pred.shape)))
    gt = gt.detach().contiguous()
    pred = pred.detach().to(gt.dtype).contiguous()
    B, _, H, W = gt.shape
    h, w = pred.shape[-2:]
    out = torch.empty(7, dtype=gt.dtype, device=gt.device)
    state = torch.empty(B, 12, dtype=torch.float64, device=gt.device)
    n = _C.lib.psfm_depth_metrics_workspace_bytes(B, H, W)
    ws = torch.empty(n, dtype=torch.uint8, device=gt.device)
    with torch.cuda.device(gt.device):
        _C.check(_C.lib.psfm_depth_metrics(B, H, W, h, w, _real(gt), int(scale_mode), int(bool(garg_crop)), int(bool(use_gt_scale)),
                                           float(min_depth), float(max_depth), _C.ptr(gt), _C.ptr(pred), _C.ptr(out),
                                           _C.ptr(state), _C.ptr(ws), n, _C.stream_ptr(gt.device)))
    return out, state


class _SobelGradientLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, gt, mask, scales):
        _C.require_cuda(pred, gt, mask)
        if pred.dim() != 4 or pred.shape[1] != 1 or pred.shape != gt.shape:
            raise ValueError('[B,1,H,W] prediction and ground truth of the same shape expected')
        pred = pred.contiguous()
        gt = gt.to(pred.dtype).contiguous()
        if mask is not None:
            if mask.numel() != pred.numel():
                raise ValueError('the mask must have one element per pixel')
            mask = (mask if mask.dtype == torch.uint8 else (mask > 0).to(torch.uint8)).contiguous()
        B, _, H, W = pred.shape
        dev = pred.device
        out = torch.empty(1, dtype=pred.dtype, device=dev)
        state = torch.empty(20, dtype=torch.float64, device=dev)
        codes = torch.empty(int(_C.lib.psfm_sobel_gradient_code_bytes(B, H, W, scales)), dtype=torch.uint8, device=dev)
        n = int(_C.lib.psfm_sobel_gradient_workspace_bytes(B, H, W, scales))
        ws = torch.empty(n, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _C.check(_C.lib.psfm_sobel_gradient_loss(B, H, W, _real(pred), scales, _C.ptr(pred), _C.ptr(gt), _C.ptr(mask),
                                                     _C.ptr(out), _C.ptr(state), _C.ptr(codes), _C.ptr(ws), n,
                                                     _C.stream_ptr(dev)))
        ctx.save_for_backward(pred, gt, mask, state, codes)
        ctx.scales = scales
        ctx.mark_non_differentiable(state)
        return out[0], state

    @staticmethod
    def backward(ctx, g, _gs=None):
        pred, gt, mask, state, codes = ctx.saved_tensors
        B, _, H, W = pred.shape
        g = g.to(pred.dtype).reshape(1).contiguous()
        d_pred = torch.empty_like(pred)
        with torch.cuda.device(pred.device):
            _C.check(_C.lib.psfm_sobel_gradient_loss_backward(B, H, W, _real(pred), ctx.scales, _C.ptr(pred), _C.ptr(gt),
                                                              _C.ptr(mask), _C.ptr(state), _C.ptr(codes), _C.ptr(g),
                                                              _C.ptr(d_pred), _C.stream_ptr(pred.device)))
        return d_pred, None, None, None


def sobel_gradient_loss_op(pred_inv_depth, gt_inv_depth, mask=None, scales=4):
    """SSISilogLoss.compute_gradient_loss on the sparse caller's filled maps (reference losses/ssi_silog_loss.py:116-174).
    Returns (loss [], state [20] float64); differentiable w.r.t. the prediction."""
    _real(pred_inv_depth)
    _no_grad_supported(gt_inv_depth, 'the ground truth')
    return _SobelGradientLoss.apply(pred_inv_depth, gt_inv_depth, mask, int(scales))
