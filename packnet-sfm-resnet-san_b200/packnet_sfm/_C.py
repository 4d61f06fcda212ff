"""ctypes binding of libpsfm_b200.so (C ABI declared in include/psfm.h).

There is no CPU fallback and no alternative backend: if the shared library is missing the
import of this module raises, and every wrapper refuses non-CUDA tensors.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('PSFM_B200_LIB', os.path.join(os.path.dirname(_HERE), 'lib', 'libpsfm_b200.so'))

MAX_SCALES = 4
MAX_CTX = 8
F32, BF16, F64, U8 = 0, 1, 2, 3
PAD = {'zeros': 0, 'border': 1}
REDUCE = {'min': 0, 'mean': 1}
CAM_PINHOLE, CAM_FISHEYE = 0, 1

EXPORTS = ('psfm_abi_version', 'psfm_last_error', 'psfm_launch_count', 'psfm_profile_enable', 'psfm_profile_read',
           'psfm_loss_workspace_bytes', 'psfm_loss_workspace_bytes_io', 'psfm_loss_forward',
           'psfm_loss_backward', 'psfm_loss_forward_backward', 'psfm_scale_inplace', 'psfm_warp_forward',
           'psfm_warp_workspace_bytes', 'psfm_warp_backward', 'psfm_resize', 'psfm_resize_backward',
           'psfm_block_workspace_bytes', 'psfm_reconstruct', 'psfm_reconstruct_backward', 'psfm_project',
           'psfm_project_backward', 'psfm_transform_points', 'psfm_transform_points_backward',
           'psfm_pose_from_vec', 'psfm_pose_from_vec_backward',
           'psfm_fisheye_reconstruct', 'psfm_fisheye_reconstruct_backward', 'psfm_fisheye_project',
           'psfm_fisheye_project_backward', 'psfm_grid_sample', 'psfm_grid_sample_backward',
           'psfm_raysurf_project', 'psfm_raysurf_project_backward',
           'psfm_supervised_loss', 'psfm_supervised_loss_backward', 'psfm_dual_head_loss', 'psfm_dual_head_loss_backward',
           'psfm_depth_metrics_workspace_bytes', 'psfm_depth_metrics',
           'psfm_sobel_gradient_code_bytes', 'psfm_sobel_gradient_workspace_bytes', 'psfm_sobel_gradient_loss',
           'psfm_sobel_gradient_loss_backward',
           'psfm_photometric_map', 'psfm_photometric_map_backward', 'psfm_normalize', 'psfm_normalize_backward',
           'psfm_smoothness_maps', 'psfm_smoothness_maps_backward', 'psfm_reduce_candidates',
           'psfm_reduce_candidates_backward', 'psfm_abs_mean', 'psfm_abs_mean_backward')


class LossCfg(C.Structure):
    _fields_ = [('B', C.c_int32), ('C', C.c_int32), ('n', C.c_int32), ('H', C.c_int32), ('W', C.c_int32),
                ('Hs', C.c_int32 * MAX_SCALES), ('Ws', C.c_int32 * MAX_SCALES),
                ('Hd', C.c_int32 * MAX_SCALES), ('Wd', C.c_int32 * MAX_SCALES),
                ('img_dtype', C.c_int32), ('real_dtype', C.c_int32), ('padding', C.c_int32),
                ('reduce_op', C.c_int32), ('automask', C.c_int32), ('premap', C.c_int32), ('camera_model', C.c_int32),
                ('ssim_w', C.c_double), ('smooth_w', C.c_double), ('C1', C.c_double), ('C2', C.c_double),
                ('clip_loss', C.c_double), ('min_depth', C.c_double), ('max_depth', C.c_double)]


class LossIO(C.Structure):
    _fields_ = [('image', C.c_void_p * MAX_SCALES),
                ('ctx', (C.c_void_p * MAX_CTX) * MAX_SCALES),
                ('inv_depth', C.c_void_p * MAX_SCALES),
                ('mask', C.c_void_p * MAX_SCALES),
                ('K', C.c_void_p), ('ref_K', C.c_void_p),
                ('pose', C.c_void_p * MAX_CTX),
                ('sel', C.c_void_p * MAX_SCALES),
                ('d_inv_depth', C.c_void_p * MAX_SCALES),
                ('d_pose', C.c_void_p * MAX_CTX),
                ('dbg', C.c_void_p * MAX_SCALES)]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            'libpsfm_b200.so not found at %s -- build it with `python __graft_entry__.py build` '
            '(or `make -C packnet-sfm-resnet-san_b200/csrc`). This package has no CPU or PyTorch fallback.' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, sz, i = C.c_void_p, C.c_size_t, C.c_int
    lib.psfm_abi_version.restype = i
    lib.psfm_last_error.restype = C.c_char_p
    lib.psfm_launch_count.restype = C.c_ulonglong
    lib.psfm_profile_enable.restype = i
    lib.psfm_profile_enable.argtypes = [i]
    lib.psfm_profile_read.restype = i
    lib.psfm_profile_read.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_int), i]
    lib.psfm_loss_workspace_bytes.restype = sz
    lib.psfm_loss_workspace_bytes.argtypes = [C.POINTER(LossCfg)]
    lib.psfm_loss_workspace_bytes_io.restype = sz
    lib.psfm_loss_workspace_bytes_io.argtypes = [C.POINTER(LossCfg), C.POINTER(LossIO)]
    lib.psfm_loss_forward.restype = i
    lib.psfm_loss_forward.argtypes = [C.POINTER(LossCfg), C.POINTER(LossIO), vp, sz, vp, vp]
    lib.psfm_loss_backward.restype = i
    lib.psfm_loss_backward.argtypes = [C.POINTER(LossCfg), C.POINTER(LossIO), vp, sz, vp, vp]
    lib.psfm_loss_forward_backward.restype = i
    lib.psfm_loss_forward_backward.argtypes = [C.POINTER(LossCfg), C.POINTER(LossIO), vp, sz, vp, vp]
    lib.psfm_scale_inplace.restype = i
    lib.psfm_scale_inplace.argtypes = [vp, vp, sz, i, vp]
    lib.psfm_warp_forward.restype = i
    lib.psfm_warp_forward.argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp]
    lib.psfm_warp_workspace_bytes.restype = sz
    lib.psfm_warp_workspace_bytes.argtypes = [i, i, i, i]
    lib.psfm_warp_backward.restype = i
    lib.psfm_warp_backward.argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, vp]
    lib.psfm_resize.restype = i
    lib.psfm_resize.argtypes = [i, i, i, i, i, i, i, vp, vp, vp]
    lib.psfm_resize_backward.restype = i
    lib.psfm_resize_backward.argtypes = [i, i, i, i, i, i, i, vp, vp, vp]
    d = C.c_double
    lib.psfm_block_workspace_bytes.restype = sz
    lib.psfm_block_workspace_bytes.argtypes = [i, sz]
    for name in ('psfm_sobel_gradient_code_bytes', 'psfm_sobel_gradient_workspace_bytes'):
        getattr(lib, name).restype = sz
        getattr(lib, name).argtypes = [i, i, i, i]
    lib.psfm_depth_metrics_workspace_bytes.restype = sz
    lib.psfm_depth_metrics_workspace_bytes.argtypes = [i, i, i]
    for name, args in (
            ('psfm_reconstruct', [i, i, i, i, vp, vp, vp, vp, vp]),
            ('psfm_reconstruct_backward', [i, i, i, i, vp, vp, vp, vp, vp]),
            ('psfm_project', [i, i, i, i, vp, vp, vp, vp, vp]),
            ('psfm_project_backward', [i, i, i, i, vp, vp, vp, vp, vp, vp, vp, sz, vp]),
            ('psfm_pose_from_vec', [i, i, vp, vp, vp]),
            ('psfm_pose_from_vec_backward', [i, i, vp, vp, vp, vp]),
            ('psfm_fisheye_reconstruct', [i, i, i, i, vp, vp, vp, vp]),
            ('psfm_fisheye_reconstruct_backward', [i, i, i, i, vp, vp, vp, vp]),
            ('psfm_fisheye_project', [i, sz, i, i, i, vp, vp, vp, vp]),
            ('psfm_fisheye_project_backward', [i, sz, i, i, i, vp, vp, vp, vp, vp]),
            ('psfm_raysurf_project', [i, i, i, i, d, vp, vp, vp, vp, vp, vp]),
            ('psfm_raysurf_project_backward', [i, i, i, i, d, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
            ('psfm_grid_sample', [i, i, i, i, i, i, i, vp, vp, vp, vp]),
            ('psfm_grid_sample_backward', [i, i, i, i, i, i, i, vp, vp, vp, vp, vp]),
            ('psfm_supervised_loss', [i, i, sz, i, vp, vp, vp, C.POINTER(d), vp, vp, vp, sz, vp]),
            ('psfm_supervised_loss_backward', [i, i, sz, i, vp, vp, vp, C.POINTER(d), vp, vp, vp, vp]),
            ('psfm_dual_head_loss', [sz, i, vp, vp, vp, C.POINTER(d), vp, vp, vp, sz, vp]),
            ('psfm_dual_head_loss_backward', [sz, i, vp, vp, vp, C.POINTER(d), vp, vp, vp, vp, vp]),
            ('psfm_sobel_gradient_loss', [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, sz, vp]),
            ('psfm_sobel_gradient_loss_backward', [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp]),
            ('psfm_depth_metrics', [i, i, i, i, i, i, i, i, i, d, d, vp, vp, vp, vp, vp, sz, vp]),
            ('psfm_transform_points', [i, sz, i, vp, vp, vp, vp]),
            ('psfm_transform_points_backward', [i, sz, i, vp, vp, vp, vp, vp, vp, sz, vp]),
            ('psfm_photometric_map', [i, i, i, i, i, d, d, d, vp, vp, vp, vp]),
            ('psfm_photometric_map_backward', [i, i, i, i, i, d, d, d, vp, vp, vp, vp, vp, vp]),
            ('psfm_normalize', [i, sz, i, vp, vp, vp, vp, sz, vp]),
            ('psfm_normalize_backward', [i, sz, i, vp, vp, vp, vp, vp, sz, vp]),
            ('psfm_smoothness_maps', [i, i, i, i, vp, vp, vp, vp, vp]),
            ('psfm_smoothness_maps_backward', [i, i, i, i, vp, vp, vp, vp, vp]),
            ('psfm_reduce_candidates', [i, i, i, i, i, i, i, C.POINTER(vp), vp, vp, vp, sz, vp]),
            ('psfm_reduce_candidates_backward', [i, i, i, i, i, i, i, vp, vp, C.POINTER(vp), vp]),
            ('psfm_abs_mean', [sz, i, vp, vp, vp, sz, vp]),
            ('psfm_abs_mean_backward', [sz, i, vp, vp, vp, vp])):
        fn = getattr(lib, name)
        fn.restype = i
        fn.argtypes = args
    if lib.psfm_abi_version() != 4:
        raise ImportError('libpsfm_b200.so ABI version mismatch: %d' % lib.psfm_abi_version())
    return lib


lib = _load()


class PsfmError(RuntimeError):
    pass


def check(rc):
    """Turn a psfm_status into a Python exception (ValueError for bad arguments, like the
    reference's asserts / ValueErrors; RuntimeError for CUDA failures)."""
    if rc == 0:
        return
    msg = lib.psfm_last_error().decode()
    if rc == -1:
        raise ValueError(msg)
    raise PsfmError(msg)


def dtype_code(t):
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    if t.dtype == torch.float64:
        return F64
    if t.dtype == torch.uint8:
        return U8
    raise ValueError('unsupported dtype %s (float32, bfloat16, float64 or uint8 image bytes)' % t.dtype)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError('packnet_sfm (B200 build) runs on CUDA tensors only; got a %s tensor. '
                               'There is no CPU fallback.' % t.device)


def stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


PROFILE_TAGS = {0: 'mean', 1: 'loss_fwd', 2: 'loss_bwd', 3: 'finish', 4: 'loss_fused', 5: 'other'}


def profile_read(max_records=65536):
    """[(tag_name, ms), ...] of the kernels bracketed since profiling was enabled (synchronises)."""
    ms = (C.c_float * max_records)()
    tags = (C.c_int * max_records)()
    n = lib.psfm_profile_read(ms, tags, max_records)
    return [(PROFILE_TAGS.get(tags[k], str(tags[k])), ms[k]) for k in range(n)]
