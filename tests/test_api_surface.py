"""CPU tests of the drop-in boundary: call signatures against the reference's (recorded by
oracle/gen_signatures.py), C-ABI exports against include/psfm.h, argument validation, and the
host-side helpers (intrinsics scaling, pose algebra, progressive scaling).  No GPU, no compute calls
into the library."""
import ctypes
import inspect
import json
import os
import re

import numpy as np
import pytest
import torch

import photometric_oracle as O
from golden_cases import GOLDEN_DIR, load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _describe(fn):
    fn = getattr(fn, '__wrapped__', fn)
    fn = getattr(fn, '__func__', fn)
    return [[p.name, str(p.kind), None if p.default is inspect.Parameter.empty else repr(p.default)]
            for p in inspect.signature(fn).parameters.values()]


# deliberate, documented differences from the fork's signatures
RENAMED = {'packnet_sfm.losses.multiview_photometric_loss:MultiViewPhotometricLoss.forward':
           {'intrinsics': 'K', 'ref_intrinsics': 'ref_K'},
           'packnet_sfm.losses.multiview_photometric_loss:MultiViewPhotometricLoss.warp_ref_image':
           {'intrinsics': 'K', 'ref_intrinsics': 'ref_K', 'poses': 'pose'}}
LEADING_ONLY = {}


def test_signatures_match_reference():
    import importlib
    with open(os.path.join(GOLDEN_DIR, 'api_signatures.json')) as f:
        ref = json.load(f)
    assert len(ref) >= 40
    for key, want in ref.items():
        mod, name = key.split(':')
        obj = importlib.import_module(mod)
        for part in name.split('.'):
            obj = inspect.getattr_static(obj, part) if inspect.isclass(obj) else getattr(obj, part)
        got = _describe(obj)
        ren = RENAMED.get(key, {})
        want = [[ren.get(n, n), k, d] for n, k, d in want][:LEADING_ONLY.get(key)]
        # every reference parameter is present at the same position with the same kind; a default the
        # reference has must be kept (a parameter the reference requires may gain one)
        assert len(got) >= len(want), (key, got, want)
        for g, w in zip(got, want):
            assert g[:2] == w[:2], (key, g, w)
            assert w[2] is None or g[2] == w[2], (key, g, w)
        # extra parameters (keyword aliases) must be optional and come after the reference's
        for g in got[len(want):]:
            assert g[2] is not None or 'VAR' in g[1], (key, g)


def test_north_star_aliases_exist():
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.utils import depth
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    assert Pose.transform_pairs is Pose.transform_points
    assert depth.inv_depths_smoothness is depth.calc_smoothness
    params = inspect.signature(MultiViewPhotometricLoss.forward).parameters
    assert list(params)[1:7] == ['image', 'context', 'inv_depths', 'K', 'ref_K', 'poses']
    assert 'intrinsics' in params and 'ref_intrinsics' in params      # the fork's keyword names still work
    assert len(list(MultiViewPhotometricLoss().state_dict())) == 0    # no parameters / buffers (checkpoints load unchanged)


def _declared_functions():
    text = open(os.path.join(ROOT, 'include', 'psfm.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(psfm_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    from packnet_sfm import _C
    names = _declared_functions()
    assert len(names) >= 30
    lib = ctypes.CDLL(_C.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), 'libpsfm_b200.so does not export %s' % n
    assert sorted(_C.EXPORTS) == names, 'ctypes table and include/psfm.h disagree'
    assert lib.psfm_abi_version() == 4


def test_abi_argument_validation_without_gpu():
    """Invalid configurations are rejected before any CUDA call, with the reference's messages."""
    from packnet_sfm import _C
    cfg = _C.LossCfg()
    cfg.B, cfg.C, cfg.n, cfg.H, cfg.W = 1, 2, 1, 8, 8
    cfg.Hs[0], cfg.Ws[0] = 8, 8
    cfg.reduce_op, cfg.automask = _C.REDUCE['mean'], 1
    assert _C.lib.psfm_loss_workspace_bytes(ctypes.byref(cfg)) == 0
    assert b'For automasking only the min photometric_reduce_op is supported' in _C.lib.psfm_last_error()
    cfg.automask, cfg.reduce_op = 0, 7
    assert _C.lib.psfm_loss_workspace_bytes(ctypes.byref(cfg)) == 0
    assert b'Unknown photometric_reduce_op' in _C.lib.psfm_last_error()
    cfg.reduce_op, cfg.C = 0, 9
    assert _C.lib.psfm_loss_workspace_bytes(ctypes.byref(cfg)) == 0
    cfg.C = 2
    assert _C.lib.psfm_loss_workspace_bytes(ctypes.byref(cfg)) > 0
    assert _C.lib.psfm_reconstruct(1, 4, 4, 5, None, None, None, None, None) == -1      # bad dtype
    assert _C.lib.psfm_block_workspace_bytes(0, 10) == 0
    # compute_depth_metrics: argument checks come before any CUDA call
    assert _C.lib.psfm_depth_metrics_workspace_bytes(0, 8, 8) == 0
    n = _C.lib.psfm_depth_metrics_workspace_bytes(2, 8, 8)
    assert n > 0
    one = ctypes.c_void_p(256)          # never dereferenced: every call below is refused
    bad = lambda *a: _C.lib.psfm_depth_metrics(*a)
    assert bad(2, 8, 8, 8, 8, 0, 2, 0, 1, 0.0, 80.0, one, one, one, one, one, n, None) == -1
    assert b'not implemented' in _C.lib.psfm_last_error()
    assert bad(2, 8, 8, 9, 8, 0, 1, 0, 1, 0.0, 80.0, one, one, one, one, one, n, None) == -1       # top-center, pred larger
    assert bad(2, 8, 8, 8, 8, 0, 0, 0, 1, 0.0, 80.0, one, one, one, one, one, n - 1, None) == -3   # workspace
    assert bad(2, 8, 8, 8, 8, 7, 0, 0, 1, 0.0, 80.0, one, one, one, one, one, n, None) == -1       # dtype
    assert bad(2, 8, 8, 8, 8, 0, 0, 0, 1, 0.0, 80.0, None, one, one, one, one, n, None) == -1      # null pointer
    with pytest.raises(ValueError):
        _C.check(-1)
    with pytest.raises(_C.PsfmError):
        _C.check(-2)


def test_python_surface_refuses_cpu_tensors_and_bad_settings():
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    from packnet_sfm.geometry.pose import Pose
    with pytest.raises(AssertionError, match='automasking'):
        MultiViewPhotometricLoss(automask_loss=True, photometric_reduce_op='mean')
    L = MultiViewPhotometricLoss(photometric_reduce_op='min')
    img = torch.rand(1, 3, 8, 8)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        L(img, [img], [torch.rand(1, 1, 8, 8)], torch.eye(3)[None], torch.eye(3)[None], [Pose.identity(1)])
    # the fork's fisheye intrinsics dicts take the same CUDA-only path
    fe = {'k': torch.zeros(1, 7), 's': torch.ones(1), 'div': torch.ones(1), 'ux': torch.ones(1), 'uy': torch.ones(1)}
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        L(img, [img], [torch.rand(1, 1, 8, 8)], fe, fe, [Pose.identity(1)])
    from packnet_sfm.ops import pack_fisheye
    packed = pack_fisheye({'k': torch.arange(7.)[None], 's': torch.tensor([7.]), 'div': torch.tensor([8.]),
                           'ux': torch.tensor([9.]), 'uy': torch.tensor([10.])}, 2, torch.float64)
    assert packed.shape == (2, 11) and packed.dtype == torch.float64
    assert packed[1].tolist() == [float(i) for i in range(11)]       # k0..k6, s, div, ux, uy (include/psfm.h)
    with pytest.raises(KeyError, match='fisheye'):
        pack_fisheye({'k': torch.zeros(1, 7)}, 1, torch.float32)
    with pytest.raises(NotImplementedError, match='Unknown photometric_reduce_op'):
        MultiViewPhotometricLoss(photometric_reduce_op='median').reduce_photometric_loss([[img]])
    with pytest.raises(AssertionError):
        Pose(torch.eye(3))


def test_host_helpers_match_reference_units():
    from packnet_sfm.geometry.camera import Camera
    from packnet_sfm.geometry.camera_utils import construct_K, scale_intrinsics
    from packnet_sfm.geometry.pose import Pose
    from packnet_sfm.geometry.pose_utils import euler2mat, invert_pose, pose_vec2mat
    u = load_golden('units')
    t = lambda k: torch.from_numpy(u[k])
    K, vec = t('K'), t('vec')
    eq = lambda a, k: np.testing.assert_allclose(a.numpy(), u[k], rtol=1e-12, atol=1e-14)
    eq(euler2mat(vec[:, 3:]), 'euler2mat')
    T = Pose.from_vec(vec, 'euler')
    eq(T.mat, 'pose_mat')
    eq(T.inverse().mat, 'pose_inv')
    eq(invert_pose(T.mat), 'invert_pose')
    eq((T @ T.inverse()).mat, 'pose_compose')
    assert pose_vec2mat(vec, 'euler').shape == (len(vec), 3, 4)
    cam = Camera(K)
    eq(cam.Kinv, 'Kinv')
    eq(cam.scaled(0.5).K, 'K_scaled_half')
    eq(cam.scaled(0.25, 0.5).K, 'K_scaled_xy')
    eq(scale_intrinsics(K.clone(), 0.5, 0.25), 'scale_intrinsics')
    assert cam.scaled(1.0) is cam and len(cam) == len(K)
    np.testing.assert_allclose(construct_K(2., 3., 4., 5.).numpy(), [[2, 0, 4], [0, 3, 5], [0, 0, 1]])
    np.testing.assert_allclose(Pose.identity(3).mat.numpy(), np.tile(np.eye(4), (3, 1, 1)))
    np.testing.assert_allclose(O.pose_vec_to_mat(vec).numpy(), T.mat.numpy(), rtol=1e-14)


def test_progressive_scaling_and_metrics():
    from packnet_sfm.losses.loss_base import LossBase, ProgressiveScaling
    assert ProgressiveScaling(0.0, 4)(0.7) == 4
    assert ProgressiveScaling(0.2, 4)(0.5) == 4          # the reference's is_list test never fires (loss_base.py:45)
    lb = LossBase()
    x = torch.ones(1, requires_grad=True) * 2
    lb.add_metric('a', x)
    assert not lb.metrics['a'].requires_grad and lb.logs == {}


def test_identity_resize_returns_same_object():
    from packnet_sfm.utils.image import gradient_x, gradient_y, interpolate_image, match_scales
    img = torch.rand(2, 3, 8, 10)
    assert interpolate_image(img, (8, 10)) is img
    assert match_scales(img, [torch.rand(2, 1, 8, 10)] * 2, 2)[1] is img
    assert gradient_x(img).shape == (2, 3, 8, 9) and gradient_y(img).shape == (2, 3, 7, 10)


def test_workspace_size_follows_the_image_sets_without_gpu():
    """psfm_loss_workspace_bytes_io (round 2): scales that share their image pointers share the image-sized scratch; the
    configuration-only query stays the upper bound.  Pure host arithmetic: no CUDA call."""
    from packnet_sfm import _C

    def make(B, H, W, Cn, n, shared):
        cfg = _C.LossCfg()
        cfg.B, cfg.C, cfg.n, cfg.H, cfg.W = B, Cn, n, H, W
        for s in range(n):
            cfg.Hs[s], cfg.Ws[s] = H, W
        cfg.automask, cfg.ssim_w, cfg.smooth_w, cfg.C1, cfg.C2 = 1, 0.85, 0.001, 1e-4, 9e-4
        io = _C.LossIO()
        for s in range(n):
            base = 0x10000 * (1 if shared else s + 1)
            io.image[s], io.inv_depth[s] = base, 0x900000 + s
            for j in range(Cn):
                io.ctx[s][j] = base + 0x100 * (j + 1)
        return cfg, io

    cfg, io = make(8, 192, 640, 2, 4, shared=True)
    upper = _C.lib.psfm_loss_workspace_bytes(ctypes.byref(cfg))
    one_set = _C.lib.psfm_loss_workspace_bytes_io(ctypes.byref(cfg), ctypes.byref(io))
    cfg2, io2 = make(8, 192, 640, 2, 4, shared=False)
    four_sets = _C.lib.psfm_loss_workspace_bytes_io(ctypes.byref(cfg2), ctypes.byref(io2))
    assert 0 < one_set < 0.3 * upper and one_set < 100 * 2 ** 20        # 86 MiB at config 2 (round 1: 337 MiB)
    assert four_sets == upper
    assert _C.lib.psfm_loss_workspace_bytes_io(ctypes.byref(cfg), None) == 0


def test_class_default_is_the_upstream_inverse_depth_semantics():
    """ADVICE (round 1): the fork's forward pre-maps sigmoids to inverse depths (:362-369); this build takes true inverse
    depths unless `sigmoid_premap=True` is passed -- on purpose (north star), and the flag reaches the kernels."""
    from packnet_sfm.losses.multiview_photometric_loss import MultiViewPhotometricLoss
    assert MultiViewPhotometricLoss().sigmoid_premap is False
    assert MultiViewPhotometricLoss()._hyper()['premap'] == 0
    L = MultiViewPhotometricLoss(sigmoid_premap=True, min_depth=0.5, max_depth=80.0)
    assert L._hyper()['premap'] == 1 and L._hyper()['min_depth'] == 0.5 and L._hyper()['max_depth'] == 80.0
