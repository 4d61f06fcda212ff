"""Record the call signatures of the reference's public API on the photometric-loss path
  --  TEST INFRASTRUCTURE, build container only.

Imports /root/reference (read-only) and writes tests/golden/api_signatures.json; the CPU test
tests/test_api_surface.py checks that the drop-in package keeps every parameter name, order and
default.  Regenerate with:  python oracle/gen_signatures.py
"""
import inspect
import json
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get('PSFM_REFERENCE', '/root/reference')

_y, _yc = types.ModuleType('yacs'), types.ModuleType('yacs.config')
_yc.CfgNode = type('CfgNode', (dict,), {})
_y.config = _yc
_m, _mc = types.ModuleType('matplotlib'), types.ModuleType('matplotlib.cm')
_mc.get_cmap = lambda *a, **k: None
_m.cm = _mc
sys.modules.update({'yacs': _y, 'yacs.config': _yc, 'matplotlib': _m, 'matplotlib.cm': _mc})
sys.path.insert(0, REF)

import warnings
warnings.filterwarnings('ignore')

API = {
    'packnet_sfm.losses.multiview_photometric_loss': [
        'SSIM', 'MultiViewPhotometricLoss.__init__', 'MultiViewPhotometricLoss.forward',
        'MultiViewPhotometricLoss.warp_ref_image', 'MultiViewPhotometricLoss.SSIM',
        'MultiViewPhotometricLoss.calc_photometric_loss', 'MultiViewPhotometricLoss.reduce_photometric_loss',
        'MultiViewPhotometricLoss.calc_smoothness_loss'],
    'packnet_sfm.losses.supervised_loss': ['SilogLoss.__init__', 'SilogLoss.forward', 'get_loss_func', 'SupervisedLoss.__init__',
                                           'SupervisedLoss.calculate_loss', 'SupervisedLoss.forward'],
    'packnet_sfm.losses.ssi_silog_loss': ['SSISilogLoss.__init__', 'SSISilogLoss.forward', 'SSISilogLoss.set_depth_range'],
    'packnet_sfm.losses.dual_head_depth_loss': ['DualHeadDepthLoss.__init__', 'DualHeadDepthLoss.forward'],
    'packnet_sfm.losses.loss_base': ['LossBase.add_metric', 'ProgressiveScaling.__init__', 'ProgressiveScaling.__call__'],
    'packnet_sfm.geometry.camera': ['Camera.__init__', 'Camera.scaled', 'Camera.reconstruct', 'Camera.project', 'Camera.to',
                                    'FisheyeCamera.__init__', 'FisheyeCamera.reconstruct', 'FisheyeCamera.project',
                                    'FisheyeCamera.to'],
    'packnet_sfm.geometry.camera_utils': ['construct_K', 'scale_intrinsics', 'view_synthesis'],
    'packnet_sfm.geometry.pose': ['Pose.__init__', 'Pose.identity', 'Pose.from_vec', 'Pose.inverse', 'Pose.repeat',
                                  'Pose.transform_pose', 'Pose.transform_points', 'Pose.__matmul__', 'Pose.item'],
    'packnet_sfm.geometry.pose_utils': ['euler2mat', 'pose_vec2mat', 'invert_pose'],
    'packnet_sfm.utils.image': ['gradient_x', 'gradient_y', 'interpolate_image', 'interpolate_scales', 'match_scales',
                                'meshgrid', 'image_grid'],
    'packnet_sfm.utils.depth': ['inv2depth', 'depth2inv', 'inv_depths_normalize', 'calc_smoothness', 'compute_depth_metrics',
                                'scale_depth'],
}


def describe(fn):
    fn = getattr(fn, '__wrapped__', fn)            # lru_cache wrappers
    fn = getattr(fn, '__func__', fn)               # classmethods
    out = []
    for p in inspect.signature(fn).parameters.values():
        d = None if p.default is inspect.Parameter.empty else repr(p.default)
        out.append([p.name, str(p.kind), d])
    return out


def main():
    import importlib
    rec = {}
    for mod, names in API.items():
        m = importlib.import_module(mod)
        for n in names:
            obj = m
            for part in n.split('.'):
                obj = inspect.getattr_static(obj, part) if inspect.isclass(obj) else getattr(obj, part)
            rec[mod + ':' + n] = describe(obj)
    path = os.path.join(ROOT, 'tests', 'golden', 'api_signatures.json')
    with open(path, 'w') as f:
        json.dump(rec, f, indent=1, sort_keys=True)
    print('wrote', path, len(rec), 'signatures')


if __name__ == '__main__':
    main()
