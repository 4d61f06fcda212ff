"""Camera helpers (reference: geometry/camera_utils.py)."""
import torch


def construct_K(fx, fy, cx, cy, dtype=torch.float, device=None):
    """[3,3] pinhole intrinsics (reference :9-13)."""
    return torch.tensor([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=dtype, device=device)


def scale_intrinsics(K, x_scale, y_scale):
    """In-place rescale of K for a resized image; returns K (reference :16-22)."""
    K[..., 0, 0] *= x_scale
    K[..., 1, 1] *= y_scale
    K[..., 0, 2] = (K[..., 0, 2] + 0.5) * x_scale - 0.5
    K[..., 1, 2] = (K[..., 1, 2] + 0.5) * y_scale - 0.5
    return K


def view_synthesis(ref_image, depth, ref_cam, cam, mode='bilinear', padding_mode='zeros'):
    """Warp `ref_image` into the view of `cam` using `depth` (reference :27-59).

    One fused kernel (back-project, pose, project, bilinear gather) instead of the reference's
    reconstruct -> project -> grid_sample chain; differentiable w.r.t. depth, ref_cam.Tcw.mat
    and ref_image.  `cam` must sit at the identity pose composed into ref_cam.Tcw the way
    MultiViewPhotometricLoss.warp_ref_image builds it; a non-identity cam.Tcw is folded in."""
    assert depth.size(1) == 1
    if mode != 'bilinear':
        raise NotImplementedError('only bilinear view synthesis is implemented')
    from packnet_sfm.geometry.camera import FisheyeCamera
    if isinstance(cam, FisheyeCamera) or isinstance(ref_cam, FisheyeCamera):
        # the reference's own chain (:50-59), every link a CUDA kernel of this library
        from packnet_sfm.block_ops import grid_sample_op
        world_points = cam.reconstruct(depth, frame='w')
        ref_coords = ref_cam.project(world_points, frame='w')
        return grid_sample_op(ref_image, ref_coords, padding_mode)
    from packnet_sfm.ops import view_synthesis_op
    # world = Twc_cam @ Xc, then Tcw_ref @ world  ==  (Tcw_ref @ Twc_cam) @ Xc
    T = ref_cam.Tcw.mat
    if not cam._identity_pose:
        T = T.bmm(cam.Twc.mat.to(T.dtype))
    return view_synthesis_op(ref_image, depth, cam.K, ref_cam.K, T, padding_mode)


def view_synthesis_generic(ref_image, depth, ref_cam, cam, mode='bilinear', padding_mode='zeros', progress=0.0):
    """Warp `ref_image` with two ray-surface cameras (reference :64-96): reconstruct with `cam`, project with
    `ref_cam` (soft arg-max kernels, geometry/camera_generic.py), bilinear gather."""
    assert depth.size(1) == 1
    if mode != 'bilinear':
        raise NotImplementedError('only bilinear view synthesis is implemented')
    from packnet_sfm.block_ops import grid_sample_op
    world_points = cam.reconstruct(depth, frame='w')
    ref_coords = ref_cam.project(world_points, progress=progress, frame='w')
    return grid_sample_op(ref_image, ref_coords, padding_mode)
