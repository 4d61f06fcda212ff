"""fp32 production kernels against the fp64 oracle: warped image max-abs <= 1e-4 (north star).

The loss / gradient tolerances (T2 with the ACTUAL flip set, T3) live in tests/test_loss_shapes_gpu.py, which runs
them at the benchmarked shapes and at the golden-case shapes (64x160 U / P / border, 192x640), bf16 images included."""
import pytest
import torch
import torch.nn.functional as F

import photometric_oracle as O
from golden_cases import make_inputs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('case,pm', [('B1_192x640_C2_U_zeros', 'zeros'), ('B1_192x640_C2_U_zeros', 'border'),
                                     ('cfg3_B1_384x640_C2_bf16', 'zeros'), ('cfg4_B1_1216x1936_C2_border', 'border')])
def test_fp32_warped_image_max_abs(case, pm):
    from packnet_sfm.geometry.camera import Camera
    from packnet_sfm.geometry.camera_utils import view_synthesis
    from packnet_sfm.geometry.pose import Pose
    inp = make_inputs(case, torch.float64)
    inp = {k: ([t.float().double() for t in v] if isinstance(v, list) else (v.float().double() if torch.is_tensor(v) else v))
           for k, v in inp.items()}      # fp32-representable values: the kernel and the oracle see the same inputs
    d, K, img = inp['inv_depths'][0], inp['K'], inp['context'][0]
    want = O.view_synthesis(img, O.inv2depth(d), K, O.pose_vec_to_mat(inp['vecs'][0]), K, pm)
    dev = 'cuda:0'
    f = lambda t: t.float().to(dev)
    got = view_synthesis(f(img), f(O.inv2depth(d)), Camera(f(K), Tcw=Pose.from_vec(f(inp['vecs'][0]), 'euler')),
                         Camera(f(K)), padding_mode=pm)
    assert (got.double().cpu() - want).abs().max().item() <= 1e-4
