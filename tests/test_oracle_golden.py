"""Pins the CPU oracle (oracle/photometric_oracle.py) to the reference: every fixture under
tests/golden/ was produced by the reference's own classes (oracle/gen_golden.py, run in the
build container where /root/reference exists).  CPU only."""
import numpy as np
import pytest
import torch

import photometric_oracle as O
from golden_cases import CASES, load_golden, make_inputs

TINY = [n for n in CASES if n.startswith('tiny')]
MID = [n for n in CASES if n.startswith('B2_')]
BIG = [n for n in CASES if n.startswith('B1_')]
FISHEYE = [n for n in CASES if n.startswith('fisheye_')]


def run_oracle(name, dtype=torch.float64):
    inp = make_inputs(name, dtype)
    invd = [d.requires_grad_() for d in inp['inv_depths']]
    vecs = [v.requires_grad_() for v in inp['vecs']]
    mats = [O.pose_vec_to_mat(v) for v in vecs]
    for m in mats:
        m.retain_grad()
    out = O.multiview_photometric_loss(inp['image'], inp['context'], invd, inp['K'], inp['K'], mats,
                                       mask=inp['mask'], fork_premap=inp['fork'], return_parts=True, **inp['hp'])
    out['loss'].backward()
    return out, invd, vecs, mats


def check_case(name, rtol=1e-11):
    gold = load_golden(name)
    out, invd, vecs, mats = run_oracle(name)
    n = len(gold['d_invd_norm'])
    np.testing.assert_allclose(out['loss'].detach().numpy(), gold['loss'], rtol=rtol)
    np.testing.assert_allclose(out['photometric_loss'].numpy(), gold['photometric_loss'], rtol=rtol)
    np.testing.assert_allclose(out['smoothness_loss'].numpy(), gold['smoothness_loss'], rtol=rtol, atol=1e-300)
    np.testing.assert_allclose(torch.stack([v.grad for v in vecs]).numpy(), gold['d_vec'], rtol=1e-8, atol=1e-13)
    np.testing.assert_allclose(torch.stack([m.grad for m in mats]).numpy(), gold['d_posemat'], rtol=1e-8, atol=1e-13)
    np.testing.assert_allclose([invd[i].grad.norm().item() for i in range(n)], gold['d_invd_norm'], rtol=1e-9)
    for i in range(n):
        if 'd_invd_s%d' % i in gold:
            np.testing.assert_allclose(invd[i].grad.numpy(), gold['d_invd_s%d' % i], rtol=1e-8, atol=1e-14)
        if 'cands_s%d' % i in gold:
            np.testing.assert_allclose(torch.cat(out['candidates'][i], 1).detach().numpy(), gold['cands_s%d' % i],
                                       rtol=1e-10, atol=1e-13)
            for j in range(len(out['warped'])):
                np.testing.assert_allclose(out['warped'][j][i].detach().numpy(), gold['warp%d_s%d' % (j, i)],
                                           rtol=1e-10, atol=1e-13)
        w0 = out['warped'][0][i].detach()
        np.testing.assert_allclose([w0.mean().item(), w0.abs().sum().item(), (w0 == 0).double().mean().item()],
                                   gold['warp0_s%d_stats' % i], rtol=1e-10)


@pytest.mark.parametrize('name', TINY + MID)
def test_oracle_matches_reference_small(name):
    check_case(name)


@pytest.mark.parametrize('name', BIG[:2])
def test_oracle_matches_reference_kitti(name):
    check_case(name)


@pytest.mark.parametrize('name', FISHEYE)
def test_oracle_matches_reference_fisheye(name):
    """the reference's FisheyeCamera inside its loss (oracle/gen_golden_fisheye.py)."""
    check_case(name)


def test_fisheye_units_against_reference():
    """FisheyeCamera.reconstruct / project (frame 'c') of the reference, incl. points behind the camera."""
    u = load_golden('fisheye_units')
    t = lambda k: torch.from_numpy(u[k])
    intr = {n: t('intr_' + n) for n in ('k', 's', 'div', 'ux', 'uy')}
    eq = lambda a, b: np.testing.assert_allclose(a.numpy(), u[b], rtol=1e-12, atol=1e-14)
    X = O.fisheye_reconstruct(t('depth'), intr)
    eq(X, 'reconstruct_c')
    Xr = O.transform_points(t('pose_mat'), X)
    eq(O.fisheye_project(Xr, intr), 'project_c')
    eq(O.fisheye_project(t('edge_points'), intr), 'project_edge')
    for pm in ('zeros', 'border'):
        eq(O.fisheye_view_synthesis(t('img'), t('depth'), intr, t('pose_mat'), intr, pm), 'view_synthesis_' + pm)
    assert O.fisheye_pack(intr).shape == (2, 11)


def test_survey_appendix_b_values():
    """SURVEY.md Appendix B row 5 (B2 64x160 C2 U zeros), as printed there."""
    g = load_golden('B2_64x160_C2_U_zeros')
    assert abs(float(g['loss'][0]) - 0.331711678682) < 1e-11
    assert abs(float(g['photometric_loss']) - 0.331645006933) < 1e-11
    np.testing.assert_allclose(g['d_vec'][0, 0], [-1.76486478e-03, 4.82144930e-03, 5.76757203e-05,
                                                  -9.86472822e-02, -4.32803122e-02, 5.18670414e-02], rtol=1e-7)


def test_units_against_reference():
    u = load_golden('units')
    t = lambda k: torch.from_numpy(u[k])
    K, vec, depth, img, img2 = t('K'), t('vec'), t('depth'), t('img'), t('img2')
    T = O.pose_vec_to_mat(vec)
    eq = lambda a, b: np.testing.assert_allclose(a.numpy() if hasattr(a, 'numpy') else a, u[b], rtol=1e-12, atol=1e-14)
    eq(O.euler_to_rotation(vec[:, 3:]), 'euler2mat')
    eq(T, 'pose_mat')
    eq(O.invert_rigid(T), 'pose_inv')
    eq(O.invert_rigid(T), 'invert_pose')
    eq(T @ O.invert_rigid(T), 'pose_compose')
    eq(O.lifting_matrix(K), 'Kinv')
    eq(O.scale_K(K, 0.5), 'K_scaled_half')
    eq(O.scale_K(K, 0.25, 0.5), 'K_scaled_xy')
    eq(O.scale_K(K, 0.5, 0.25), 'scale_intrinsics')
    eq(O.reconstruct(depth, K), 'reconstruct_c')
    eq(O.reconstruct(depth, K, O.invert_rigid(T)), 'reconstruct_w')
    X = O.reconstruct(depth, K)
    eq(O.project(X, K, T), 'project_w')
    eq(O.project(X, K), 'project_c')
    eq(O.transform_points(T, X), 'transform_points')
    for pm in ('zeros', 'border'):
        eq(O.view_synthesis(img, depth, K, T, K, pm), 'view_synthesis_' + pm)
    eq(O.ssim_index(img, img2), 'ssim')
    eq(O.ssim_distance(img, img2), 'ssim_loss')
    invd = [t('invd0'), t('invd1')]
    imgs = O.match_scales(img, invd, 2)
    eq(imgs[1], 'match_scale1')
    sx, sy = O.smoothness_maps(invd, imgs, 2)
    eq(sx[0], 'smooth_x0'); eq(sy[0], 'smooth_y0'); eq(sx[1], 'smooth_x1'); eq(sy[1], 'smooth_y1')
    eq(O.inv2depth(invd[0]), 'inv2depth')


def test_grid_sample_probe_values():
    """SURVEY Appendix A.6 probe, produced by torch in the build container."""
    u = load_golden('units')
    np.testing.assert_allclose(u['gs_out_zeros'].ravel(), [1.75, 3.64, 3.0025, 11.0], atol=1e-12)
    np.testing.assert_allclose(u['gs_out_border'].ravel(), [7.0, 5.2, 3.0025, 11.0], atol=1e-12)
    np.testing.assert_allclose(u['gs_dgrid_zeros'].reshape(4, 2), [[-10.5, 1.0], [7.8, 2.8], [1.5, 4.0], [-16.5, -11.0]],
                               atol=1e-12)
    np.testing.assert_allclose(u['gs_dgrid_border'].reshape(4, 2), [[0, 4], [0, 4], [1.5, 4], [0, 0]], atol=1e-12)


def test_supervised_oracle_matches_reference():
    """oracle/supervised_oracle.py against the reference's own SupervisedLoss (fp64)."""
    import supervised_oracle as S
    from golden_cases import load_golden
    from supervised_cases import CASES as SUP, EXTRA, key_of, make_inputs as sup_inputs
    gold = load_golden('supervised')
    for case in SUP:
        preds, gt, masks = sup_inputs(case)
        preds = [p.requires_grad_() for p in preds]
        loss = S.supervised_loss(preds, gt, case[0], case[1], masks, **EXTRA.get(case[0], {}))
        loss.backward()
        k = key_of(case)
        np.testing.assert_allclose(loss.item(), gold[k + '/loss'][0], rtol=1e-12)
        for i, p in enumerate(preds):
            got = torch.zeros_like(p) if p.grad is None else p.grad
            np.testing.assert_allclose(got.numpy(), gold[k + '/d_pred%d' % i], rtol=1e-10, atol=1e-16)


def test_dual_head_oracle_matches_reference():
    """oracle/supervised_oracle.py::dual_head_loss against the reference's own DualHeadDepthLoss (fp64)."""
    import supervised_oracle as S
    from golden_cases import load_golden
    from supervised_cases import DUAL_CASES
    gold = load_golden('dual_head')
    for name, (kw, shape, gt_size) in DUAL_CASES.items():
        ip, fp, gt = S.synth_dual_head(*shape, gt_size=gt_size)
        ip, fp = ip.requires_grad_(), fp.requires_grad_()
        res = S.dual_head_loss(ip, fp, gt, **kw)
        if res[0].requires_grad:
            res[0].backward()
        np.testing.assert_allclose([float(x) for x in res], gold[name + '/losses'], rtol=1e-12, atol=1e-300)
        for t, key in ((ip, 'd_integer'), (fp, 'd_fractional')):
            got = torch.zeros_like(t) if t.grad is None else t.grad
            np.testing.assert_allclose(got.numpy(), gold[name + '/' + key], rtol=1e-10, atol=1e-16)


def test_depth_metrics_oracle_matches_reference():
    """compute_depth_metrics restatement vs the reference's own function (oracle/gen_golden_metrics.py), incl. its float32
    a1..a3 path: bit-exact on every catalogue case."""
    import metrics_oracle as M
    gold = load_golden('depth_metrics')
    assert sorted(gold.keys()) == sorted(M.CASES)
    for name, (B, _, _, crop, scale_output, use_gt_scale, lo, hi, _) in M.CASES.items():
        gt, pred = M.synth_metrics(name)
        got = M.compute_depth_metrics(gt, pred, lo, hi, crop, scale_output, use_gt_scale).numpy()
        np.testing.assert_array_equal(got, gold[name], err_msg=name)
    with pytest.raises(NotImplementedError):
        M.scale_depth(torch.ones(1, 1, 2, 2), torch.ones(1, 1, 4, 4), 'bottom-left')


def test_ssi_gradient_term_oracle_matches_reference():
    """SSISilogLoss with gradient_weight = 0.2 (multi-scale Sobel term, losses/ssi_silog_loss.py:116-174) vs the reference's
    own SupervisedLoss (oracle/gen_golden_supervised.py::gradient_term)."""
    import supervised_oracle as S
    from supervised_cases import GRAD_CASES, GRAD_KW, make_grad_inputs
    gold = load_golden('ssi_gradient')
    for name, (B, H, W, use_mask, gs) in GRAD_CASES.items():
        preds, gt, masks = make_grad_inputs(name)
        preds = [p.requires_grad_() for p in preds]
        loss = S.supervised_loss(preds, gt, 'sparse-ssi-silog', 1, masks=masks, **dict(GRAD_KW, gradient_scales=gs))
        np.testing.assert_allclose(float(loss), gold[name + '/loss'], rtol=1e-12, atol=1e-300, err_msg=name)
        if torch.is_tensor(loss) and loss.requires_grad:
            loss.backward()
            np.testing.assert_allclose(preds[0].grad.numpy(), gold[name + '/d_pred'], rtol=1e-10, atol=1e-14, err_msg=name)
        else:
            assert not gold[name + '/d_pred'].any()


def test_raysurf_oracle_matches_the_reference_fixtures():
    """oracle/raysurf_oracle.py against GenericMultiViewPhotometricLoss / GenericCamera outputs."""
    import numpy as np
    import photometric_oracle as O
    import raysurf_oracle as RO
    for name, (skw, hp) in RO.RAYSURF_CASES.items():
        g = load_golden(name)
        cfg = dict(RO.RAYSURF_LOSS)
        cfg.update(hp)
        image, context, invd, res, vecs, tmpl = RO.synth_raysurf(skw['H'], skw['W'], skw['C'])
        invd, res = invd.clone().requires_grad_(), res.clone().requires_grad_()
        vecs = [v.clone().requires_grad_() for v in vecs]
        out = RO.generic_photometric_loss(image, context, invd, res, tmpl, [O.pose_vec_to_mat(v) for v in vecs],
                                          skw['progress'], **cfg)
        out['loss'].backward()
        rel = lambda a, b: float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))
        assert abs(out['loss'].item() - g['loss'].item()) <= 1e-12 * abs(g['loss'].item())
        assert np.abs(out['warped'][0].detach().numpy() - g['warped0']).max() <= 1e-11
        assert rel(invd.grad.numpy(), g['d_inv_depth']) <= 1e-10
        assert rel(np.stack([v.grad.numpy() for v in vecs]), g['d_vec']) <= 1e-10
        if np.linalg.norm(g['d_residual']) > 0:
            assert rel(res.grad.numpy(), g['d_residual']) <= 1e-10
