"""Pins the oracle's SURVEY 8f functions (extra losses, Normalizer dim=3, ReferenceSkeletonsDeNormalizer, the
absolute_loc movements output) to outputs of the reference's own code (tests/golden/make_golden_widen.py ->
tests/golden/reference_widen.npz).  CPU only."""
import os

import numpy as np
import pytest
import torch

import oracle as O
from _util import assert_close

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope='module')
def gold():
    z = np.load(os.path.join(HERE, 'golden', 'reference_widen.npz'))
    return {k: torch.from_numpy(z[k]) for k in z.files}


@pytest.fixture(scope='module')
def ref32():
    loc, rot, names, par = O.reference_tensors(torch.float32)
    ident = torch.eye(3).reshape(1, 1, 3, 3).repeat(loc.shape[0], loc.shape[1], 1, 1)
    abs_loc, _, _ = O.p3d_pose_forward(ident, loc, rot, par)
    proj = O.project(abs_loc, torch.zeros((4, 3)), torch.eye(3).repeat(4, 1, 1), O.reference_camera())
    return dict(loc=loc, rot=rot, par=par, abs_loc=abs_loc, proj=proj)


def test_extra_losses_and_gradients(gold, ref32):
    sid = gold['skeleton_id'].long()
    d6 = gold['loss/d6'].clone().requires_grad_(True)
    mats = O.rotation_6d_to_matrix(d6)
    mats.retain_grad()
    assert_close('mats', mats, gold['loss/mats'], 1e-6)
    out = O.full_step(mats, ref32['loc'][sid], ref32['rot'][sid], ref32['par'], O.training_camera(),
                      gold['loss/gt_projection_2d_transformed'], gold['loss/gt_absolute_pose_loc'])
    assert_close('absolute_pose_rot', out['absolute_pose_rot'], gold['loss/absolute_pose_rot'], 1e-5)
    r3d = O.loss_rot_3d(out['absolute_pose_rot'], gold['loss/gt_absolute_pose_rot'])
    w = dict(zip(('loc_2d', 'loc_3d', 'rot_3d'), gold['loss/weights'].tolist()))
    vals = {
        'rot_3d': r3d,
        'weighted': O.loss_weighted_loc_2d_loc_rot_3d(out['loc_2d'], out['loc_3d'], r3d, w),
        'cum_pose_changes': O.loss_cum_pose_changes(mats, gold['loss/gt_pose_changes']),
        'pose_changes': O.loss_pose_changes(mats, gold['loss/gt_pose_changes']),
    }
    assert_close('loc_rot_3d', out['loc_3d'] + r3d, gold['loss/loc_rot_3d'], 1e-5)
    assert_close('loc_2d_loc_rot_3d', O.loss_weighted_loc_2d_loc_rot_3d(out['loc_2d'], out['loc_3d'], r3d),
                 gold['loss/loc_2d_loc_rot_3d'], 1e-5)
    for name, val in vals.items():
        assert_close(name, val, gold['loss/' + name], 1e-5)
        d6.grad = None
        mats.grad = None
        val.backward(retain_graph=True)
        assert_close('grad_mats_' + name, mats.grad, gold['loss/grad_mats_' + name], 2e-5)
        assert_close('grad_d6_' + name, d6.grad, gold['loss/grad_d6_' + name], 2e-5)


def test_normalizer_dim3_and_reference_denormalizer(gold, ref32):
    sid = gold['skeleton_id'].long()
    x3 = gold['den/x3'].clone().requires_grad_(True)
    assert_close('norm3', O.normalize(x3, 3), gold['den/norm3'], 1e-6)
    a3 = O.reference_denormalize(x3, ref32['abs_loc'], sid, 3, autonormalize=True)
    assert_close('from_abs', a3, gold['den/from_abs'], 1e-6)
    (a3 * gold['den/w3']).sum().backward()
    assert_close('grad_x3', x3.grad, gold['den/grad_x3'], 2e-5)
    x3b = gold['den/x3'].clone().requires_grad_(True)
    (O.normalize(x3b, 3) * gold['den/w3']).sum().backward()
    assert_close('grad_x3_norm3', x3b.grad, gold['den/grad_x3_norm3'], 2e-5)
    assert_close('from_abs_plain', O.reference_denormalize(gold['den/norm3'], ref32['abs_loc'], sid, 3),
                 gold['den/from_abs_plain'], 1e-6)
    x2 = gold['den/x2'].clone().requires_grad_(True)
    p2 = O.reference_denormalize(x2, ref32['proj'][..., :2], sid, 2, autonormalize=True)
    assert_close('from_projection', p2, gold['den/from_projection'], 1e-6)
    (p2 * gold['den/w2']).sum().backward()
    assert_close('grad_x2', x2.grad, gold['den/grad_x2'], 2e-5)
    p2c = O.reference_denormalize(gold['den/x2'][..., :2].contiguous(), ref32['proj'][..., :2], sid, 2, True)
    assert_close('from_projection_2ch', p2c, gold['den/from_projection_2ch'], 1e-6)


def test_absolute_loc_movements_output(gold, ref32):
    sid = gold['skeleton_id'].long()
    x = gold['absloc/x'].clone().requires_grad_(True)
    abs_loc = O.abs_from_abs_loc(x, ref32['abs_loc'], sid)
    assert_close('absolute_pose_loc', abs_loc, gold['absloc/absolute_pose_loc'], 1e-6)
    n, l = x.shape[:2]
    world_loc, world_rot = O.calculate_world_from_changes((n, l))
    cam = O.training_camera()
    proj = torch.stack([O.project(abs_loc[:, i], world_loc[:, i], world_rot[:, i], cam) for i in range(l)], dim=1)
    assert_close('projection_2d', proj, gold['absloc/projection_2d'], 1e-6)
    (proj * gold['absloc/w']).sum().backward()
    assert_close('grad_x', x.grad, gold['absloc/grad_x'], 2e-5)


def test_quaternion_oracle_against_scipy():
    """pytorch3d is not installed anywhere reachable: the restated quaternion_to_matrix is pinned to scipy for unit
    quaternions and to its own scale invariance (M(q) == M(c q)) for the rest."""
    from scipy.spatial.transform import Rotation
    g = torch.Generator().manual_seed(11)
    q = torch.randn((200, 4), generator=g, dtype=torch.float64)
    unit = q / q.norm(dim=-1, keepdim=True)
    want = Rotation.from_quat(unit[:, [1, 2, 3, 0]].numpy()).as_matrix()
    assert np.abs(O.quaternion_to_matrix(unit).numpy() - want).max() < 1e-12
    assert np.abs(O.quaternion_to_matrix(q * 3.7).numpy() - want).max() < 1e-12


def test_metrics_against_reference_golden(gold):
    pa, ga = gold['metric/pred_abs_loc'], gold['metric/gt_abs_loc']
    assert_close('mpjpe', O.metric_mpjpe(pa, ga), gold['metric/mpjpe'], 1e-6)
    assert_close('mrpe', O.metric_mrpe(pa, ga, gold['metric/pred_wlc'], gold['metric/gt_wlc']), gold['metric/mrpe'],
                 1e-6)
    wl = torch.cumsum(gold['metric/pred_wlc'], dim=1) * 1.5
    assert_close('mrpe world_loc', O.metric_mrpe(pa, ga, None, gold['metric/gt_wlc'], wl),
                 gold['metric/mrpe_world_loc'], 1e-6)
    p2, g2 = gold['metric/pred_projection_2d'], gold['metric/gt_projection_2d']
    pt, gt = gold['metric/pred_projection_2d_transformed'], gold['metric/gt_projection_2d_transformed']
    for name, args in (('bbox', (p2, g2, g2, 'bbox', 0.05, True)), ('hn', (p2, g2, g2, 'hn', 0.05, True)),
                       ('bbox_nomask', (p2, g2, g2, 'bbox', 0.05, False)),
                       ('hn_transformed', (pt, gt, g2, 'hn', 0.1, True))):
        correct, total = O.metric_pck(*args)
        want = gold['metric/pck_' + name]
        assert (correct, total) == (int(want[0]), int(want[1])), name


def test_bbox_extractors_against_reference_golden(gold):
    x = gold['ext/x']
    assert_close('bbox', O.normalize_with_extractor(x, 'bbox'), gold['ext/bbox'], 1e-6)
    assert_close('hips_neck_bbox', O.normalize_with_extractor(x, 'hips_neck_bbox'), gold['ext/hips_neck_bbox'], 1e-6)
    assert_close('hips_neck_bbox 2ch', O.normalize_with_extractor(x[..., :2].contiguous(), 'hips_neck_bbox'),
                 gold['ext/hips_neck_bbox_2ch'], 1e-6)
