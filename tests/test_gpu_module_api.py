"""The reference's own walker_control / projection tests re-stated against the drop-in classes on the GPU
(reference tests/walker_control/test_pose.py, test_p3d_pose.py, test_p3d_pose_projection.py, tests/utils/test_world.py)
plus the module-level step against the oracle."""
import numpy as np
import pytest
import torch

import oracle as O
from _util import assert_close, assert_close_fp32_aware, make_case, make_targets, oracle_step

pytestmark = pytest.mark.gpu
DEV = torch.device('cuda', 0)


@pytest.fixture(scope='module')
def api():
    from pedestrians_video_2_carla_b200.carla_utils.transforms import Rotation
    from pedestrians_video_2_carla_b200.data.carla import reference
    from pedestrians_video_2_carla_b200.data.carla.utils import load, yaml_to_pose_dict
    from pedestrians_video_2_carla_b200.loss import LossModes
    from pedestrians_video_2_carla_b200.modules.flow_output_types import MovementsModelOutputType
    from pedestrians_video_2_carla_b200.modules.layers.projection import ProjectionModule
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.normalizer import Normalizer
    from pedestrians_video_2_carla_b200.utils.world import calculate_world_from_changes
    from pedestrians_video_2_carla_b200.walker_control.p3d_pose import P3dPose
    from pedestrians_video_2_carla_b200.walker_control.p3d_pose_projection import P3dPoseProjection
    return dict(locals())


def test_golden_yaml_pair_through_p3d_pose(api):
    """reference tests/walker_control/test_pose.py:31-52 for P3dPose: relative YAML -> .absolute == absolute YAML
    (loc atol 1e-5 m, angles atol 1e-2 deg, root skipped)."""
    p = api['P3dPose'](device=DEV)
    p.relative = api['yaml_to_pose_dict'](api['load']('sk_female_relative')['transforms'])
    want = api['yaml_to_pose_dict'](api['load']('sk_female_absolute')['transforms'], is_abs=True)
    got = p.absolute
    for i, (name, t) in enumerate(got.items()):
        if i == 0:
            continue
        w = want[name]
        assert np.allclose([t.location.x, t.location.y, t.location.z], [w.location.x, w.location.y, w.location.z],
                           atol=1e-5), name
        for a, b in ((t.rotation.pitch, w.rotation.pitch), (t.rotation.yaw, w.rotation.yaw),
                     (t.rotation.roll, w.rotation.roll)):
            assert abs((a - b + 180) % 360 - 180) < 1e-2, name


def test_relative_roundtrip_and_move(api):
    """test_pose.py:16-28 (dict -> tensors -> dict) and test_p3d_pose.py:21-55 (move == change . relative)."""
    p = api['P3dPose'](device=DEV)
    src = api['yaml_to_pose_dict'](api['load']('adult_male')['transforms'])
    p.relative = src
    back = p.relative
    for name in src:
        a, b = src[name], back[name]
        assert np.allclose([a.location.x, a.location.y, a.location.z], [b.location.x, b.location.y, b.location.z], atol=1e-6)
        for x, y in ((a.rotation.pitch, b.rotation.pitch), (a.rotation.yaw, b.rotation.yaw), (a.rotation.roll, b.rotation.roll)):
            assert abs((x - y + 180) % 360 - 180) < 1e-3
    loc0, rot0 = p.tensors
    Rotation = api['Rotation']
    moves = {'crl_arm__L': Rotation(yaw=-30), 'crl_foreArm__L': Rotation(pitch=-30, roll=10), 'crl_shoulder__L': Rotation(yaw=5)}
    p.move(moves)
    _, rot1 = p.tensors
    names = list(p.empty.keys())
    e = torch.zeros(26, 3, dtype=torch.float64)
    for k, r in moves.items():
        e[names.index(k)] = torch.tensor([np.deg2rad(-r.roll), np.deg2rad(-r.pitch), np.deg2rad(-r.yaw)])
    want = torch.bmm(O.euler_angles_to_matrix(e), rot0.double().cpu())
    assert_close('move', rot1, want)


def test_forward_batch_invariants(api):
    """test_p3d_pose.py:58-101: identity changes keep the relative rotations and every batch item equal; random
    (non-orthonormal) changes make them differ."""
    p = api['reference'].get_poses(DEV)[0]
    loc, rot = p.tensors
    n = 5
    loc, rot = loc.unsqueeze(0).repeat(n, 1, 1), rot.unsqueeze(0).repeat(n, 1, 1, 1)
    ident = torch.eye(3, device=DEV).reshape(1, 1, 3, 3).repeat(n, 26, 1, 1)
    al, ar, rr = p(ident, loc, rot)
    assert al.shape == (n, 26, 3) and ar.shape == (n, 26, 3, 3) and rr.shape == (n, 26, 3, 3)
    assert torch.equal(rr, rot)
    assert all(torch.equal(al[0], al[i]) for i in range(1, n))
    changes = torch.rand((n, 26, 3, 3), device=DEV) * 0.8
    al2, _, rr2 = p(changes, loc, rot)
    assert not torch.allclose(rr2, rot)
    assert not any(torch.allclose(al2[0], al2[i]) for i in range(1, n))


def test_reference_tensors_match_reference_code(api, golden):
    """data/carla/reference.py outputs of the reference's own code (tests/golden/reference_run.npz)."""
    R = api['reference']
    loc, rot = R.get_relative_tensors(DEV)
    assert_close('ref_rel_loc', loc, torch.as_tensor(golden['ref_rel_loc']))
    assert_close('ref_rel_rot', rot, torch.as_tensor(golden['ref_rel_rot']))
    al, ar = R.get_absolute_tensors(DEV)
    assert_close('ref_abs_loc', al, torch.as_tensor(golden['ref_abs_loc']))
    assert_close('ref_abs_rot', ar, torch.as_tensor(golden['ref_abs_rot']))
    assert_close('ref_projections', R.get_projections(DEV), torch.as_tensor(golden['ref_projections']))
    d = R.get_projections(DEV, as_dict=True)
    assert list(d.keys()) == list(R.CARLA_REFERENCE_SKELETON_TYPES)


def test_projection_matches_reference_code_and_invariants(api, golden):
    pp = api['P3dPoseProjection'](device=DEV, pedestrian=object())
    out = pp(torch.as_tensor(golden['proj_x']).to(DEV), torch.as_tensor(golden['proj_loc']).to(DEV),
             torch.as_tensor(golden['proj_rot']).to(DEV))
    assert_close('projection', out, torch.as_tensor(golden['proj_out']))
    # test_p3d_pose_projection.py:75-128: same input in every batch item -> same output
    al, _ = api['reference'].get_absolute_tensors(DEV)
    x = al[:1].repeat(4, 1, 1)
    o = pp(x, torch.zeros(4, 3, device=DEV), torch.eye(3, device=DEV).repeat(4, 1, 1))
    assert all(torch.equal(o[0], o[i]) for i in range(1, 4))
    hips = o[0, 1].cpu().numpy()
    assert abs(hips[0] - 400.0) < 1e-3 and 300 < hips[1] < 600   # hips centred horizontally, below the image centre


def test_world_from_changes(api):
    """tests/utils/test_world.py: shapes, early-out broadcast, initial values respected."""
    f = api['calculate_world_from_changes']
    wl, wr = f((3, 5), DEV)
    assert wl.shape == (3, 5, 3) and wr.shape == (3, 5, 3, 3)
    assert torch.equal(wl, torch.zeros_like(wl)) and torch.equal(wr, torch.eye(3, device=DEV).expand(3, 5, 3, 3))
    g = torch.Generator().manual_seed(0)
    dl = torch.randn((3, 5, 3), generator=g)
    dr = O.euler_angles_to_matrix(torch.rand((3, 5, 3), generator=g) - 0.5)
    il = torch.randn((3, 3), generator=g)
    wl, wr = f((3, 5), DEV, dl.to(DEV), dr.to(DEV), il.to(DEV), None)
    owl, owr = O.calculate_world_from_changes((3, 5), dl, dr, il)
    assert_close('world_loc', wl, owl)
    assert_close('world_rot', wr, owr)
    wl, _ = f((3, 5), DEV, None, dr.to(DEV))
    assert torch.equal(wl, torch.zeros_like(wl))


@pytest.mark.parametrize('mode', ['pose_changes', 'relative_rot'])
def test_projection_module_and_losses_against_oracle(api, mode):
    """ProjectionModule.forward -> Normalizer -> LossModes (the reference's unfused call sequence,
    modules/flow/pose_lifting.py:121-195 + base.py:440-469) with autograd back to the 6-D model output."""
    c = make_case(6, 8, 21, mode=mode)
    gt2, gt3 = make_targets(c, mode)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), mode, gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    c32 = dict(c, loc=c['loc'].float(), rot=c['rot'].float(), wl=c['wl'].float(), wr=c['wr'].float())
    x32 = c['d6'].float().requires_grad_(True)
    o32 = oracle_step(c32, O.rotation_6d_to_matrix(x32), mode, gt=(gt2.float(), gt3.float()))
    o32['loc_2d_3d'].backward()

    from pedestrians_video_2_carla_b200 import ops
    MT = api['MovementsModelOutputType']
    module = api['ProjectionModule'](movements_output_type=MT[mode])
    types = api['reference'].CARLA_REFERENCE_SKELETON_TYPES
    meta = {'age': [types[i][0] for i in c['sid'].tolist()], 'gender': [types[i][1] for i in c['sid'].tolist()]}
    frames = torch.zeros((c['n'], c['l'], 26, 2), device=DEV)
    module.on_batch_start((frames, None, meta), 0)
    x = c['d6'].float().to(DEV).requires_grad_(True)
    proj, out = module(ops.rotation_6d_to_matrix(x), c['wl'].float().to(DEV), c['wr'].float().to(DEV))
    for k in ('relative_pose_loc', 'relative_pose_rot', 'absolute_pose_loc', 'absolute_pose_rot', 'world_loc', 'world_rot'):
        assert_close(k, out[k], want[k])
    assert_close('projection_2d', proj, want['projection_2d'])
    sliced = dict(projection_2d=proj, projection_2d_transformed=api['Normalizer']()(proj),
                  absolute_pose_loc=out['absolute_pose_loc'],
                  targets={'projection_2d_transformed': gt2.float().to(DEV), 'absolute_pose_loc': gt3.float().to(DEV)})
    assert_close('projection_2d_transformed', sliced['projection_2d_transformed'], want['projection_2d_transformed'])
    from pedestrians_video_2_carla_b200.data.carla.skeleton import CARLA_SKELETON
    losses = {}
    # loss_modes=[loc_2d_3d]: its requirements first, then the mode itself (modules/flow/base.py:71-93)
    for m in (api['LossModes'][k] for k in ('loc_2d', 'loc_3d', 'loc_2d_3d')):
        fn, criterion, *deps = m.value
        if isinstance(fn, type):
            fn = fn(criterion=criterion, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON, mask_missing_joints=True)
        losses[m.name] = fn(criterion=criterion, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON,
                            mask_missing_joints=True, requirements={k: losses[k] for k in (deps[0] if deps else ())},
                            loss_weights=None, **sliced)
    for k in ('loc_2d', 'loc_3d', 'loc_2d_3d'):
        assert_close(k, losses[k], want[k])
    losses['loc_2d_3d'].backward()
    assert_close_fp32_aware('grad d6 (unfused module path)', x.grad, d6.grad, x32.grad)

    # the fused entry of the same module gives the same numbers in one call
    y = c['d6'].float().to(DEV).requires_grad_(True)
    proj2, out2 = module.fused_step(y, sliced['targets'], c['wl'].float().to(DEV), c['wr'].float().to(DEV))
    out2['loss']['loc_2d_3d'].backward()
    assert_close('fused projection', proj2, want['projection_2d'])
    assert_close('fused loss', out2['loss']['loc_2d_3d'], want['loc_2d_3d'])
    assert_close_fp32_aware('fused grad', y.grad, d6.grad, x32.grad)
