"""Host-side mirror of the reference API: CPU-checkable pieces (loading, bone order, camera constants) here; the
GPU behaviour of the same classes is in test_gpu_module_api.py (modelled on the reference's tests/walker_control)."""
import numpy as np
import pytest
import torch

import oracle as O
from pedestrians_video_2_carla_b200.data.carla.skeleton import CARLA_PARENTS, CARLA_SKELETON, get_neck_point, get_root_point
from pedestrians_video_2_carla_b200.data.carla.utils import load, yaml_to_pose_dict
from pedestrians_video_2_carla_b200.walker_control.p3d_pose_projection import P3dPoseProjection, look_at_view_transform
from pedestrians_video_2_carla_b200.walker_control.pose import Pose, bone_order


def test_bone_order_matches_reference_dfs():
    names, parents = bone_order(load('structure')['structure'])
    _, _, onames, opar = O.reference_tensors()
    assert names == onames and parents == opar == list(CARLA_PARENTS)
    assert [m.name for m in CARLA_SKELETON] == names
    assert get_root_point() == O.HIPS and get_neck_point() == O.NECK
    p = Pose()
    assert list(p.empty.keys()) == names and all(v is None for v in p.empty.values())


@pytest.mark.parametrize('kind', ['adult_female', 'adult_male', 'child_female', 'child_male'])
def test_yaml_to_pose_dict(kind):
    """cm -> m, hips zeroed for relative poses (reference data/carla/utils.py:40-77)."""
    pose = yaml_to_pose_dict(load(kind)['transforms'])
    want = O.yaml_to_pose(O.load_yaml(O.SKELETON_FILES[kind])['transforms'])
    for name, t in pose.items():
        assert np.allclose([t.location.x, t.location.y, t.location.z], want[name][0])
        assert np.allclose([t.rotation.pitch, t.rotation.yaw, t.rotation.roll], want[name][1])
    hips = pose['crl_hips__C'].location
    assert (hips.x, hips.y, hips.z) == (0.0, 0.0, 0.0)


def test_absolute_yaml_is_hips_relative():
    pose = yaml_to_pose_dict(load('sk_female_absolute')['transforms'], is_abs=True)
    hips = pose['crl_hips__C'].location
    assert (hips.x, hips.y, hips.z) == (0.0, 0.0, 0.0)


def test_camera_constants_match_oracle():
    """p3d_pose_projection.py:37-69: f = 400 px, principal point (400, 300), training / reference cameras."""
    pp = P3dPoseProjection(device='cuda:0', pedestrian=object())
    cam = O.training_camera()
    assert np.allclose(pp.camera._R, cam.R.numpy(), atol=1e-12) and np.allclose(pp.camera._T, cam.T.numpy(), atol=1e-12)
    assert pp.camera.focal_length == pytest.approx(400.0) and pp.camera.principal_point == (400.0, 300.0)
    assert pp.image_size == (800, 600)
    assert np.allclose(pp.camera._R, [[0, 0, -1], [1, 0, 0], [0, -1, 0]], atol=1e-12)
    assert np.allclose(pp.camera._T, [0, -1.2, 3.1], atol=1e-12)
    ref = P3dPoseProjection(device='cuda:0', look_at=(0, 0, 0), camera_position=(3.1, 0, 0))
    rc = O.reference_camera()
    assert np.allclose(ref.camera._R, rc.R.numpy(), atol=1e-12) and np.allclose(ref.camera._T, rc.T.numpy(), atol=1e-12)
    pp.update_camera((5.0, 0.5, 1.0))
    R, T = look_at_view_transform((5.0, 0.5, -1.0), (0.0, 0.0, -1.2), (0, 0, -1))
    assert np.allclose(pp.camera._R, R) and np.allclose(pp.camera._T, T)
    with pytest.raises(AssertionError):
        P3dPoseProjection(device='cuda:0')


def test_projection_module_rejects_bad_input_ranks():
    from pedestrians_video_2_carla_b200.modules.flow_output_types import MovementsModelOutputType
    from pedestrians_video_2_carla_b200.modules.layers.projection import ProjectionModule
    m = ProjectionModule()
    with pytest.raises(RuntimeError):          # reference projection.py:90-92
        m.forward(torch.zeros(2, 3, 26, 6))
    m = ProjectionModule(movements_output_type=MovementsModelOutputType.absolute_loc)
    with pytest.raises(RuntimeError):          # reference projection.py:93-95
        m.forward(torch.zeros(2, 26, 3))
    m = ProjectionModule(movements_output_type=MovementsModelOutputType.absolute_loc_rot)
    with pytest.raises(RuntimeError):          # reference projection.py:96-98
        m.forward(torch.zeros(2, 3, 26, 3))


def test_loss_modes_contract():
    from pedestrians_video_2_carla_b200.loss import LossModes
    assert [m.name for m in LossModes] == ['loc_2d', 'loc_3d', 'rot_3d', 'cum_pose_changes', 'pose_changes',
                                           'loc_2d_3d', 'loc_2d_loc_rot_3d', 'weighted_loc_2d_loc_rot_3d',
                                           'loc_rot_3d']
    assert LossModes.loc_rot_3d.value[2] == ('loc_3d', 'rot_3d')
    assert LossModes.weighted_loc_2d_loc_rot_3d.value[2] == ('loc_2d', 'loc_3d', 'rot_3d')
    w = LossModes.weighted_loc_2d_loc_rot_3d.value[0]
    req = {'loc_2d': torch.tensor(1.0), 'loc_3d': torch.tensor(2.0), 'rot_3d': torch.tensor(4.0)}
    assert float(w(requirements=req, loss_weights={'loc_3d': 0.5})) == 6.0
    assert w(requirements={'loc_2d': torch.tensor(1.0)}, loss_weights={}) is None
    assert LossModes.loc_2d_3d.value[2] == ('loc_2d', 'loc_3d')
    fn = LossModes.loc_2d_3d.value[0]
    assert float(fn(requirements={'loc_2d': torch.tensor(1.5), 'loc_3d': torch.tensor(2.0)})) == 3.5
    assert LossModes.loc_3d.value[0](absolute_pose_loc=torch.zeros(1), targets={}) is None   # loc_3d.py:35-39
