"""Generate golden input/output vectors by running the REFERENCE'S OWN Python code.

Run in the build container only (`python tests/golden/make_golden.py`); the GPU box has no
/root/reference.  The reference's hot-path modules are imported unmodified from
/root/reference/src.  Their third-party imports that are not installed here are replaced
by minimal stand-ins registered in sys.modules BEFORE the import:

  * pytorch3d 0.6.0 (transforms + PerspectiveCameras/look_at_view_transform): restated from
    the published formulas (same restatement the oracle documents) -- so the camera closed
    form itself stays "parity unpinned"; everything the reference wrote itself (FK recursion,
    frame loops, world scan, axis swaps, normaliser, mask, losses) is REAL reference code.
  * pedestrians_scenarios.karma: the Skeleton enum base, CARLA_SKELETON (DFS order of
    structure.yaml), deepcopy helpers, PoseDict.
  * cameratransform, pytorch_lightning.utilities.warnings, modules.movements.movements
    (only its MovementsModelOutputType re-export is needed): import-only stand-ins.

Outputs: tests/golden/reference_run.npz (+ .json manifest).
"""
import enum
import json
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = '/root/reference/src'
SEED = 22742  # reference default seed, modeling.py:120


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    parent, _, child = name.rpartition('.')
    if parent and parent in sys.modules:
        setattr(sys.modules[parent], child, m)
    return m


# ------------------------------------------------------------------ pytorch3d stand-in
def _axis(axis, a):
    c, s, o, z = torch.cos(a), torch.sin(a), torch.ones_like(a), torch.zeros_like(a)
    flat = {'X': (o, z, z, z, c, -s, z, s, c), 'Y': (c, z, s, z, o, z, -s, z, c), 'Z': (c, -s, z, s, c, z, z, z, o)}[axis]
    return torch.stack(flat, -1).reshape(a.shape + (3, 3))


def euler_angles_to_matrix(e, convention):
    m = [_axis(c, a) for c, a in zip(convention, torch.unbind(e, -1))]
    return torch.matmul(torch.matmul(m[0], m[1]), m[2])


def matrix_to_euler_angles(m, convention):
    assert convention == 'XYZ'
    return torch.stack((torch.atan2(-m[..., 1, 2], m[..., 2, 2]), torch.asin(m[..., 0, 2]),
                        torch.atan2(-m[..., 0, 1], m[..., 0, 0])), -1)


def rotation_6d_to_matrix(d6):
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = F.normalize(a1, dim=-1)
    b2 = F.normalize(a2 - (b1 * a2).sum(-1, keepdim=True) * b1, dim=-1)
    return torch.stack((b1, b2, torch.cross(b1, b2, dim=-1)), dim=-2)


def matrix_to_rotation_6d(m):
    return m[..., :2, :].clone().reshape(m.shape[:-2] + (6,))


def look_at_view_transform(eye, at, up):
    eye, at, up = (torch.tensor(v, dtype=torch.float32) for v in (eye, at, up))
    z = F.normalize(at - eye, eps=1e-5)
    x = F.normalize(torch.cross(up, z, dim=1), eps=1e-5)
    y = F.normalize(torch.cross(z, x, dim=1), eps=1e-5)
    close = torch.isclose(x, torch.tensor(0.0), atol=5e-3).all(dim=1, keepdim=True)
    if close.any():
        x = torch.where(close, F.normalize(torch.cross(y, z, dim=1), eps=1e-5), x)
    R = torch.cat((x[:, None, :], y[:, None, :], z[:, None, :]), dim=1).transpose(1, 2)
    T = -torch.bmm(R.transpose(1, 2), eye[:, :, None])[:, :, 0]
    return R, T


class PerspectiveCameras(object):
    """Closed form of pytorch3d 0.6.0 PerspectiveCameras(in_ndc=False).transform_points_screen (eps=None)."""

    def __init__(self, device, in_ndc, focal_length, principal_point, image_size, R, T):
        assert in_ndc is False
        self.device = device
        self.f = float(np.asarray(focal_length).reshape(-1)[0])
        self.pp = principal_point[0]
        self.R, self.T = R.to(device), T.to(device)

    def transform_points_screen(self, points):
        c = torch.matmul(points, self.R) + self.T[:, None, :]
        z = c[..., 2]
        return torch.stack((self.pp[0] - self.f * c[..., 0] / z, self.pp[1] - self.f * c[..., 1] / z, 1.0 / z), -1)


def install_stubs():
    _mod('pytorch3d')
    _mod('pytorch3d.transforms', euler_angles_to_matrix=euler_angles_to_matrix,
         matrix_to_euler_angles=matrix_to_euler_angles, rotation_6d_to_matrix=rotation_6d_to_matrix,
         matrix_to_rotation_6d=matrix_to_rotation_6d)
    _mod('pytorch3d.transforms.rotation_conversions', euler_angles_to_matrix=euler_angles_to_matrix,
         matrix_to_euler_angles=matrix_to_euler_angles, rotation_6d_to_matrix=rotation_6d_to_matrix,
         matrix_to_rotation_6d=matrix_to_rotation_6d)
    _mod('pytorch3d.transforms.transform3d', Rotate=object, Translate=object)
    _mod('pytorch3d.renderer')
    _mod('pytorch3d.renderer.cameras', PerspectiveCameras=PerspectiveCameras,
         look_at_view_transform=look_at_view_transform)

    import yaml
    structure = yaml.safe_load(open(os.path.join(
        REF_SRC, 'pedestrians_video_2_carla/data/carla/files/structure.yaml')))['structure']
    names = []

    def add(node):
        (n, ch), = node.items()
        names.append(n)
        for c in (ch or []):
            add(c)
    add(structure[0])

    class Skeleton(enum.Enum):
        @classmethod
        def get_hips_point(cls):
            return cls.crl_hips__C

        @classmethod
        def get_neck_point(cls):
            return cls.crl_neck__C

    CARLA_SKELETON = Skeleton('CARLA_SKELETON', [(n, i) for i, n in enumerate(names)])

    def deepcopy_location(l):
        return type(l)(x=l.x, y=l.y, z=l.z)

    def deepcopy_rotation(r):
        return type(r)(pitch=r.pitch, yaw=r.yaw, roll=r.roll)

    def deepcopy_transform(t):
        return type(t)(location=deepcopy_location(t.location), rotation=deepcopy_rotation(t.rotation))

    _mod('pedestrians_scenarios')
    _mod('pedestrians_scenarios.karma')
    _mod('pedestrians_scenarios.karma.pose')
    _mod('pedestrians_scenarios.karma.pose.skeleton', Skeleton=Skeleton, CARLA_SKELETON=CARLA_SKELETON)
    _mod('pedestrians_scenarios.karma.pose.types', PoseDict=dict)
    _mod('pedestrians_scenarios.karma.utils')
    _mod('pedestrians_scenarios.karma.utils.rotations', mul_carla_rotations=None)
    _mod('pedestrians_scenarios.karma.utils.deepcopy', deepcopy_location=deepcopy_location,
         deepcopy_rotation=deepcopy_rotation, deepcopy_transform=deepcopy_transform)
    _mod('cameratransform')
    _mod('pytorch_lightning')
    _mod('pytorch_lightning.utilities')
    _mod('pytorch_lightning.utilities.warnings', rank_zero_warn=lambda *a, **k: None)

    sys.path.insert(0, REF_SRC)
    import pedestrians_video_2_carla.modules.flow.output_types as ot
    _mod('pedestrians_video_2_carla.modules.movements')
    _mod('pedestrians_video_2_carla.modules.movements.movements',
         MovementsModelOutputType=ot.MovementsModelOutputType)
    return names


def main():
    names = install_stubs()
    import warnings
    warnings.simplefilter('ignore')
    from pedestrians_video_2_carla.data.carla import reference as ref
    from pedestrians_video_2_carla.data.carla.skeleton import CARLA_SKELETON
    from pedestrians_video_2_carla.data.carla.utils import load, yaml_to_pose_dict
    from pedestrians_video_2_carla.loss import base_pose_loss, loc_2d, loc_2d_3d, loc_3d
    from pedestrians_video_2_carla.modules.flow.output_types import MovementsModelOutputType
    from pedestrians_video_2_carla.modules.layers.projection import ProjectionModule
    from pedestrians_video_2_carla.transforms.pose.normalization.hips_neck_extractor import HipsNeckExtractor
    from pedestrians_video_2_carla.transforms.pose.normalization.normalizer import Normalizer
    from pedestrians_video_2_carla.utils.world import calculate_world_from_changes
    from pedestrians_video_2_carla.walker_control.p3d_pose import P3dPose
    from pedestrians_video_2_carla.walker_control.p3d_pose_projection import P3dPoseProjection

    dev = torch.device('cpu')
    out = {}
    g = torch.Generator().manual_seed(SEED)

    # -- 1. reference skeleton tensors / absolute / projections (data/carla/reference.py:52-117)
    rel_loc, rel_rot = ref.get_relative_tensors(dev)
    abs_loc, abs_rot = ref.get_absolute_tensors(dev)
    out['ref_rel_loc'], out['ref_rel_rot'] = rel_loc, rel_rot
    out['ref_abs_loc'], out['ref_abs_rot'] = abs_loc, abs_rot
    out['ref_projections'] = ref.get_projections(dev)

    # -- 2. P3dPose.forward, non-orthonormal changes (tests/walker_control/test_p3d_pose.py:58-101)
    pose = P3dPose(device=dev)
    pose.relative = yaml_to_pose_dict(load('sk_female_relative.yaml')['transforms'])[0]
    loc0, rot0 = pose.tensors
    n = 4
    changes = torch.rand((n, 26, 3, 3), generator=g) * 0.8
    a_loc, a_rot, r_rot = pose.forward(changes, loc0[None].repeat(n, 1, 1), rot0[None].repeat(n, 1, 1, 1))
    out['fwd_changes'], out['fwd_abs_loc'], out['fwd_abs_rot'], out['fwd_rel_rot'] = changes, a_loc, a_rot, r_rot

    # -- 3. world scan (utils/world.py:16-63)
    wl = torch.randn((3, 7, 3), generator=g) * 0.1
    wr = euler_angles_to_matrix(torch.randn((3, 7, 3), generator=g) * 0.2, 'XYZ')
    w_loc, w_rot = calculate_world_from_changes((3, 7), dev, wl, wr)
    out['world_loc_change'], out['world_rot_change'], out['world_loc'], out['world_rot'] = wl, wr, w_loc, w_rot

    # -- 4. ProjectionModule + Normalizer + losses, fwd + bwd through the reference's autograd graph
    ages = ['adult', 'child', 'adult', 'child', 'adult', 'adult']
    genders = ['female', 'male', 'male', 'female', 'female', 'male']
    N, L = len(ages), 5
    for mode in ('pose_changes', 'relative_rot'):
        d6 = torch.randn((N, L, 26, 6), generator=g)
        if mode == 'pose_changes':   # small changes around identity, like a trained model would emit
            d6 = d6 * 0.15 + torch.tensor([1., 0, 0, 0, 1, 0])
        d6.requires_grad_(True)
        mats = rotation_6d_to_matrix(d6)
        mats.retain_grad()
        wlc = torch.randn((N, L, 3), generator=g) * 0.05
        wrc = euler_angles_to_matrix(torch.randn((N, L, 3), generator=g) * 0.05, 'XYZ')
        # NB: the reference cannot backprop into world changes: utils/world.py:56-61 writes world_rot
        # in place while bmm saved a view of it (autograd raises); only ZeroTrajectory exists upstream.
        pm = ProjectionModule(movements_output_type=MovementsModelOutputType[mode])
        frames = torch.zeros((N, L, 26, 2))
        pm.on_batch_start((frames, None, {'age': ages, 'gender': genders}), 0)
        proj, d = pm(mats, wlc, wrc)
        norm = Normalizer(HipsNeckExtractor(CARLA_SKELETON))
        proj_t = norm(proj)
        # targets: an independent draw through the same reference code, some joints zeroed (never hips)
        with torch.no_grad():
            t6 = torch.randn((N, L, 26, 6), generator=g)
            if mode == 'pose_changes':
                t6 = t6 * 0.15 + torch.tensor([1., 0, 0, 0, 1, 0])
            tproj, td = pm(rotation_6d_to_matrix(t6), wlc.detach(), wrc.detach())
            tproj_t = Normalizer(HipsNeckExtractor(CARLA_SKELETON))(tproj)
            drop = torch.rand((N, L, 26), generator=g) < 0.1
            drop[..., CARLA_SKELETON.crl_hips__C.value] = False
            tproj_t[drop] = 0.0
        targets = {'projection_2d_transformed': tproj_t, 'absolute_pose_loc': td['absolute_pose_loc']}
        mse = torch.nn.MSELoss(reduction='mean')
        l2d = loc_2d.Loc2DPoseLoss(criterion=mse, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON,
                                   mask_missing_joints=True)(targets=targets, projection_2d=proj,
                                                             projection_2d_transformed=proj_t)
        l3d = loc_3d.calculate_loss_loc_3d(criterion=mse, input_nodes=CARLA_SKELETON,
                                           output_nodes=CARLA_SKELETON,
                                           absolute_pose_loc=d['absolute_pose_loc'], targets=targets)
        loss = loc_2d_3d.calculate_loss_loc_2d_3d(requirements={'loc_2d': l2d, 'loc_3d': l3d})
        loss.backward()
        k = mode + '/'
        out[k + 'd6'], out[k + 'mats'] = d6, mats
        out[k + 'world_loc_change'], out[k + 'world_rot_change'] = wlc, wrc
        out[k + 'projection_2d'], out[k + 'projection_2d_transformed'] = proj, proj_t
        for kk, v in d.items():
            out[k + kk] = v
        out[k + 'gt_projection_2d_transformed'] = tproj_t
        out[k + 'gt_absolute_pose_loc'] = td['absolute_pose_loc']
        out[k + 'loc_2d'], out[k + 'loc_3d'], out[k + 'loc_2d_3d'] = l2d, l3d, loss
        out[k + 'grad_d6'], out[k + 'grad_mats'] = d6.grad, mats.grad
    skel_id = [ref.CARLA_REFERENCE_SKELETON_TYPES.index((a, s)) for a, s in zip(ages, genders)]

    # -- 5. P3dPoseProjection alone with the training camera (p3d_pose_projection.py:115-152)
    from pedestrians_video_2_carla.walker_control.controlled_pedestrian import ControlledPedestrian
    ped = ControlledPedestrian(world=None, age='adult', gender='female', device=dev, reference_pose=P3dPose)
    pp = P3dPoseProjection(device=dev, pedestrian=ped)
    x = abs_loc[:1].repeat(3, 1, 1) + torch.randn((3, 26, 3), generator=g) * 0.05
    ploc = torch.randn((3, 3), generator=g) * 0.3
    prot = euler_angles_to_matrix(torch.randn((3, 3), generator=g) * 0.3, 'XYZ')
    out['proj_x'], out['proj_loc'], out['proj_rot'], out['proj_out'] = x, ploc, prot, pp(x, ploc, prot)
    out['camera_R'], out['camera_T'] = pp.camera.R[0], pp.camera.T[0]

    arrays = {k: v.detach().numpy() for k, v in out.items()}
    arrays['skeleton_id'] = np.asarray(skel_id, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, 'reference_run.npz'), **arrays)
    manifest = {'seed': SEED, 'bones': names, 'ages': ages, 'genders': genders,
                'keys': {k: list(v.shape) for k, v in arrays.items()},
                'reference_commit': '410bba4', 'torch': torch.__version__,
                'note': 'generated by tests/golden/make_golden.py from the reference code with third-party stubs'}
    json.dump(manifest, open(os.path.join(HERE, 'reference_run.json'), 'w'), indent=1)
    print('wrote', len(arrays), 'arrays;', os.path.getsize(os.path.join(HERE, 'reference_run.npz')), 'bytes')
    for k in ('pose_changes/loc_2d', 'pose_changes/loc_3d', 'relative_rot/loc_2d', 'relative_rot/loc_3d'):
        print(k, float(arrays[k]))


if __name__ == '__main__':
    main()
