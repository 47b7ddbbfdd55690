"""Reference skeletons as tensors (reference data/carla/reference.py:12-117): the four (age, gender) UE4 skeletons,
their absolute poses and projections, cached per device.  `get_relative_tensors` is the 4-row table the kernels index
with a per-clip skeleton id instead of stacking one clone per clip."""
from functools import lru_cache

import torch

from ...walker_control.p3d_pose import P3dPose
from ...walker_control.p3d_pose_projection import P3dPoseProjection
from .utils import load, yaml_to_pose_dict

CARLA_REFERENCE_SKELETON_TYPES = (('adult', 'female'), ('adult', 'male'), ('child', 'female'), ('child', 'male'))


def skeleton_index(age, gender):
    return CARLA_REFERENCE_SKELETON_TYPES.index((age, gender))


@lru_cache(maxsize=10)
def get_poses(device=torch.device('cuda', 0), as_dict=False):
    structure = load('structure')['structure']
    poses = []
    for age, gender in CARLA_REFERENCE_SKELETON_TYPES:
        p = P3dPose(structure=structure, device=device)
        p.relative = yaml_to_pose_dict(load('{}_{}'.format(age, gender))['transforms'])
        poses.append(p)
    return dict(zip(CARLA_REFERENCE_SKELETON_TYPES, poses)) if as_dict else poses


@lru_cache(maxsize=10)
def get_relative_tensors(device=torch.device('cuda', 0), as_dict=False):
    tensors = [p.tensors for p in get_poses(device)]
    if as_dict:
        return dict(zip(CARLA_REFERENCE_SKELETON_TYPES, tensors))
    loc, rot = zip(*tensors)
    return torch.stack(loc), torch.stack(rot)


@lru_cache(maxsize=10)
def get_absolute_tensors(device=torch.device('cuda', 0), as_dict=False):
    poses = get_poses(device)
    loc, rot = get_relative_tensors(device)
    movements = torch.eye(3, device=device).reshape(1, 1, 3, 3).repeat(len(poses), loc.shape[1], 1, 1)
    abs_loc, abs_rot, _ = poses[0](movements, loc, rot)
    if as_dict:
        return {k: (abs_loc[i], abs_rot[i]) for i, k in enumerate(CARLA_REFERENCE_SKELETON_TYPES)}
    return abs_loc, abs_rot


@lru_cache(maxsize=10)
def get_projections(device=torch.device('cuda', 0), as_dict=False):
    reference_abs, _ = get_absolute_tensors(device)
    pp = P3dPoseProjection(device=device, look_at=(0, 0, 0), camera_position=(3.1, 0, 0))
    n = len(reference_abs)
    proj = pp(reference_abs, torch.zeros((n, 3), device=device), torch.eye(3, device=device).repeat(n, 1, 1))
    if as_dict:
        return {k: proj[i] for i, k in enumerate(CARLA_REFERENCE_SKELETON_TYPES)}
    return proj
