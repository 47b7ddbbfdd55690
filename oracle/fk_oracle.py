"""Plain-torch CPU restatement of the reference hot path (TEST INFRASTRUCTURE, see __init__).

Every function cites the reference file:line it follows (paths relative to
/root/reference/src/pedestrians_video_2_carla unless noted).  Op ORDER follows the
reference (per-bone recursion, per-frame Python loops) so that it doubles as the timed
CPU baseline; dtype is a parameter so tests can run it in float64 and use torch
autograd as the gradient reference.
"""
import math
import os
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F
import yaml

__all__ = [
    'FILES_DIR', 'load_yaml', 'bone_names', 'parents_of', 'SKELETON_FILES', 'SKELETON_TYPES',
    'euler_angles_to_matrix', 'matrix_to_euler_angles', 'rotation_6d_to_matrix', 'matrix_to_rotation_6d',
    'yaml_to_pose', 'pose_to_tensors', 'reference_tensors', 'move_to_relative', 'relative_to_absolute',
    'p3d_pose_forward', 'calculate_world_from_changes', 'look_at_view_transform', 'Camera',
    'training_camera', 'reference_camera', 'project', 'abs_from_pose_changes', 'abs_from_relative_rot',
    'projection_module_forward', 'hips_neck_normalize', 'missing_joints_mask', 'loss_loc_2d',
    'loss_loc_3d', 'loss_loc_2d_3d', 'full_step', 'HIPS', 'NECK',
    'normalize', 'get_shift_scale', 'denormalize', 'reference_denormalize', 'cumulative_changes', 'loss_rot_3d',
    'quaternion_to_matrix', 'normalize_with_extractor', 'extractor_shift_scale', 'metric_mpjpe', 'metric_mrpe', 'metric_pck', 'get_bboxes', 'loss_cum_pose_changes', 'loss_pose_changes', 'loss_weighted_loc_2d_loc_rot_3d', 'abs_from_abs_loc',
]

FILES_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..',
                         'pedestrians_video_2_carla_b200', 'data', 'carla', 'files')

# data/carla/utils.py:27-33
SKELETON_FILES = {
    'adult_female': 'sk_female_relative.yaml',
    'adult_male': 'sk_male_relative.yaml',
    'child_female': 'sk_girl_relative.yaml',
    'child_male': 'sk_kid_relative.yaml',
}
# data/carla/reference.py:12-17
SKELETON_TYPES = (('adult', 'female'), ('adult', 'male'), ('child', 'female'), ('child', 'male'))

# CARLA_SKELETON values = DFS index of structure.yaml; hips = crl_hips__C, neck = crl_neck__C
HIPS, NECK = 1, 8


def load_yaml(name, files_dir=None):
    """data/base/utils.py:14-25 + data/carla/utils.py:16-37."""
    with open(os.path.join(files_dir or FILES_DIR, name), 'r') as f:
        return yaml.safe_load(f)


def bone_names(structure):
    """DFS order of the bone tree: walker_control/pose.py:41-47."""
    out = []

    def add(node):
        (name, children), = node.items()
        out.append(name)
        for c in (children or []):
            add(c)
    add(structure[0])
    return out


def parents_of(structure):
    """Parent index per bone (root = -1), same DFS as p3d_pose.py:116-149."""
    names = bone_names(structure)
    par = [-1] * len(names)

    def walk(node, p):
        (name, children), = node.items()
        i = names.index(name)
        par[i] = p
        for c in (children or []):
            walk(c, i)
    walk(structure[0], -1)
    return par


# ----------------------------------------------------------------------------- pytorch3d 0.6.0 restated
def _axis_angle_rotation(axis, angle):
    c, s = torch.cos(angle), torch.sin(angle)
    one, zero = torch.ones_like(angle), torch.zeros_like(angle)
    if axis == 'X':
        flat = (one, zero, zero, zero, c, -s, zero, s, c)
    elif axis == 'Y':
        flat = (c, zero, s, zero, one, zero, -s, zero, c)
    else:
        flat = (c, -s, zero, s, c, zero, zero, zero, one)
    return torch.stack(flat, -1).reshape(angle.shape + (3, 3))


def euler_angles_to_matrix(euler, convention='XYZ'):
    """pytorch3d.transforms.euler_angles_to_matrix: R = R_c0(e0) @ R_c1(e1) @ R_c2(e2)
    (call sites walker_control/p3d_pose.py:54,259, p3d_pose_projection.py:112)."""
    m = [_axis_angle_rotation(c, e) for c, e in zip(convention, torch.unbind(euler, -1))]
    return torch.matmul(torch.matmul(m[0], m[1]), m[2])


def matrix_to_euler_angles(m, convention='XYZ'):
    """pytorch3d.transforms.matrix_to_euler_angles for 'XYZ' (p3d_pose.py:78):
    (atan2(-M12, M22), asin(M02), atan2(-M01, M00))."""
    assert convention == 'XYZ'
    return torch.stack((torch.atan2(-m[..., 1, 2], m[..., 2, 2]),
                        torch.asin(m[..., 0, 2]),
                        torch.atan2(-m[..., 0, 1], m[..., 0, 0])), -1)


def rotation_6d_to_matrix(d6):
    """pytorch3d.transforms.rotation_6d_to_matrix (modules/movements/movements.py:113-116)."""
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = F.normalize(a1, dim=-1)
    b2 = a2 - (b1 * a2).sum(-1, keepdim=True) * b1
    b2 = F.normalize(b2, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def matrix_to_rotation_6d(m):
    """pytorch3d.transforms.matrix_to_rotation_6d (modules/movements/seq2seq.py:330)."""
    return m[..., :2, :].clone().reshape(m.shape[:-2] + (6,))


def look_at_view_transform(eye, at, up):
    """pytorch3d.renderer.cameras.look_at_view_transform / look_at_rotation
    (walker_control/p3d_pose_projection.py:43-47).  Returns R (3,3) with the camera axes
    as COLUMNS and T = -R^T eye; points map as X_cam = X_world @ R + T."""
    eye = torch.as_tensor(eye, dtype=torch.float64)
    at = torch.as_tensor(at, dtype=torch.float64)
    up = torch.as_tensor(up, dtype=torch.float64)
    z = F.normalize(at - eye, dim=0, eps=1e-5)
    x = F.normalize(torch.linalg.cross(up, z), dim=0, eps=1e-5)
    y = F.normalize(torch.linalg.cross(z, x), dim=0, eps=1e-5)
    if bool(torch.isclose(x, torch.zeros_like(x), atol=5e-3).all()):
        x = F.normalize(torch.linalg.cross(y, z), dim=0, eps=1e-5)
    R = torch.stack((x, y, z), dim=1)
    T = -(R.T @ eye)
    return R, T


# ----------------------------------------------------------------------------- reference skeleton loading
def yaml_to_pose(transforms, is_abs=False):
    """data/carla/utils.py:40-77: cm -> m; relative: hips location := 0; absolute: subtract hips
    location from every bone.  Returns OrderedDict bone -> ((x, y, z), (pitch, yaw, roll))."""
    pose = OrderedDict()
    for name, t in transforms.items():
        pose[name] = ([t['location'][k] / 100.0 for k in 'xyz'],
                      [t['rotation'][k] for k in ('pitch', 'yaw', 'roll')])
    hips = list(pose['crl_hips__C'][0])
    if is_abs:
        for name in pose:
            pose[name][0][:] = [a - b for a, b in zip(pose[name][0], hips)]
    else:
        pose['crl_hips__C'][0][:] = [0.0, 0.0, 0.0]
    return pose


def pose_to_tensors(pose, names, dtype=torch.float32):
    """walker_control/p3d_pose.py:34-54: loc -> (x, y, -z); rot -> euler_XYZ(deg2rad(-roll, -pitch, -yaw))."""
    locs, rots = [], []
    for n in names:
        (x, y, z), (pitch, yaw, roll) = pose[n]
        locs.append((x, y, -z))
        rots.append((np.deg2rad(-roll), np.deg2rad(-pitch), np.deg2rad(-yaw)))
    return (torch.tensor(locs, dtype=dtype),
            euler_angles_to_matrix(torch.tensor(rots, dtype=dtype), 'XYZ'))


def reference_tensors(dtype=torch.float32, files_dir=None):
    """data/carla/reference.py:20-63: (4,B,3), (4,B,3,3) for SKELETON_TYPES order, plus names, parents."""
    structure = load_yaml('structure.yaml', files_dir)['structure']
    names = bone_names(structure)
    locs, rots = [], []
    for age, gender in SKELETON_TYPES:
        pose = yaml_to_pose(load_yaml(SKELETON_FILES['%s_%s' % (age, gender)], files_dir)['transforms'])
        l, r = pose_to_tensors(pose, names, dtype)
        locs.append(l)
        rots.append(r)
    return torch.stack(locs), torch.stack(rots), names, parents_of(structure)


# ----------------------------------------------------------------------------- P3dPose
def move_to_relative(changes, prev_rel_rot):
    """p3d_pose.py:98-114."""
    bs = changes.shape[0]
    return torch.bmm(changes.reshape(-1, 3, 3), prev_rel_rot.reshape(-1, 3, 3)).reshape(bs, -1, 3, 3)


def _transform_descendants(abs_loc, abs_rot, rel_loc, rel_rot, children_of, idx, prev_transform):
    """p3d_pose.py:116-149, one bone per call, 4x4 homogeneous row-vector transforms."""
    pad_rel_loc = F.pad(rel_loc[:, idx:idx + 1], pad=(0, 1, 0, 0), mode='constant', value=1)
    loc = torch.bmm(pad_rel_loc, prev_transform)
    abs_loc[:, idx] = loc[:, 0, :3]
    abs_rot[:, idx] = torch.bmm(rel_rot[:, idx], prev_transform[:, :3, :3])
    new_transform = torch.eye(4, dtype=rel_loc.dtype).reshape(1, 4, 4).repeat(abs_loc.shape[0], 1, 1)
    new_transform[:, :3, :3] = abs_rot[:, idx]
    new_transform[:, 3, :3] = abs_loc[:, idx]
    for c in children_of[idx]:
        _transform_descendants(abs_loc, abs_rot, rel_loc, rel_rot, children_of, c, new_transform)


def relative_to_absolute(loc, rot, parents):
    """p3d_pose.py:151-184."""
    children_of = [[c for c, p in enumerate(parents) if p == b] for b in range(len(parents))]
    abs_loc = torch.zeros_like(loc)
    abs_rot = torch.zeros_like(rot)
    initial = torch.eye(4, dtype=loc.dtype).reshape(1, 4, 4).repeat(loc.shape[0], 1, 1)
    _transform_descendants(abs_loc, abs_rot, loc, rot, children_of, parents.index(-1), initial)
    return abs_loc, abs_rot


def p3d_pose_forward(changes, prev_rel_loc, prev_rel_rot, parents):
    """p3d_pose.py:186-213 -> (abs_loc, abs_rot, rel_rot)."""
    rot = move_to_relative(changes, prev_rel_rot)
    a_loc, a_rot = relative_to_absolute(prev_rel_loc, rot, parents)
    return a_loc, a_rot, rot


# ----------------------------------------------------------------------------- utils/world.py
def calculate_world_from_changes(shape, loc_change=None, rot_change=None, initial_loc=None, initial_rot=None,
                                 dtype=torch.float32):
    """utils/world.py:16-63."""
    n, l = shape[0], shape[1]
    if initial_loc is None:
        initial_loc = torch.zeros((n, 3), dtype=dtype)
    if initial_rot is None:
        initial_rot = torch.eye(3, dtype=dtype).reshape(1, 3, 3).repeat(n, 1, 1)
    if loc_change is None and rot_change is None:
        return initial_loc.unsqueeze(1).repeat(1, l, 1), initial_rot.unsqueeze(1).repeat(1, l, 1, 1)
    if loc_change is None:
        loc_change = torch.zeros((n, l, 3), dtype=dtype)
    if rot_change is None:
        rot_change = torch.eye(3, dtype=dtype).reshape(1, 1, 3, 3).repeat(n, l, 1, 1)
    world_loc = torch.empty((n, l + 1, 3), dtype=dtype)
    world_rot = torch.empty((n, l + 1, 3, 3), dtype=dtype)
    world_loc[:, 0] = initial_loc
    world_rot[:, 0] = initial_rot
    for i in range(l):
        world_rot[:, i + 1] = torch.bmm(world_rot[:, i], rot_change[:, i])
        world_loc[:, i + 1] = world_loc[:, i] + loc_change[:, i]
    return world_loc[:, 1:], world_rot[:, 1:]


# ----------------------------------------------------------------------------- P3dPoseProjection
class Camera(object):
    """Pinhole constants of p3d_pose_projection.py:37-69 (+ pose_projection.py:18-39)."""

    def __init__(self, eye, at, image_size=(800, 600), fov=90.0, lens_x_size=0.08):
        self.R, self.T = look_at_view_transform(eye, at, (0.0, 0.0, -1.0))
        sensor_width_mm = lens_x_size * 1000
        focal_length_mm = sensor_width_mm / (2 * np.tan(np.deg2rad(fov) / 2))
        self.f = float(focal_length_mm * 10)            # p3d_pose_projection.py:63
        self.px, self.py = image_size[0] / 2, image_size[1] / 2
        self.image_size = image_size


def training_camera(distance=3.1, elevation=1.2):
    """carla_utils/setup.py:37-62 + pose_projection.py:67-75 + p3d_pose_projection.py:38-41."""
    return Camera(eye=(distance, 0.0, -elevation), at=(0.0, 0.0, -elevation))


def reference_camera():
    """data/carla/reference.py:95-99 (look_at=(0,0,0), camera_position=(3.1,0,0))."""
    return Camera(eye=(3.1, 0.0, -0.0), at=(0.0, 0.0, 0.0))


def project(x, loc, rot, cam):
    """p3d_pose_projection.py:115-152 with PerspectiveCameras(in_ndc=False).transform_points_screen
    of pytorch3d 0.6.0 reduced to closed form: u = px - f X/Z, v = py - f Y/Z, out_z = 1/Z.

    PARITY UNPINNED (pytorch3d is not installable here; the reference pins `pytorch3d==0.6.0`, Dockerfile:16,98-99 /
    environment.yml:225 -- the v0.6.0 release tag; no commit hash is recorded by the reference).  Read from that
    release's renderer/cameras.py + renderer/utils semantics: for a camera defined in screen space (in_ndc=False)
    `transform_points_screen` composes  view (R, T)  ->  the screen-space projection  x_s = f X/Z + px  ->
    screen-to-NDC WITHOUT the axis flip  x_n = (x_s - cx) / s  ->  NDC-to-screen WITH the flip (+X left, +Y up in
    NDC)  u = cx - s x_n = 2 cx - (f X/Z + px),  cx = W/2, and hands back 1/Z in the third channel.  The reference's
    principal point IS the image centre (px = W/2, p3d_pose_projection.py:55-63), so  u = px - f X/Z.  The one open
    point: a pytorch3d release that centres NDC at (W-1)/2 would shift raw `projection_2d` by exactly ONE PIXEL in u and
    v; the hips/neck-normalised projection, every loss and every gradient on this path are invariant to that shift."""
    n = x.shape[0]
    dt = x.dtype
    p3d_2_world = torch.tensor(((0., -1., 0.), (1., 0., 0.), (0., 0., 1.)), dtype=dt).expand(n, -1, -1)
    world_x = torch.bmm(x, p3d_2_world)
    world_transform = torch.eye(4, dtype=dt).reshape(1, 4, 4).repeat(n, 1, 1)
    world_transform[:, :3, :3] = rot
    world_transform[:, 3, :3] = loc
    world_pos = torch.bmm(F.pad(world_x, pad=(0, 1, 0, 0), mode='constant', value=1), world_transform)[..., :3]
    c = torch.matmul(world_pos, cam.R.to(dt)) + cam.T.to(dt)
    z = c[..., 2]
    return torch.stack((cam.px - cam.f * c[..., 0] / z, cam.py - cam.f * c[..., 1] / z, 1.0 / z), -1)


# ----------------------------------------------------------------------------- ProjectionModule
def abs_from_pose_changes(changes, ref_loc, ref_rot, parents):
    """modules/layers/projection.py:170-195.  changes (N,L,B,3,3); ref_* per clip (N,B,..)."""
    n, l, b = changes.shape[:3]
    abs_loc = torch.empty((n, l, b, 3), dtype=changes.dtype)
    abs_rot = torch.empty((n, l, b, 3, 3), dtype=changes.dtype)
    rel_loc = ref_loc.unsqueeze(1).repeat(1, l, 1, 1)
    rel_rot = torch.empty((n, l, b, 3, 3), dtype=changes.dtype)
    prev = ref_rot
    for i in range(l):
        abs_loc[:, i], abs_rot[:, i], rel_rot[:, i] = p3d_pose_forward(changes[:, i], ref_loc, prev, parents)
        prev = rel_rot[:, i]
    return rel_loc, rel_rot, abs_loc, abs_rot


def abs_from_relative_rot(rel_rot, ref_loc, parents):
    """modules/layers/projection.py:144-168."""
    n, l, b = rel_rot.shape[:3]
    abs_loc = torch.empty((n, l, b, 3), dtype=rel_rot.dtype)
    abs_rot = torch.empty((n, l, b, 3, 3), dtype=rel_rot.dtype)
    for i in range(l):
        abs_loc[:, i], abs_rot[:, i] = relative_to_absolute(ref_loc, rel_rot[:, i], parents)
    return ref_loc.unsqueeze(1).repeat(1, l, 1, 1), rel_rot, abs_loc, abs_rot


def projection_module_forward(pose_inputs, ref_loc, ref_rot, parents, cam, world_loc_change=None,
                              world_rot_change=None, mode='pose_changes'):
    """modules/layers/projection.py:73-123 -> (projections, dict)."""
    if mode == 'pose_changes':
        rel_loc, rel_rot, abs_loc, abs_rot = abs_from_pose_changes(pose_inputs, ref_loc, ref_rot, parents)
    else:
        rel_loc, rel_rot, abs_loc, abs_rot = abs_from_relative_rot(pose_inputs, ref_loc, parents)
    world_loc, world_rot = calculate_world_from_changes(abs_loc.shape, world_loc_change, world_rot_change,
                                                        dtype=abs_loc.dtype)
    proj = torch.empty_like(abs_loc)
    for i in range(abs_loc.shape[1]):
        proj[:, i] = project(abs_loc[:, i], world_loc[:, i], world_rot[:, i], cam)
    return proj, {'relative_pose_loc': rel_loc, 'relative_pose_rot': rel_rot, 'absolute_pose_loc': abs_loc,
                  'absolute_pose_rot': abs_rot, 'world_loc': world_loc, 'world_rot': world_rot}


# ----------------------------------------------------------------------------- normaliser + losses
def hips_neck_normalize(sample, hips=HIPS, neck=NECK, near_zero=1e-5):
    """transforms/pose/normalization/normalizer.py:20-41 + extractor.py:23-36 + hips_neck_extractor.py:6-13."""
    xy = sample[..., 0:2]
    shift = xy[..., (hips,), :].mean(dim=-2)
    scale_pt = xy[..., (neck,), :].mean(dim=-2)
    scale = torch.linalg.norm(scale_pt - shift, dim=shift.ndim - 1, ord=2)
    out = torch.empty_like(sample)
    out[..., 0:2] = (xy - shift.unsqueeze(-2)) / scale[(...,) + (None,) * 2]
    if sample.shape[-1] > 2:
        out[..., 2] = sample[..., 2]
    out = torch.nan_to_num(out, nan=0, posinf=0, neginf=0)
    if out.shape[-1] > 2:
        out[..., 0:2] = out[..., 0:2].where(out[..., 2:] >= near_zero, torch.tensor(0.0, dtype=out.dtype))
    return out


def missing_joints_mask(gt, hips=HIPS):
    """utils/tensors.py:29-40."""
    mask = torch.all(gt != 0, dim=-1)
    if hips is not None:
        mask[..., hips] = 1
    return mask


def loss_loc_2d(pred, gt, mask_missing_joints=True, hips=HIPS):
    """loss/loc_2d.py:69-89 + loss/base_pose_loss.py:36-66 with nn.MSELoss(reduction='mean')."""
    p, g = pred[..., 0:2], gt[..., 0:2]
    if mask_missing_joints:
        m = missing_joints_mask(g, hips)
        p, g = p[m], g[m]
    return F.mse_loss(p, g, reduction='mean')


def loss_loc_3d(abs_loc, gt_abs_loc):
    """loss/loc_3d.py:12-40."""
    return F.mse_loss(abs_loc, gt_abs_loc, reduction='mean')


def loss_loc_2d_3d(l2d, l3d):
    """loss/loc_2d_3d.py:6-17."""
    return l2d + l3d


def full_step(pose_changes, ref_loc, ref_rot, parents, cam, gt_proj_transformed, gt_abs_loc,
              world_loc_change=None, world_rot_change=None, mode='pose_changes', mask_missing_joints=True,
              normalize=True):
    """One pose_lifting step's hot path (modules/flow/pose_lifting.py:121-195 + flow/base.py:440-469):
    ProjectionModule.forward -> Normalizer -> loc_2d, loc_3d, loc_2d_3d.  Returns dict."""
    proj, out = projection_module_forward(pose_changes, ref_loc, ref_rot, parents, cam, world_loc_change,
                                          world_rot_change, mode)
    proj_t = hips_neck_normalize(proj) if normalize else proj
    l2d = loss_loc_2d(proj_t, gt_proj_transformed, mask_missing_joints)
    l3d = loss_loc_3d(out['absolute_pose_loc'], gt_abs_loc)
    out.update(projection_2d=proj, projection_2d_transformed=proj_t, loc_2d=l2d, loc_3d=l3d,
               loc_2d_3d=loss_loc_2d_3d(l2d, l3d))
    return out


# ----------------------------------------------------------------------------- SURVEY 8f rows 1-2
def get_shift_scale(sample, hips=HIPS, neck=NECK):
    """transforms/pose/normalization/extractor.py:23-36 with hips_neck_extractor.py:6-13."""
    shift = sample[..., (hips,), :].mean(dim=-2)
    scale_pt = sample[..., (neck,), :].mean(dim=-2)
    scale = torch.linalg.norm(scale_pt - shift, dim=shift.ndim - 1, ord=2)
    scale = scale[((slice(None),) * scale.ndim + (None,) * 2)[:(sample.ndim - 2)]]
    return shift, scale


def normalize(sample, dim=2, hips=HIPS, neck=NECK, near_zero=1e-5):
    """transforms/pose/normalization/normalizer.py:20-41 for dim = 2 or 3."""
    shift, scale = get_shift_scale(sample[..., 0:dim], hips, neck)
    out = torch.empty_like(sample)
    out[..., 0:dim] = (sample[..., 0:dim] - torch.unsqueeze(shift, -2)) / scale[(...,) + (None,) * 2]
    if dim == 2 and sample.shape[-1] > 2:
        out[..., 2] = sample[..., 2]
    out = torch.nan_to_num(out, nan=0, posinf=0, neginf=0)
    if dim == 2 and out.shape[-1] > 2:
        out[..., 0:2] = out[..., 0:2].where(out[..., 2:] >= near_zero, torch.tensor(0.0, dtype=out.dtype))
    return out


def denormalize(sample, scale, shift, dim=2):
    """transforms/pose/normalization/denormalizer.py:13-29."""
    out = torch.empty_like(sample)
    d = scale[(slice(None),) * scale.ndim + (None,) * (sample.ndim - scale.ndim)]
    h = shift[(slice(None),) * scale.ndim + (None,) * (sample.ndim - shift.ndim) + (slice(None),)]
    out[..., 0:dim] = (sample[..., 0:dim] * d) + h
    if dim == 2 and sample.shape[-1] > 2:
        out[..., 2] = sample[..., 2]
    return out


def reference_denormalize(frames, reference, skeleton_id, dim, autonormalize=False, hips=HIPS, neck=NECK):
    """reference_skeletons_denormalizer.py:41-91 (from_projection: dim 2, reference = reference projections[..., :2];
    from_abs: dim 3, reference = reference absolute locations); `reference` holds one entry per skeleton type and
    `skeleton_id` picks the clip's (the reference stacks them per clip, :60-63 / :84-88)."""
    if autonormalize:
        frames = normalize(frames, dim, hips, neck)
    per_clip = torch.stack([reference[int(i)] for i in skeleton_id], dim=0)
    shift, scale = get_shift_scale(per_clip, hips, neck)       # denormalizer.py:31-35
    return denormalize(frames, scale, shift, dim)


def abs_from_abs_loc(pose_inputs, reference_abs_loc, skeleton_id):
    """modules/layers/projection.py:125-136."""
    return reference_denormalize(pose_inputs, reference_abs_loc, skeleton_id, 3, autonormalize=True)


def cumulative_changes(changes):
    """loss/cum_pose_changes.py:24-52: prev = prev . changes[t] per (clip, bone), frame by frame."""
    n, l, b = changes.shape[:3]
    prev = torch.eye(3, dtype=changes.dtype).reshape(1, 3, 3).repeat(n * b, 1, 1)
    mats = changes.transpose(0, 1).reshape(l, -1, 3, 3)
    out = []
    for i in range(l):
        prev = torch.bmm(prev, mats[i])
        out.append(prev)
    return torch.stack(out, dim=0).reshape(l, n, b, 3, 3).transpose(0, 1)


def loss_cum_pose_changes(pose_inputs, gt_pose_changes):
    """loss/cum_pose_changes.py:9-56 with nn.MSELoss('mean')."""
    return F.mse_loss(cumulative_changes(pose_inputs), cumulative_changes(gt_pose_changes), reduction='mean')


def loss_pose_changes(pose_inputs, gt_pose_changes):
    """loss/pose_changes.py:7-29 with nn.MSELoss('sum') (loss/__init__.py:33)."""
    return F.mse_loss(pose_inputs, gt_pose_changes, reduction='sum')


def loss_rot_3d(abs_rot, gt_abs_rot):
    """loss/rot_3d.py:9-37."""
    return F.mse_loss(abs_rot, gt_abs_rot, reduction='mean')


def loss_weighted_loc_2d_loc_rot_3d(l2d, l3d, r3d, weights=None):
    """loss/weighted_loc_2d_loc_rot_3d.py:6-25 (unit weights: loc_2d_loc_rot_3d.py:6-21)."""
    w = weights or {}
    return w.get('loc_2d', 1.0) * l2d + w.get('loc_3d', 1.0) * l3d + w.get('rot_3d', 1.0) * r3d


def quaternion_to_matrix(quaternions):
    """pytorch3d 0.6.0 transforms/rotation_conversions.py quaternion_to_matrix (real part first; restated from the
    published formula -- the reference has no quaternion call site, the north star names the entry).  Pinned in
    tests/ against scipy.spatial.transform.Rotation for unit quaternions."""
    r, i, j, k = torch.unbind(quaternions, -1)
    two_s = 2.0 / (quaternions * quaternions).sum(-1)
    o = torch.stack((
        1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
        two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
        two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(quaternions.shape[:-1] + (3, 3))


# ----------------------------------------------------------------------------- metrics
def metric_mpjpe(pred_abs_loc, gt_abs_loc):
    """metrics/mpjpe.py:29-49 for one update: millimetres."""
    per_clip = torch.mean(torch.linalg.norm(pred_abs_loc - gt_abs_loc, dim=-1, ord=2), dim=(-2, -1))
    return 1000.0 * torch.sum(per_clip) / per_clip.numel()


def metric_mrpe(pred_abs_loc, gt_abs_loc, pred_world_loc_changes, gt_world_loc_changes, pred_world_loc=None,
                hips=HIPS, neck=NECK):
    """metrics/mrpe.py:38-80 for one update: millimetres."""
    if pred_world_loc_changes is not None:
        pw, _ = calculate_world_from_changes(pred_world_loc_changes.shape, pred_world_loc_changes,
                                             dtype=pred_abs_loc.dtype)
    else:
        pw = pred_world_loc
    gw, _ = calculate_world_from_changes(pw.shape, gt_world_loc_changes, dtype=pred_abs_loc.dtype)
    ph, _ = get_shift_scale(pred_abs_loc, hips, neck)
    th, _ = get_shift_scale(gt_abs_loc, hips, neck)
    per_clip = torch.mean(torch.linalg.norm((pw + ph) - (gw + th), dim=-1, ord=2), dim=-1)
    return 1000.0 * torch.sum(per_clip) / per_clip.numel()


def get_bboxes(sample, near_zero=1e-5):
    """utils/tensors.py:12-26."""
    missing = torch.all(sample[..., 0:2] < near_zero, dim=-1)
    lo = sample.clone()
    lo[missing] = float('inf')
    hi = sample.clone()
    hi[missing] = float('-inf')
    return torch.stack((lo.min(dim=-2)[0], hi.max(dim=-2)[0]), dim=-2)


def metric_pck(pred, gt, gt_projection_2d=None, norm='bbox', threshold=0.05, mask_missing_joints=True, hips=HIPS,
               neck=NECK, near_zero=1e-5):
    """metrics/pck.py:54-97 for one update -> (correct, total)."""
    if mask_missing_joints and gt_projection_2d is not None:
        mask = missing_joints_mask(gt_projection_2d, hips)
    else:
        mask = torch.ones(gt.shape[:-1], dtype=torch.bool)
    if norm == 'hn':
        normalize = get_shift_scale(gt, hips, neck)[1]
    else:
        boxes = get_bboxes(gt, near_zero)
        normalize = torch.linalg.norm(boxes[..., 1, :] - boxes[..., 0, :], dim=-1, ord=2)
    normalize = normalize.clone()
    mask[normalize < near_zero] = 0
    normalize[normalize < near_zero] = 1
    normalize = normalize[(slice(None),) * normalize.ndim + (None,) * (pred.ndim - normalize.ndim)]
    dist = torch.linalg.norm((pred - gt) / normalize, dim=-1, ord=2)
    return int(torch.sum(dist[mask] < threshold)), int(dist[mask].numel())


def extractor_shift_scale(sample, extractor='hips_neck_bbox', hips=HIPS, neck=NECK, near_zero=1e-5):
    """Extractor.get_shift_scale of the 2-D extractors: BBoxExtractor (bbox_extractor.py:6-18),
    HipsNeckBBoxFallbackExtractor (hips_neck_bbox_fallback_extractor.py:21-46) or HipsNeckExtractor ('hips_neck');
    these are the values Normalizer.shift / .scale hold after a call (normalizer.py:39-49)."""
    xy = sample[..., 0:2]
    if extractor == 'hips_neck':
        shift, scale = get_shift_scale(xy, hips, neck)
        return shift, scale
    boxes = get_bboxes(xy, near_zero)
    bb_shift = boxes.mean(dim=-2)
    bb_scale_pt = torch.stack((boxes.mean(dim=-2)[..., 0], boxes.min(dim=-2)[0][..., 1]), dim=-1)
    bb_scale = torch.linalg.norm(bb_scale_pt - bb_shift, dim=bb_shift.ndim - 1, ord=2)
    if extractor == 'bbox':
        shift, scale = bb_shift, bb_scale
    else:
        hn_shift = xy[..., (hips,), :].mean(dim=-2)
        hn_neck = xy[..., (neck,), :].mean(dim=-2)
        hn_scale = torch.linalg.norm(hn_neck - hn_shift, dim=hn_shift.ndim - 1, ord=2)
        missing_hips = torch.all(hn_shift < near_zero, dim=-1)
        shift = hn_shift.clone()
        if torch.any(missing_hips):      # :27-31: chained indexing assigns into a copy -- no effect on `shift`
            shift[missing_hips][:, 0] = bb_shift[missing_hips][:, 0] + bb_scale[missing_hips] * 0.0
            shift[missing_hips][:, 1] = bb_shift[missing_hips][:, 1] + bb_scale[missing_hips] * -0.1059
        missing_neck = torch.all(hn_neck < near_zero, dim=-1)
        scale = hn_scale.clone()
        if torch.any(missing_hips):
            scale[missing_hips] = bb_scale[missing_hips] * 0.5748
        if torch.any(missing_neck):
            scale[missing_neck] = bb_scale[missing_neck] * 0.5748
    return shift, scale


def normalize_with_extractor(sample, extractor='hips_neck_bbox', hips=HIPS, neck=NECK, near_zero=1e-5):
    """Normalizer (normalizer.py:20-41, dim=2) with BBoxExtractor (bbox_extractor.py:6-18) or
    HipsNeckBBoxFallbackExtractor (hips_neck_bbox_fallback_extractor.py:21-46)."""
    xy = sample[..., 0:2]
    shift, scale = extractor_shift_scale(sample, extractor, hips, neck, near_zero)
    out = torch.empty_like(sample)
    out[..., 0:2] = (xy - torch.unsqueeze(shift, -2)) / scale[(...,) + (None,) * 2]
    if sample.shape[-1] > 2:
        out[..., 2] = sample[..., 2]
    out = torch.nan_to_num(out, nan=0, posinf=0, neginf=0)
    if out.shape[-1] > 2:
        out[..., 0:2] = out[..., 0:2].where(out[..., 2:] >= near_zero, torch.tensor(0.0, dtype=out.dtype))
    return out
