"""ProjectionModule: the hot-path orchestrator, drop-in for reference modules/layers/projection.py
(on_batch_start(batch, batch_idx); forward(pose_inputs, world_loc_change, world_rot_change) -> (projections, dict)).

Where the reference builds one ControlledPedestrian per clip every batch and loops over frames in Python
(projection.py:52-71, 109-114, 144-195), this module keeps the four reference skeletons as a device table, maps
(age, gender) to a row index per clip, and runs the whole clip batch in ONE kernel launch (FK + projection fused).
`fused_step` additionally fuses the normaliser, loc_2d / loc_3d / loc_2d_3d and their backward for rotation_6d input.
"""
import torch
from torch import nn

from .. import flow_output_types as T
from ... import _lib, ops
from ...data.carla import reference as R
from ...data.carla.skeleton import get_neck_point, get_root_point
from ...transforms.pose.normalization.reference_skeletons_denormalizer import ReferenceSkeletonsDeNormalizer
from ...utils.world import calculate_world_from_changes
from ...walker_control.p3d_pose import P3dPose
from ...walker_control.p3d_pose_projection import P3dPoseProjection


class ProjectionModule(nn.Module):
    def __init__(self, movements_output_type=T.MovementsModelOutputType.pose_changes,
                 trajectory_output_type=T.TrajectoryModelOutputType.changes, **kwargs):
        super().__init__()
        self.movements_output_type = movements_output_type
        self.trajectory_output_type = trajectory_output_type
        self._absolute = movements_output_type in (T.MovementsModelOutputType.absolute_loc,
                                                   T.MovementsModelOutputType.absolute_loc_rot)
        if self._absolute:   # projection.py:30-35: self-normalise, then rescale to the clip's reference skeleton
            self._denormalize = ReferenceSkeletonsDeNormalizer()
        self._mode = (_lib.MODE_POSE_CHANGES if movements_output_type == T.MovementsModelOutputType.pose_changes
                      else _lib.MODE_RELATIVE_ROT)
        self._pose = None
        self._projection = None
        self._ref_index = None
        self._ref = None
        self._world_locations = None
        self._world_rotations = None

    @property
    def pose_projection(self):
        return self._projection

    def on_batch_start(self, batch, batch_idx):
        """batch = (frames, targets, meta) with meta['age'][i], meta['gender'][i] (projection.py:52-71).
        A per-clip `meta['reference_pose']` (custom P3dPose objects) is honoured by stacking their tensors."""
        frames, _, meta = batch
        device = frames.device
        n = len(frames)
        if self._pose is None or self._pose._device != device:
            self._pose = R.get_poses(device)[0]              # one P3dPose instance serves every skeleton type
            self._projection = P3dPoseProjection(device=device, pedestrian=self._pose)
        if 'reference_pose' in meta and not isinstance(meta['reference_pose'][0], type):
            loc, rot = zip(*[p.tensors for p in meta['reference_pose']])
            self._ref, self._ref_index = (torch.stack(loc), torch.stack(rot)), None
        elif torch.is_tensor(meta.get('skeleton_index')):
            # (device data generators) the row of the reference table per clip, already on the device: no host loop
            self._ref = R.get_relative_tensors(device)
            self._ref_index = meta['skeleton_index'].to(device=device, dtype=torch.int32)
        else:
            self._ref = R.get_relative_tensors(device)
            idx = [R.skeleton_index(meta['age'][i], meta['gender'][i]) for i in range(n)]
            self._ref_index = torch.tensor(idx, dtype=torch.int32).to(device, non_blocking=True)
        self._world_locations = torch.zeros((n, 3), device=device)
        self._world_rotations = torch.eye(3, device=device).reshape(1, 3, 3).repeat(n, 1, 1)

    def get_reference_tensors(self):
        """(N,B,3), (N,B,3,3) per-clip reference tensors, as the reference returns them (projection.py:197-206)."""
        loc, rot = self._ref
        if self._ref_index is None:
            return loc, rot
        i = self._ref_index.long()
        return loc[i], rot[i]

    def _world(self, shape, device, loc_in, rot_in):
        if self.trajectory_output_type == T.TrajectoryModelOutputType.changes:
            return calculate_world_from_changes(shape, device, loc_in, rot_in, self._world_locations,
                                                self._world_rotations)
        n, l = shape[0], shape[1]
        if loc_in is None:
            loc_in = torch.zeros((n, l, 3), device=device)
        if rot_in is None:
            rot_in = torch.eye(3, device=device).reshape(1, 1, 3, 3).repeat(n, l, 1, 1)
        return loc_in, rot_in

    def _forward_absolute(self, pose_inputs_batch, world_loc_change_batch, world_rot_change_batch):
        """absolute_loc / absolute_loc_rot movements output (projection.py:125-142): the model emits absolute joint
        locations in its own scale; one kernel self-normalises them (hips -> 0, |neck - hips| -> 1) and rescales to
        the clip's reference skeleton, one kernel projects.  No FK: relative tensors are None as in the reference."""
        abs_rot = None
        if self.movements_output_type == T.MovementsModelOutputType.absolute_loc_rot:
            if not isinstance(pose_inputs_batch, tuple):
                raise RuntimeError('Absolute location with rotation should be a Tuple of tensors.')
            pose_inputs_batch, abs_rot = pose_inputs_batch
        if pose_inputs_batch.ndim < 4:
            raise RuntimeError('Absolute location should have shape of (N - batch_size, L - clip_length, B - bones, '
                               '3 - absolute location coordinates)')
        n, l, b = pose_inputs_batch.shape[:3]
        device = pose_inputs_batch.device
        if self._ref_index is None:
            raise RuntimeError('absolute_loc movements need (age, gender) reference skeletons, not custom poses')
        abs_loc = self._denormalize.from_abs(pose_inputs_batch, None, autonormalize=True, rows=self._ref_index)
        world_loc, world_rot = self._world((n, l), device, world_loc_change_batch, world_rot_change_batch)
        proj = ops.project(abs_loc.reshape(n * l, b, 3), world_loc.reshape(n * l, 3), world_rot.reshape(n * l, 3, 3),
                           self._projection.camera.struct).reshape(n, l, b, 3)
        return proj, {
            'relative_pose_loc': None,
            'relative_pose_rot': None,
            'absolute_pose_loc': abs_loc,
            'absolute_pose_rot': abs_rot,
            'world_loc': world_loc,
            'world_rot': world_rot,
        }

    def forward(self, pose_inputs_batch, world_loc_change_batch=None, world_rot_change_batch=None):
        if self._absolute:
            return self._forward_absolute(pose_inputs_batch, world_loc_change_batch, world_rot_change_batch)
        if pose_inputs_batch.ndim < 5:
            raise RuntimeError('Pose changes should have shape of (N - batch_size, L - clip_length, B - bones, 3, 3 '
                               '- rotations as rotation matrices)')
        n, l, b = pose_inputs_batch.shape[:3]
        device = pose_inputs_batch.device
        world_loc, world_rot = self._world((n, l), device, world_loc_change_batch, world_rot_change_batch)
        ref_loc, ref_rot = self._ref
        rel_rot, abs_rot, abs_loc, proj = ops.fk_clip(
            pose_inputs_batch, ref_loc, ref_rot if self._mode == _lib.MODE_POSE_CHANGES else None, self._ref_index,
            world_loc, world_rot, self._pose.tree, self._projection.camera.struct, self._mode, True)
        if rel_rot is None:
            rel_rot = pose_inputs_batch
        rel_loc = self.get_reference_tensors()[0].unsqueeze(1).repeat(1, l, 1, 1)
        return proj, {
            'relative_pose_loc': rel_loc,
            'relative_pose_rot': rel_rot,
            'absolute_pose_loc': abs_loc,
            'absolute_pose_rot': abs_rot,
            'world_loc': world_loc,
            'world_rot': world_rot,
        }

    def fused_step(self, pose_inputs_6d, targets, world_loc_change_batch=None, world_rot_change_batch=None,
                   mask_missing_joints=True, transform='hips_neck', loss_weights=None, loss_mode='loc_2d_3d'):
        """The whole pose_lifting hot path in one call: raw rotation_6d model output (N,L,B,6) -> projections,
        the module's dict, projection_2d_transformed, loc_2d / loc_3d / loc_2d_3d, with the gradient wrt the 6-D
        input attached to `loc_2d_3d` (modules/flow/pose_lifting.py:121-195 + modules/flow/base.py:440-469).
        `targets` holds 'projection_2d_transformed' (or 'projection_2d' when transform is None),
        'absolute_pose_loc' and, for the loss modes with rot_3d, 'absolute_pose_rot'.  out['loss'][loss_mode] is the
        differentiable value of the chosen mode; loc_2d / loc_3d / rot_3d beside it are reported values."""
        n, l, b = pose_inputs_6d.shape[:3]
        device = pose_inputs_6d.device
        world_loc = world_rot = None
        if world_loc_change_batch is not None or world_rot_change_batch is not None:
            world_loc, world_rot = self._world((n, l), device, world_loc_change_batch, world_rot_change_batch)
        # modules/flow/pose_lifting.py:167-170 applies dm.transform_callable: only the hips/neck normaliser is fused.
        # 'hips_neck_bbox' (the CLI default, HipsNeckBBoxFallbackExtractor) differs from it only when hips or neck are
        # UNDETECTED (pixel coordinates below near_zero), which cannot happen for a projected skeleton in front of the
        # camera; 'bbox' or a user callable is a different normalisation and is refused here (unfused path:
        # forward() + Normalizer(BBoxExtractor()) + the loss mirrors).
        if transform not in (None, 'none', 'hips_neck', 'hips_neck_bbox'):
            raise NotImplementedError("fused_step fuses transform None / 'hips_neck' / 'hips_neck_bbox' only, got %r; "
                                      'use forward() + the transform + the loss mirrors for anything else' % (transform,))
        normalize = transform not in (None, 'none')
        gt2 = targets['projection_2d_transformed' if normalize else 'projection_2d']
        w = loss_weights or {}
        ref_loc, ref_rot = self._ref
        # loss_mode picks the terms of the differentiable sum (loss/__init__.py:28-48); weights as in
        # weighted_loc_2d_loc_rot_3d (missing weights are 1)
        terms = {'loc_2d': ('loc_2d',), 'loc_3d': ('loc_3d',), 'rot_3d': ('rot_3d',), 'loc_2d_3d': ('loc_2d', 'loc_3d'),
                 'loc_rot_3d': ('loc_3d', 'rot_3d'), 'loc_2d_loc_rot_3d': ('loc_2d', 'loc_3d', 'rot_3d'),
                 'weighted_loc_2d_loc_rot_3d': ('loc_2d', 'loc_3d', 'rot_3d')}.get(loss_mode)
        if terms is None:
            raise NotImplementedError('fused_step fuses loc_2d, loc_3d, rot_3d and their sums, got loss_mode %r' % (loss_mode,))
        weighted = loss_mode == 'weighted_loc_2d_loc_rot_3d'
        wt = {k: (float(w.get(k, 1.0)) if weighted else 1.0) if k in terms else 0.0 for k in ('loc_2d', 'loc_3d', 'rot_3d')}
        gt_rot = targets.get('absolute_pose_rot') if 'rot_3d' in terms else None
        if 'rot_3d' in terms and gt_rot is None:
            raise RuntimeError("loss mode %r needs targets['absolute_pose_rot']" % (loss_mode,))
        res = ops.fused_step(pose_inputs_6d, self._pose.tree, self._projection.camera.struct, ref_loc, ref_rot, gt2,
                             targets['absolute_pose_loc'], ref_index=self._ref_index, world_loc=world_loc,
                             world_rot=world_rot, mode=self._mode, hips=get_root_point(), neck=get_neck_point(),
                             normalize=normalize, mask_missing_joints=mask_missing_joints,
                             weight_2d=wt['loc_2d'], weight_3d=wt['loc_3d'], gt_rot=gt_rot, weight_rot=wt['rot_3d'])
        if world_loc is None:
            world_loc = self._world_locations.unsqueeze(1).repeat(1, l, 1)
            world_rot = self._world_rotations.unsqueeze(1).repeat(1, l, 1, 1)
        out = {
            'relative_pose_loc': self.get_reference_tensors()[0].unsqueeze(1).repeat(1, l, 1, 1),
            'relative_pose_rot': res.relative_pose_rot,
            'absolute_pose_loc': res.absolute_pose_loc,
            'absolute_pose_rot': res.absolute_pose_rot,
            'world_loc': world_loc,
            'world_rot': world_rot,
            'projection_2d_transformed': res.projection_2d_transformed,
            'loss': {'loc_2d': res.loc_2d, 'loc_3d': res.loc_3d, 'rot_3d': res.rot_3d, loss_mode: res.total},
            'loss_sums': res.losses[3:7],
        }
        return res.projection_2d, out
