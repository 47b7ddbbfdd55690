"""P3dPose: batched forward kinematics over a bone tree on the GPU, drop-in for reference
walker_control/p3d_pose.py (same constructor, forward / relative_to_absolute / tensors / relative / absolute / move /
pose_to_tensors / tensors_to_pose).  The reference walks the tree recursively with ~10 ATen ops per bone
(p3d_pose.py:116-184); here each call is one sm_100a kernel launch through the C ABI."""
import time
from collections import OrderedDict

import numpy as np
import torch

from .. import _lib, ops
from ..carla_utils.transforms import Location, Rotation, Transform
from .pose import Pose


class P3dPose(Pose, torch.nn.Module):
    def __init__(self, device=None, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._device = torch.device(device) if device is not None else torch.device('cuda', 0)
        self._tree = _lib.make_tree(self._parents)
        self.__relative_p3d_locations = None
        self.__relative_p3d_rotations = None
        self._last_rel_mod = time.time_ns()
        self.__last_rel = None
        self.__last_rel_get = None
        self._last_abs = None
        self._last_abs_mod = None

    @property
    def tree(self):
        return self._tree

    # ------------------------------------------------------------------ dict <-> tensor (p3d_pose.py:34-96)
    def pose_to_tensors(self, pose):
        """{bone: Transform} -> (locations (B,3), rotation matrices (B,3,3)) in p3d space: (x, y, -z) and
        euler_XYZ(deg2rad(-roll, -pitch, -yaw))."""
        locations = [(p.location.x, p.location.y, -p.location.z) for p in pose.values()]
        rotations = [(np.deg2rad(-p.rotation.roll), np.deg2rad(-p.rotation.pitch), np.deg2rad(-p.rotation.yaw))
                     for p in pose.values()]
        loc = torch.tensor(locations, device=self._device, dtype=torch.float32)
        rot = ops.euler_angles_to_matrix(torch.tensor(rotations, device=self._device, dtype=torch.float32), 'XYZ')
        return loc, rot

    def tensors_to_pose(self, p3d_locations, p3d_rotations):
        """Inverse of pose_to_tensors (p3d_pose.py:56-96)."""
        eul = -torch.rad2deg(ops.matrix_to_euler_angles(p3d_rotations, 'XYZ')).cpu().numpy()
        loc = p3d_locations.detach().cpu().numpy()
        return OrderedDict(
            (name, Transform(Location(float(l[0]), float(l[1]), -float(l[2])),
                             Rotation(pitch=float(e[1]), yaw=float(e[2]), roll=float(e[0]))))
            for name, l, e in zip(self._bone_names, loc, eul))

    # ------------------------------------------------------------------ FK
    def relative_to_absolute(self, loc, rot):
        """(N,B,3), (N,B,3,3) relative -> absolute (p3d_pose.py:151-184)."""
        _, abs_rot, abs_loc, _ = ops.fk_clip(rot.unsqueeze(1), loc, None, None, None, None, self._tree, None,
                                             _lib.MODE_RELATIVE_ROT, False)
        return abs_loc[:, 0], abs_rot[:, 0]

    def forward(self, changes_matrix, prev_relative_loc, prev_relative_rot):
        """(N,B,3,3) rotation changes, (N,B,3) relative locations, (N,B,3,3) previous relative rotations ->
        (absolute locations, absolute rotations, updated relative rotations) (p3d_pose.py:186-213)."""
        rel_rot, abs_rot, abs_loc, _ = ops.fk_clip(changes_matrix.unsqueeze(1), prev_relative_loc, prev_relative_rot,
                                                   None, None, None, self._tree, None, _lib.MODE_POSE_CHANGES, False)
        return abs_loc[:, 0], abs_rot[:, 0], rel_rot[:, 0]

    # ------------------------------------------------------------------ state (p3d_pose.py:215-285)
    @property
    def relative(self):
        if self._last_rel_mod != self.__last_rel_get:
            self.__last_rel = self.tensors_to_pose(self.__relative_p3d_locations, self.__relative_p3d_rotations)
            self.__last_rel_get = self._last_rel_mod
        return self.__last_rel

    @relative.setter
    def relative(self, new_pose_dict):
        self.__relative_p3d_locations, self.__relative_p3d_rotations = self.pose_to_tensors(new_pose_dict)
        self._last_rel_mod = time.time_ns()

    @property
    def absolute(self):
        if self._last_abs_mod != self._last_rel_mod:
            loc, rot = self.relative_to_absolute(self.__relative_p3d_locations.unsqueeze(0),
                                                 self.__relative_p3d_rotations.unsqueeze(0))
            self._last_abs = self.tensors_to_pose(loc[0], rot[0])
            self._last_abs_mod = self._last_rel_mod
        return OrderedDict(self._last_abs)

    def move(self, rotations):
        """{bone: Rotation} change in CARLA degrees, merged into the relative rotations (p3d_pose.py:243-263)."""
        changes = torch.zeros((len(self._bone_names), 3), dtype=torch.float32)
        for bone_name, r in rotations.items():
            changes[self._bone_names.index(bone_name)] = torch.tensor(
                (np.deg2rad(-r.roll), np.deg2rad(-r.pitch), np.deg2rad(-r.yaw)), dtype=torch.float32)
        changes_matrix = ops.euler_angles_to_matrix(changes.to(self._device), 'XYZ')
        # __move_to_relative (p3d_pose.py:98-114): rel_rot = change . prev_rel_rot, a one-frame scan
        self.__relative_p3d_rotations = self.forward(changes_matrix.unsqueeze(0),
                                                     self.__relative_p3d_locations.unsqueeze(0),
                                                     self.__relative_p3d_rotations.unsqueeze(0))[2][0]
        self._last_rel_mod = time.time_ns()

    @property
    def tensors(self):
        return (self.__relative_p3d_locations.detach().clone(), self.__relative_p3d_rotations.detach().clone())

    @tensors.setter
    def tensors(self, new_tensors):
        self.__relative_p3d_locations, self.__relative_p3d_rotations = new_tensors
        self._last_rel_mod = time.time_ns()
