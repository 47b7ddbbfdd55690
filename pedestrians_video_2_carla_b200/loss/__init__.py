"""The losses of the hot path with the reference's call contract (reference loss/__init__.py:18-53,
modules/flow/base.py:440-469): each entry is (callable-or-class, criterion, dependencies); callables take
(criterion=, input_nodes=, output_nodes=, mask_missing_joints=, requirements=, loss_weights=, **sliced)."""
from enum import Enum

from torch import nn

from .cum_pose_changes import calculate_loss_cum_pose_changes
from .loc_2d import Loc2DPoseLoss
from .loc_2d_3d import calculate_loss_loc_2d_3d
from .loc_2d_loc_rot_3d import calculate_loss_loc_2d_loc_rot_3d
from .loc_3d import calculate_loss_loc_3d
from .loc_rot_3d import calculate_loss_loc_rot_3d
from .pose_changes import calculate_loss_pose_changes
from .rot_3d import calculate_loss_rot_3d
from .weighted_loc_2d_loc_rot_3d import calculate_loss_weighted_loc_2d_loc_rot_3d


class LossModes(Enum):
    loc_2d = (Loc2DPoseLoss, nn.MSELoss(reduction='mean'))
    loc_3d = (calculate_loss_loc_3d, nn.MSELoss(reduction='mean'))
    rot_3d = (calculate_loss_rot_3d, nn.MSELoss(reduction='mean'))
    cum_pose_changes = (calculate_loss_cum_pose_changes, nn.MSELoss(reduction='mean'))
    pose_changes = (calculate_loss_pose_changes, nn.MSELoss(reduction='sum'))
    loc_2d_3d = (calculate_loss_loc_2d_3d, None, ('loc_2d', 'loc_3d'))
    loc_2d_loc_rot_3d = (calculate_loss_loc_2d_loc_rot_3d, None, ('loc_2d', 'loc_3d', 'rot_3d'))
    weighted_loc_2d_loc_rot_3d = (calculate_loss_weighted_loc_2d_loc_rot_3d, None, ('loc_2d', 'loc_3d', 'rot_3d'))
    loc_rot_3d = (calculate_loss_loc_rot_3d, None, ('loc_3d', 'rot_3d'))
