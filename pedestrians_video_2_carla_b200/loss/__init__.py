"""The losses of the hot path with the reference's call contract (reference loss/__init__.py:18-53,
modules/flow/base.py:440-469): each entry is (callable-or-class, criterion, dependencies); callables take
(criterion=, input_nodes=, output_nodes=, mask_missing_joints=, requirements=, loss_weights=, **sliced)."""
from enum import Enum

from torch import nn

from .loc_2d import Loc2DPoseLoss
from .loc_2d_3d import calculate_loss_loc_2d_3d
from .loc_3d import calculate_loss_loc_3d


class LossModes(Enum):
    loc_2d = (Loc2DPoseLoss, nn.MSELoss(reduction='mean'))
    loc_3d = (calculate_loss_loc_3d, nn.MSELoss(reduction='mean'))
    loc_2d_3d = (calculate_loss_loc_2d_3d, None, ('loc_2d', 'loc_3d'))
