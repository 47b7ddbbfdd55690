"""pose_changes (reference loss/pose_changes.py:7-29 with nn.MSELoss(reduction='sum'), loss/__init__.py:33): the SUM
of squared differences between predicted and target pose changes.  The reduction kernel returns the mean; the element
count is a constant of the shapes (no mask)."""
from .. import ops
from ._criterion import require_mse


def calculate_loss_pose_changes(criterion=None, pose_inputs=None, targets=None, **kwargs):
    gt = targets['pose_changes']
    shape = pose_inputs.shape[:-2] + (9,)
    reduction = require_mse(criterion, 'pose_changes', ('sum', 'mean'))
    mean = ops.masked_mse(pose_inputs.reshape(shape), gt.reshape(shape), 9, -1)
    return mean * float(pose_inputs.numel()) if reduction == 'sum' else mean
