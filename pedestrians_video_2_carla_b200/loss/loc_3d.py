"""loc_3d (reference loss/loc_3d.py:12-40): MSE('mean') of absolute joint locations; None when the data has no
'absolute_pose_loc' target."""
import warnings

from .. import ops
from ._criterion import require_mse


def calculate_loss_loc_3d(criterion=None, input_nodes=None, output_nodes=None, absolute_pose_loc=None, targets=None,
                          **kwargs):
    require_mse(criterion, 'loc_3d')
    if input_nodes is not output_nodes:
        raise NotImplementedError('cross-skeleton index mapping is outside the accelerated path')
    try:
        gt = targets['absolute_pose_loc']
    except KeyError:
        warnings.warn("The 'loc_3d' loss is not supported for this data, missing 'absolute_pose_loc' in targets.")
        return None
    return ops.masked_mse(absolute_pose_loc, gt, 3, -1)
