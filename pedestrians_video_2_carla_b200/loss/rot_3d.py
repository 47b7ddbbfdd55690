"""rot_3d (reference loss/rot_3d.py:9-37): MSE('mean') of the absolute bone rotation matrices; None when the
prediction or the 'absolute_pose_rot' target is missing.  One two-stage reduction kernel over 9 channels per bone; its
gradient enters the FK backward as g_abs_rot."""
from .. import ops
from ._criterion import require_mse


def calculate_loss_rot_3d(criterion=None, input_nodes=None, output_nodes=None, absolute_pose_rot=None, targets=None,
                          **kwargs):
    require_mse(criterion, 'rot_3d')
    if absolute_pose_rot is None or 'absolute_pose_rot' not in targets:
        return None
    if input_nodes is not output_nodes:
        raise NotImplementedError('cross-skeleton index mapping is outside the accelerated path')
    gt = targets['absolute_pose_rot']
    shape = absolute_pose_rot.shape[:-2] + (9,)
    return ops.masked_mse(absolute_pose_rot.reshape(shape), gt.reshape(shape), 9, -1)
