"""cum_pose_changes (reference loss/cum_pose_changes.py:9-56): MSE('mean') between the per-bone cumulative products of
the predicted and the target pose changes over the frames of a clip.  The reference's L bmm pairs + stack + transposes
are one scan kernel per operand (the target's has no autograd edge) and one reduction."""
from .. import ops
from ._criterion import require_mse


def calculate_loss_cum_pose_changes(criterion=None, pose_inputs=None, targets=None, **kwargs):
    require_mse(criterion, 'cum_pose_changes')
    pred = ops.cumulative_changes(pose_inputs)
    gt = ops.cumulative_changes(targets['pose_changes'].detach())
    shape = pred.shape[:-2] + (9,)
    return ops.masked_mse(pred.reshape(shape), gt.reshape(shape), 9, -1)
