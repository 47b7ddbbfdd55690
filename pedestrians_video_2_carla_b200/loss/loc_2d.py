"""loc_2d (reference loss/loc_2d.py:69-89 + loss/base_pose_loss.py:36-66): MSE('mean') between the first two channels
of the (transformed) projection and of the target, over the joints kept by utils/tensors.py:29-40.  The boolean-mask
gather + mse_loss of the reference is one masked two-stage reduction kernel here; 2-channel predictions that want a
gradient (the autoencoder flow's `pose_2d` output, modules/flow/autoencoder.py:102-112) get loss AND gradient from one
cooperative launch (ops.loc2d_loss -> p2c_loc2d_step)."""
from .. import ops
from ._criterion import require_mse
from ..data.carla.skeleton import CARLA_SKELETON, get_root_point


class Loc2DPoseLoss(object):
    def __init__(self, criterion=None, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON,
                 mask_missing_joints=True, **kwargs):
        if input_nodes is not output_nodes:
            raise NotImplementedError('cross-skeleton index mapping (get_common_indices) is outside the accelerated '
                                      'path: input and output skeletons must match')
        require_mse(criterion, 'loc_2d')
        self._mask = mask_missing_joints
        self._hips = get_root_point() if input_nodes is CARLA_SKELETON else input_nodes.get_hips_point().value

    def __call__(self, targets=None, projection_2d=None, projection_2d_transformed=None, **kwargs):
        assert 'projection_2d' in targets or 'projection_2d_transformed' in targets, \
            'Either projection_2d or projection_2d_transformed must be provided.'
        gt = targets['projection_2d_transformed'] if 'projection_2d_transformed' in targets else targets['projection_2d']
        assert projection_2d is not None or projection_2d_transformed is not None, \
            'Either projection_2d or projection_2d_transformed must be provided.'
        pred = projection_2d_transformed if projection_2d_transformed is not None else projection_2d
        return ops.loc2d_loss(pred, gt, self._hips if self._mask else -1)
