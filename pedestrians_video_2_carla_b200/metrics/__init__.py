"""Evaluation metrics of the pose_lifting flow on the fused step's output buffers (reference metrics/mpjpe.py,
mrpe.py, pck.py).  Same `update(predictions, targets)` / `compute()` / `reset()` protocol as the torchmetrics classes
they replace; the running state is two doubles on the device that the kernels accumulate into (no host sync in
`update`), so `dist_reduce_fx="sum"` is one all-reduce of `state` (see `sync`)."""
import torch

from .. import _lib
from .._lib import check, ptr, stream
from ..data.carla.skeleton import CARLA_SKELETON, get_neck_point, get_root_point
from ..utils.world import calculate_world_from_changes


def _f32(t, name):
    if not t.is_cuda:
        raise RuntimeError('p2c_b200: %s must be a CUDA tensor (no CPU fallback); got %s' % (name, t.device))
    return t.detach().float().contiguous()


class _KernelMetric(object):
    def __init__(self, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON, **kwargs):
        if input_nodes is not output_nodes:
            raise NotImplementedError('cross-skeleton index mapping is outside the accelerated path')
        self.input_nodes = input_nodes
        if input_nodes is CARLA_SKELETON:
            self._hips, self._neck = get_root_point(), get_neck_point()
        else:
            self._hips, self._neck = input_nodes.get_hips_point().value, input_nodes.get_neck_point().value
        self.state = None
        self._ws = None

    def _buffers(self, device):
        if self.state is None or self.state.device != device:
            self.state = torch.zeros(2, dtype=torch.float64, device=device)
            self._ws = torch.empty(_lib.load().p2c_metric_workspace_floats(), device=device)
        return self._ws, self.state

    def reset(self):
        if self.state is not None:
            self.state.zero_()

    def sync(self, group=None):
        """dist_reduce_fx='sum' of the torchmetrics original: all-reduce the two accumulators."""
        import torch.distributed as dist
        if self.state is not None and dist.is_available() and dist.is_initialized():
            dist.all_reduce(self.state, group=group)

    def __call__(self, predictions, targets):
        self.update(predictions, targets)
        return self.compute()


class MPJPE(_KernelMetric):
    """Mean per-joint position error in millimetres (metrics/mpjpe.py:9-49)."""

    def update(self, predictions, targets):
        try:
            pred, gt = predictions['absolute_pose_loc'], targets['absolute_pose_loc']
        except KeyError:
            return
        assert pred.shape == gt.shape
        pred, gt = _f32(pred, 'prediction'), _f32(gt, 'target')
        n, l, j, c = pred.shape
        ws, state = self._buffers(pred.device)
        check(_lib.load().p2c_mpjpe_sum(ptr(pred), ptr(gt), n * l, j, c, ptr(ws), ptr(state), 1, stream()), 'mpjpe_sum')

    def compute(self):
        # state = (sum of joint distances, number of them); the reference averages per clip over joints and frames and
        # then over clips -- the same number, every clip having L * J joints
        return (1000.0 * self.state[0] / self.state[1]).float()


class MRPE(_KernelMetric):
    """Mean root (hips) position error in millimetres (metrics/mrpe.py:9-80)."""

    def update(self, predictions, targets):
        try:
            pred, gt = predictions['absolute_pose_loc'], targets['absolute_pose_loc']
            assert pred.shape[:2] == gt.shape[:2]
            if predictions['world_loc_changes'] is not None:
                pw, _ = calculate_world_from_changes(predictions['world_loc_changes'].shape, pred.device,
                                                     predictions['world_loc_changes'])
            else:
                pw = predictions['world_loc']
            gw, _ = calculate_world_from_changes(pw.shape, pw.device, targets['world_loc_changes'])
        except KeyError:
            return
        assert pw.shape == gw.shape
        pred, gt, pw, gw = _f32(pred, 'prediction'), _f32(gt, 'target'), _f32(pw, 'world'), _f32(gw, 'world')
        n, l, j, c = pred.shape
        ws, state = self._buffers(pred.device)
        check(_lib.load().p2c_mrpe_sum(ptr(pred), ptr(gt), ptr(pw), ptr(gw), n * l, j, c, self._hips, ptr(ws),
                                       ptr(state), 1, stream()), 'mrpe_sum')

    def compute(self):
        return (1000.0 * self.state[0] / self.state[1]).float()


class PCK(_KernelMetric):
    """Percentage of correct keypoints (metrics/pck.py:13-97): joints whose distance to the target, divided by the
    target's hips-neck distance ('hn') or bounding-box diagonal (default), is below `threshold`."""

    def __init__(self, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON, mask_missing_joints=True,
                 key='projection_2d', threshold=0.05, get_normalization_tensor=None, **kwargs):
        super().__init__(input_nodes, output_nodes)
        if callable(get_normalization_tensor):
            raise NotImplementedError('custom normalisation callables are outside the accelerated path')
        self.key, self.threshold = key, threshold
        self._norm = 0 if get_normalization_tensor == 'hn' else 1
        self.mask_missing_joints = mask_missing_joints
        self.near_zero = 1e-5

    def update(self, predictions, targets):
        mask_src = targets['projection_2d'] if self.mask_missing_joints and 'projection_2d' in targets else None
        try:
            pred, gt = predictions[self.key], targets[self.key]
        except KeyError:
            return
        assert pred.shape == gt.shape
        pred, gt = _f32(pred, 'prediction'), _f32(gt, 'target')
        n, l, j, c = pred.shape
        cm = 0
        if mask_src is not None:
            mask_src = _f32(mask_src, 'projection_2d target')
            cm = mask_src.shape[-1]
        ws, state = self._buffers(pred.device)
        check(_lib.load().p2c_pck_count(ptr(pred), ptr(gt), ptr(mask_src), n * l, j, c, cm, self._hips, self._hips,
                                        self._neck, self._norm, self.threshold, self.near_zero, ptr(ws), ptr(state),
                                        1, stream()), 'pck_count')

    def compute(self):
        return (self.state[0] / self.state[1]).float()
