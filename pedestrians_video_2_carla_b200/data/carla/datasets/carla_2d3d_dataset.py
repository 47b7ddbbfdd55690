"""Carla2D3DIterableDataset.generate_batch on the device (reference data/carla/datasets/carla_2d3d_dataset.py:145-210):
the synthetic-target generator that feeds the pose_lifting flow -- sparse random Euler changes per frame ->
euler_angles_to_matrix -> ProjectionModule -> transform -> (inputs, targets, meta).

Where the reference draws on the host (a Python loop over batch x clip with np.random.choice, :157-162) and runs its
frame-by-frame projection on CPU workers, this generator draws with ONE device kernel (counter-based Philox stream,
csrc/generator.cu) and runs the same kernels as the training step, so a batch never crosses PCIe and nothing
synchronises with the host.  The random stream is therefore NOT the reference's (numpy / CPU torch generators have no
device equivalent): parity is defined on the deterministic part -- given the drawn Euler changes,
every output equals the oracle's ProjectionModule / Normalizer on them (tests/test_gpu_widen.py) -- and on the
distribution's invariants (exactly `random_changes_each_frame` bones change per frame, |angle| <= max_change, only yaw
in the world rotation).  Deformation / augmentation mixins (missing-joint noise) are host-side data policy and are
not reproduced."""
import math

import torch

from .... import ops
from ....modules.layers.projection import ProjectionModule
from ....transforms.pose.normalization.hips_neck_bbox_fallback_extractor import HipsNeckBBoxFallbackExtractor
from ....transforms.pose.normalization.normalizer import Normalizer
from ..reference import CARLA_REFERENCE_SKELETON_TYPES
from ..skeleton import CARLA_SKELETON


class _LazyLabels(object):
    """meta['age'] / meta['gender'] of a generated batch: a sequence of strings backed by the device-side skeleton
    index; the one device->host copy happens only if somebody indexes or iterates it."""

    def __init__(self, index, field):
        self._index, self._field, self._host = index, field, None

    def _labels(self):
        if self._host is None:
            self._host = [CARLA_REFERENCE_SKELETON_TYPES[i][self._field] for i in self._index.tolist()]
        return self._host

    def __len__(self):
        return int(self._index.numel())

    def __getitem__(self, i):
        return self._labels()[i]

    def __iter__(self):
        return iter(self._labels())


class Carla2D3DIterableDataset(torch.utils.data.IterableDataset):
    def __init__(self, batch_size=64, clip_length=30, random_changes_each_frame=3, max_change_in_deg=5,
                 max_world_rot_change_in_deg=0, max_initial_world_rot_change_in_deg=0, points=CARLA_SKELETON,
                 transform='hips_neck_bbox', needs_confidence=False, device=None, seed=None, **kwargs):
        super().__init__()
        self.nodes = points
        self.batch_size, self.clip_length = batch_size, clip_length
        self.random_changes_each_frame = random_changes_each_frame
        self.max_change_in_rad = math.radians(max_change_in_deg)
        self.max_world_rot_change_in_rad = math.radians(max_world_rot_change_in_deg)
        self.max_initial_world_rot_change_in_rad = math.radians(max_initial_world_rot_change_in_deg)
        self.return_confidence = needs_confidence
        self.device = torch.device(device if device is not None else 'cuda:0')
        if self.device.type != 'cuda':
            raise RuntimeError('p2c_b200: the synthetic generator runs on a CUDA device (no CPU fallback)')
        self.seed = int(seed) if seed is not None else int(torch.seed())
        self._batches = 0     # the batch counter is the generator's only state (counter-based draws)
        if transform in (None, 'none'):
            self.transform = None
        elif transform == 'hips_neck':
            self.transform = Normalizer()
        elif transform == 'hips_neck_bbox':       # the reference's default (data/base/base_datamodule.py:61,276)
            self.transform = Normalizer(HipsNeckBBoxFallbackExtractor())
        else:
            self.transform = transform             # any callable (sample) -> sample
        self.projection = ProjectionModule()

    def draw_changes(self, batch_size):
        """(N, L, B, 3) Euler changes: `random_changes_each_frame` distinct bones per frame get angles uniform in
        +-max_change (reference :157-162: np.random.choice without replacement + torch.rand); (N, L, 3) world yaw
        changes; (N,) skeleton types (:181-182).  ONE kernel (ops.draw_sparse_changes), no host round trip."""
        out = ops.draw_sparse_changes(self.seed, self._batches, batch_size, self.clip_length, len(self.nodes),
                                      self.random_changes_each_frame, self.max_change_in_rad,
                                      self.max_world_rot_change_in_rad, self.max_initial_world_rot_change_in_rad,
                                      len(CARLA_REFERENCE_SKELETON_TYPES), self.device)
        self._batches += 1
        return out

    def generate_batch(self, batch_size=None):
        n = self.batch_size if batch_size is None else batch_size
        pose_changes, world_rot_change, skeleton_index = self.draw_changes(n)
        pose_changes_batch = ops.euler_angles_to_matrix(pose_changes, 'XYZ')
        world_rot_change_batch = ops.euler_angles_to_matrix(world_rot_change, 'XYZ')
        world_loc_change_batch = torch.zeros((n, self.clip_length, 3), device=self.device)
        # meta['age'] / meta['gender'] read like the reference's lists but only reach for the host when someone looks;
        # the projection layer takes the device-side row index directly
        meta = {'age': _LazyLabels(skeleton_index, 0), 'gender': _LazyLabels(skeleton_index, 1),
                'skeleton_index': skeleton_index}
        with torch.no_grad():
            self.projection.on_batch_start((pose_changes_batch, None, meta), 0)
            orig_projection_2d, projection_outputs = self.projection(pose_changes_batch, world_loc_change_batch,
                                                                     world_rot_change_batch)
            targets = {'projection_2d': orig_projection_2d[..., :2]}
            projection_2d = orig_projection_2d
            if self.transform is not None:
                projection_2d = self.transform(orig_projection_2d)
                # (a stride-3 view of the 3-channel records: ops.fused_step takes it as such, no copy)
                targets['projection_2d_transformed'] = projection_2d[..., :2]
        if not self.return_confidence:
            projection_2d = projection_2d[..., :2]
        targets.update(pose_changes=pose_changes_batch, world_loc_changes=world_loc_change_batch,
                       world_rot_changes=world_rot_change_batch, **projection_outputs)
        targets['euler_pose_changes'] = pose_changes
        return projection_2d, targets, meta

    def __iter__(self):
        while True:
            inputs, targets, meta = self.generate_batch(self.batch_size)
            for idx in range(self.batch_size):
                yield (inputs[idx], {k: v[idx] for k, v in targets.items()}, {k: v[idx] for k, v in meta.items()})
