"""Carla2D3DIterableDataset.generate_batch on the device (reference data/carla/datasets/carla_2d3d_dataset.py:145-210):
the synthetic-target generator that feeds the pose_lifting flow -- sparse random Euler changes per frame ->
euler_angles_to_matrix -> ProjectionModule -> transform -> (inputs, targets, meta).

Where the reference draws on the host (a Python loop over batch x clip with np.random.choice, :157-162) and runs its
frame-by-frame projection on CPU workers, this generator draws with the device RNG and runs the same kernels as the
training step, so a batch never crosses PCIe.  The random stream is therefore NOT the reference's (numpy / CPU torch
generators have no device equivalent): parity is defined on the deterministic part -- given the drawn Euler changes,
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
from ..skeleton import CARLA_SKELETON


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
        self.generator = torch.Generator(device=self.device)
        if seed is not None:
            self.generator.manual_seed(seed)
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
        +-max_change (reference :157-162: np.random.choice without replacement + torch.rand); (N, L, 3) world yaw."""
        n, l, b = batch_size, self.clip_length, len(self.nodes)
        g, dev = self.generator, self.device
        keys = torch.rand((n, l, b), device=dev, generator=g)
        chosen = keys.topk(self.random_changes_each_frame, dim=-1).indices       # distinct bones: a random subset
        angles = (torch.rand((n, l, self.random_changes_each_frame, 3), device=dev, generator=g) * 2 - 1) \
            * self.max_change_in_rad
        pose_changes = torch.zeros((n, l, b, 3), device=dev)
        pose_changes.scatter_(2, chosen.unsqueeze(-1).expand(-1, -1, -1, 3), angles)
        world_rot_change = torch.zeros((n, l, 3), device=dev)
        if self.max_initial_world_rot_change_in_rad > 0:
            world_rot_change[:, 0, 2] = (torch.rand((n,), device=dev, generator=g) * 2 - 1) \
                * self.max_initial_world_rot_change_in_rad
        if self.max_world_rot_change_in_rad != 0.0 and l > 1:
            world_rot_change[:, 1:, 2] = (torch.rand((n, l - 1), device=dev, generator=g) * 2 - 1) \
                * self.max_world_rot_change_in_rad
        return pose_changes, world_rot_change

    def generate_batch(self, batch_size=None):
        n = self.batch_size if batch_size is None else batch_size
        pose_changes, world_rot_change = self.draw_changes(n)
        pose_changes_batch = ops.euler_angles_to_matrix(pose_changes, 'XYZ')
        world_rot_change_batch = ops.euler_angles_to_matrix(world_rot_change, 'XYZ')
        world_loc_change_batch = torch.zeros((n, self.clip_length, 3), device=self.device)
        pick = torch.randint(0, 4, (n,), device=self.device, generator=self.generator).tolist()
        age = [('adult', 'child')[i >> 1] for i in pick]
        gender = [('male', 'female')[i & 1] for i in pick]
        with torch.no_grad():
            self.projection.on_batch_start((pose_changes_batch, None, {'age': age, 'gender': gender}), 0)
            orig_projection_2d, projection_outputs = self.projection(pose_changes_batch, world_loc_change_batch,
                                                                     world_rot_change_batch)
            targets = {'projection_2d': orig_projection_2d[..., :2]}
            projection_2d = orig_projection_2d
            if self.transform is not None:
                projection_2d = self.transform(orig_projection_2d)
                targets['projection_2d_transformed'] = projection_2d[..., :2]
        if not self.return_confidence:
            projection_2d = projection_2d[..., :2]
        targets.update(pose_changes=pose_changes_batch, world_loc_changes=world_loc_change_batch,
                       world_rot_changes=world_rot_change_batch, **projection_outputs)
        targets['euler_pose_changes'] = pose_changes
        return projection_2d, targets, {'age': age, 'gender': gender}

    def __iter__(self):
        while True:
            inputs, targets, meta = self.generate_batch(self.batch_size)
            for idx in range(self.batch_size):
                yield (inputs[idx], {k: v[idx] for k, v in targets.items()}, {k: v[idx] for k, v in meta.items()})
