"""reference utils/world.py:16-63: absolute world transform per frame from per-frame changes."""
import torch

from .. import ops
from .tensors import eye_batch


def calculate_world_from_changes(shape, device, world_loc_change_batch=None, world_rot_change_batch=None,
                                 initial_world_loc=None, initial_world_rot=None):
    """-> (world_loc (N,L,3), world_rot (N,L,3,3)).  Both changes None: the initial transform broadcast over the clip
    (world.py:32-37); otherwise one scan kernel instead of L bmm + add steps (world.py:56-61)."""
    batch_size, clip_length = shape[0], shape[1]
    if initial_world_loc is None:
        initial_world_loc = torch.zeros((batch_size, 3), device=device)
    if initial_world_rot is None:
        initial_world_rot = eye_batch(batch_size, device=device)
    if world_loc_change_batch is None and world_rot_change_batch is None:
        return (initial_world_loc.unsqueeze(1).repeat(1, clip_length, 1),
                initial_world_rot.unsqueeze(1).repeat(1, clip_length, 1, 1))
    if world_loc_change_batch is None:
        world_loc_change_batch = torch.zeros((batch_size, clip_length, 3), device=device)
    if world_rot_change_batch is None:
        world_rot_change_batch = eye_batch(batch_size, clip_length, device=device)
    return ops.world_scan(world_loc_change_batch, world_rot_change_batch, initial_world_loc, initial_world_rot)
