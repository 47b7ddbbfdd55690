"""reference utils/tensors.py:29-53 (these two helpers are index/NaN bookkeeping, not kernels)."""
import torch


def get_missing_joints_mask(common_gt, hips=None, input_indices=None):
    mask = torch.all(common_gt != 0, dim=-1)
    if hips is not None:
        mask[..., hips if not hasattr(hips, 'value') else hips.value] = 1
    return mask


def nan_to_zero(t):
    return torch.nan_to_num(t, nan=0, posinf=0, neginf=0)


def eye_batch(batch_size, clip_length=None, device=None):
    shape = (batch_size, 1, 1) if clip_length is None else (batch_size, clip_length, 1, 1)
    return torch.eye(3, device=device).reshape((1,) * (len(shape) - 2) + (3, 3)).repeat(shape)
