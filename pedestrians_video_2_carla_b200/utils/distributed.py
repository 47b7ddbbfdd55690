"""Batch sharding of the hot path across ranks (one process per GPU).  Clips are independent, so there is no
data-path collective; the only exchange is the all-reduce of the four loss sums that p2c_fused_step leaves in
losses[3:7] = (sum_sq_2d, count_2d, sum_sq_3d, count_3d).  Local losses / gradients keep DDP semantics (local means,
DDP averages parameter gradients outside the path); `global_losses` gives the batch-global values for logging."""
import torch
import torch.distributed as dist


def shard_range(n_clips, rank, world_size):
    """Contiguous slice [lo, hi) of the clip axis owned by `rank` (remainder spread over the first ranks)."""
    base, rem = divmod(n_clips, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def global_losses(loss_sums, weight_2d=1.0, weight_3d=1.0, group=None):
    """loss_sums: tensor (4,) on this rank -> dict of batch-global loc_2d, loc_3d, loc_2d_3d after ONE all-reduce
    (NCCL over NVLink on GPUs; gloo in the CPU tests)."""
    s = loss_sums.detach().clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
    l2, l3 = s[0] / s[1], s[2] / s[3]
    return {'loc_2d': l2, 'loc_3d': l3, 'loc_2d_3d': weight_2d * l2 + weight_3d * l3, 'sums': s}
