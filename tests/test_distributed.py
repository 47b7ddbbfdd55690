"""N > 1 host logic on CPU: world_size-2 gloo run of the batch sharding + loss-sum all-reduce.  Each
rank evaluates its shard with the oracle (standing in for the kernel, which needs a GPU) and the all-reduced sums must
reproduce the unsharded losses."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def _sums(c, gt2, gt3, lo, hi):
    import oracle as O
    from _util import oracle_step
    sub = dict(c, n=hi - lo, sid=c['sid'][lo:hi], wl=c['wl'][lo:hi], wr=c['wr'][lo:hi])
    out = oracle_step(sub, O.rotation_6d_to_matrix(c['d6'][lo:hi]), gt=(gt2[lo:hi], gt3[lo:hi]))
    mask = O.missing_joints_mask(gt2[lo:hi][..., :2], c['hips'])
    c2 = float(mask.sum()) * 2
    c3 = float(gt3[lo:hi].numel())
    return torch.tensor([float(out['loc_2d']) * c2, c2, float(out['loc_3d']) * c3, c3], dtype=torch.float64)


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from _util import make_case, make_targets
    from pedestrians_video_2_carla_b200.utils.distributed import global_losses, shard_range
    c = make_case(7, 4, 99)          # 7 clips over 2 ranks: uneven shards (4 + 3)
    gt2, gt3 = make_targets(c)
    lo, hi = shard_range(c['n'], rank, world)
    g = global_losses(_sums(c, gt2, gt3, lo, hi))
    ret[rank] = (lo, hi, float(g['loc_2d']), float(g['loc_3d']), float(g['loc_2d_3d']))
    dist.destroy_process_group()


def test_sharded_loss_sums_allreduce_gloo():
    import oracle as O
    from _util import make_case, make_targets, oracle_step
    from pedestrians_video_2_carla_b200.utils.distributed import shard_range
    assert [shard_range(7, r, 2) for r in range(2)] == [(0, 4), (4, 7)]
    assert [shard_range(8, r, 4) for r in range(4)] == [(0, 2), (2, 4), (4, 6), (6, 8)]
    assert shard_range(1, 1, 2) == (1, 1)
    c = make_case(7, 4, 99)
    gt2, gt3 = make_targets(c)
    want = oracle_step(c, O.rotation_6d_to_matrix(c['d6']), gt=(gt2, gt3))
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, 29611, ret), nprocs=2, join=True)
    assert (ret[0][0], ret[0][1], ret[1][0], ret[1][1]) == (0, 4, 4, 7)
    for r in (0, 1):
        assert abs(ret[r][2] - float(want['loc_2d'])) < 1e-9 * max(1.0, abs(float(want['loc_2d'])))
        assert abs(ret[r][3] - float(want['loc_3d'])) < 1e-9 * max(1.0, abs(float(want['loc_3d'])))
        assert abs(ret[r][4] - float(want['loc_2d_3d'])) < 1e-9 * max(1.0, abs(float(want['loc_2d_3d'])))


def _metric_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    import oracle as O
    from pedestrians_video_2_carla_b200.metrics import MPJPE
    from pedestrians_video_2_carla_b200.utils.distributed import shard_range
    g = torch.Generator().manual_seed(5)
    gt = torch.randn((7, 4, 26, 3), generator=g, dtype=torch.float64)
    pred = gt + torch.randn((7, 4, 26, 3), generator=g, dtype=torch.float64) * 0.05
    lo, hi = shard_range(7, rank, world)
    m = MPJPE()
    # the accumulators a rank's kernel launches would have left (the oracle stands in for the kernel on CPU):
    # (sum of joint distances, number of them) over the rank's clips
    dists = torch.linalg.norm(pred[lo:hi] - gt[lo:hi], dim=-1)
    m.state = torch.tensor([float(dists.sum()), float(dists.numel())], dtype=torch.float64)
    m.sync()                                          # dist_reduce_fx="sum" of the torchmetrics original
    ret[rank] = float(m.compute())
    dist.destroy_process_group()


def test_metric_state_allreduce_gloo():
    """MPJPE over uneven shards (4 + 3 clips): the all-reduced accumulators give the unsharded metric on every rank."""
    import oracle as O
    g = torch.Generator().manual_seed(5)
    gt = torch.randn((7, 4, 26, 3), generator=g, dtype=torch.float64)
    pred = gt + torch.randn((7, 4, 26, 3), generator=g, dtype=torch.float64) * 0.05
    want = float(O.metric_mpjpe(pred, gt))
    ret = mp.Manager().dict()
    mp.spawn(_metric_worker, args=(2, 29613, ret), nprocs=2, join=True)
    for r in (0, 1):
        assert abs(ret[r] - want) < 1e-5 * want
