"""The in-kernel exchange of the loss sums between data-parallel ranks: the step kernels store each rank's
{sum_sq_2d, count_2d, sum_sq_3d, count_3d} into every rank's block; the reader sums them in rank order.  Parity bar:
the summed slots equal the oracle's UNSHARDED loc_2d / loc_3d.  One GPU: the ranks are emulated as blocks on one device
(same kernels, no IPC).  Two GPUs (skipped below 2 devices): two processes, blocks mapped through CUDA IPC."""
import os
import socket

import pytest
import torch

import oracle as O
from _util import assert_close, make_case, make_targets, oracle_step

pytestmark = pytest.mark.gpu


def _step(ops, lib, c, gt2, gt3, lo, hi, dev, exchange, **kw):
    cam = c['cam']
    f = lambda t: t.float().to(dev)   # noqa: E731
    return ops.fused_step(f(c['d6'][lo:hi]), lib.make_tree(c['par']),
                          lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py), f(c['loc']), f(c['rot']),
                          f(gt2[lo:hi]), f(gt3[lo:hi]), ref_index=c['sid'][lo:hi].int().to(dev), exchange=exchange, **kw)


@pytest.mark.parametrize('n,l,world', [(64, 16, 2), (700, 16, 4), (9, 5, 3)])
def test_emulated_ranks_sum_to_the_unsharded_losses(n, l, world):
    """Tile kernel (one tile per CTA and many tiles) and the fallback kernels (odd pose counts) all publish."""
    from pedestrians_video_2_carla_b200 import _lib, ops
    from pedestrians_video_2_carla_b200.exchange import LossExchange
    dev = 'cuda:0'
    c = make_case(n, l, 31 + n, world=False)
    gt2, gt3 = make_targets(c)
    want = oracle_step(c, O.rotation_6d_to_matrix(c['d6']), gt=(gt2, gt3))
    xs = LossExchange.local(world, dev)
    try:
        bounds = [n * r // world for r in range(world + 1)]
        for rep in range(3):   # three steps: the sequence numbers advance
            for r in range(world):
                _step(ops, _lib, c, gt2, gt3, bounds[r], bounds[r + 1], dev, xs[r])
        for r in range(world):
            xs[r].flush()             # a step's sums are sent by the next step's kernel: send the last ones
        torch.cuda.synchronize()
        for r in range(world):
            assert xs[r].steps == 3
            for seq in (0, 3, 2, 1):
                got = xs[r].read(seq).cpu()
                assert int(got[4]) == world and int(got[7]) == (seq or 3)
                assert_close('global loc_2d', got[5], want['loc_2d'])
                assert_close('global loc_3d', got[6], want['loc_3d'])
                assert float(got[3]) == n * l * 26 * 3
            assert int(xs[r].read(4).cpu()[4]) == 0     # a step nobody has published yet: nothing counted, no wait
        # every rank holds bit-identical totals (fixed rank order)
        assert all(torch.equal(xs[0].read().cpu(), xs[r].read().cpu()) for r in range(1, world))
    finally:
        for x in xs:
            x.close()


def test_exchange_rejects_bad_ranks():
    from pedestrians_video_2_carla_b200 import _lib, ops
    from pedestrians_video_2_carla_b200.exchange import LossExchange
    c = make_case(4, 4, 3, world=False)
    gt2, gt3 = make_targets(c)
    (x,) = LossExchange.local(1, 'cuda:0')
    try:
        x.rank = 5
        with pytest.raises(RuntimeError):
            _step(ops, _lib, c, gt2, gt3, 0, 4, 'cuda:0', x)
    finally:
        x.close()


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _rank_main(rank, world, port, n, l, seed, queue):
    import torch.distributed as dist
    from pedestrians_video_2_carla_b200 import _lib, ops
    from pedestrians_video_2_carla_b200.exchange import LossExchange
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    dev = 'cuda:%d' % rank
    c = make_case(n, l, seed, world=False)
    gt2, gt3 = make_targets(c)
    x = LossExchange.distributed()
    lo, hi = n * rank // world, n * (rank + 1) // world
    for _ in range(2):
        _step(ops, _lib, c, gt2, gt3, lo, hi, dev, x)
    x.flush()
    torch.cuda.synchronize()
    dist.barrier()            # the peers' steps are complete: their stores are in this rank's block
    queue.put((rank, x.read().cpu().tolist()))
    dist.barrier()
    x.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs (one process per GPU, CUDA IPC)')
def test_two_processes_two_gpus():
    import torch.multiprocessing as mp
    n, l, seed, world = 96, 16, 4141, 2
    c = make_case(n, l, seed, world=False)
    gt2, gt3 = make_targets(c)
    want = oracle_step(c, O.rotation_6d_to_matrix(c['d6']), gt=(gt2, gt3))
    ctx = mp.get_context('spawn')
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, n, l, seed, queue)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(queue.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for r in range(world):
        v = got[r]
        assert int(v[4]) == world and int(v[7]) == 2
        assert_close('global loc_2d (rank %d)' % r, torch.tensor(v[5]), want['loc_2d'])
        assert_close('global loc_3d (rank %d)' % r, torch.tensor(v[6]), want['loc_3d'])
    assert got[0] == got[1]
