"""LossExchange: how the ranks of a data-parallel job (one process per GPU, clips sharded by batch -- README.md:74-75
run shape) see each other's loss sums.

The reference lets Lightning reduce the logged losses across DDP ranks (modules/flow/base.py:467-468).  Here the step
kernel that finishes a rank's four sums {sum_sq_2d, count_2d, sum_sq_3d, count_3d} stores them straight into a small
block in EVERY rank's HBM (peer stores over NVLink, blocks mapped once through CUDA IPC): no collective kernel on the
step path, nothing waits for a peer, and the logger forms the global losses when it reads (`read`).  torch.distributed
is used once, at set-up, to pass the 64-byte IPC handles around.

Single-process use (tests, one GPU): `LossExchange.local(world)` makes `world` blocks on one device that name each other
directly -- the same kernels, no IPC."""
import ctypes as C

import torch

from . import _lib
from ._lib import check


class LossExchange(object):
    def __init__(self, rank, world, block, peers, owned, opened):
        self.rank, self.world = rank, world
        self.block = block            # this rank's block (device pointer, int)
        self._peers = peers           # every rank's block as this process addresses it
        self._owned, self._opened = owned, opened
        self.steps = 0                # steps published through the block so far (host-side count)

    # ------------------------------------------------------------------ construction
    @staticmethod
    def _create(want_handle):
        lib = _lib.load()
        p = C.c_void_p()
        handle = C.create_string_buffer(64) if want_handle else None
        check(lib.p2c_exchange_create(C.byref(p), handle), 'exchange_create')
        return p.value, (handle.raw if want_handle else None)

    @staticmethod
    def _set_peers(block, peers):
        arr = (C.c_void_p * len(peers))(*peers)
        check(_lib.load().p2c_exchange_set_peers(C.c_void_p(block), arr, len(peers)), 'exchange_set_peers')

    @classmethod
    def distributed(cls, group=None):
        """One block per rank; handles travel once through torch.distributed (any backend)."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        if world > int(_lib.load().p2c_exchange_max_ranks()):
            raise RuntimeError('p2c_b200: the exchange block holds at most %d ranks' % _lib.load().p2c_exchange_max_ranks())
        block, handle = cls._create(True)
        handles = [None] * world
        dist.all_gather_object(handles, handle, group=group)
        peers, opened = [], []
        for r in range(world):
            if r == rank:
                peers.append(block)
                continue
            q = C.c_void_p()
            check(_lib.load().p2c_exchange_open(handles[r], C.byref(q)), 'exchange_open (rank %d)' % r)
            peers.append(q.value)
            opened.append(q.value)
        cls._set_peers(block, peers)
        torch.cuda.synchronize()
        dist.barrier(group=group)     # every table is in place before any rank's first step
        return cls(rank, world, block, peers, [block], opened)

    @classmethod
    def local(cls, world, device=None):
        """`world` emulated ranks on ONE device in this process: returns a list of exchanges (one per rank)."""
        if device is not None:
            torch.cuda.set_device(device)
        blocks = [cls._create(False)[0] for _ in range(world)]
        for b in blocks:
            cls._set_peers(b, blocks)
        torch.cuda.synchronize()
        return [cls(r, world, blocks[r], blocks, [blocks[r]], []) for r in range(world)]

    # ------------------------------------------------------------------ use
    def attach(self, args):
        """Name this exchange in a p2c_fused_step argument block (ops.fused_step(exchange=...) does it for you)."""
        args.exchange = C.c_void_p(self.block)
        args.exchange_rank, args.exchange_world = self.rank, self.world
        return args

    def flush(self):
        """Send the sums of the last step (the step kernels send a step's sums at the start of the NEXT step's kernel,
        off their own tail).  Enqueue on the steps' stream before the synchronisation point that precedes read()."""
        check(_lib.load().p2c_exchange_flush(C.c_void_p(self.block), self.rank, self.world, _lib.stream()), 'exchange_flush')

    def read(self, seq=0, out=None):
        """(8,) device tensor {sum_sq_2d, count_2d, sum_sq_3d, count_3d, ranks arrived, global loc_2d, global loc_3d,
        seq} of step `seq` (1-based; 0 = the last step this rank published).  Enqueued on the current stream; never
        waits for a peer -- every rank calls flush(), synchronises and meets the others (the barrier that ends an epoch
        or a logging interval) before anyone reads; check [4] == world."""
        if out is None:
            out = torch.empty(8, device='cuda')
        check(_lib.load().p2c_exchange_read(C.c_void_p(self.block), self.world, int(seq) & 0xffffffff, _lib.ptr(out),
                                            _lib.stream()), 'exchange_read')
        return out

    def close(self):
        lib = _lib.load()
        for q in self._opened:
            lib.p2c_exchange_close(C.c_void_p(q))
        for b in self._owned:
            lib.p2c_exchange_destroy(C.c_void_p(b))
        self._opened, self._owned = [], []
