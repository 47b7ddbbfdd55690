#!/usr/bin/env python
"""bench.py -- poses/sec of the FK + projection + loc_2d_3d hot path, forward + backward (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm (oracle port) on host cores

A "step" is one pass of the hot path over one batch: rotation_6d model output (N,L,B,6) -> relative-rotation scan
-> FK over the CARLA_SKELETON tree -> pinhole projection -> hips/neck normalisation -> loc_2d + loc_3d + loc_2d_3d
-> gradient w.r.t. the 6-D input.  Default workload = BASELINE.json configs[1]: 256 clips x 16 frames per GPU
(weak scaling: every rank runs its own 256x16 shard; the only exchange is the four loss sums, which the step kernel
itself stores into every rank's HBM over NVLink -- no collective on the step path).  One JSON line, printed by rank 0.
`sweep` holds the other BASELINE configs: C3 (autoencoder flow: 2-D masked MSE fwd+bwd), C4 (4096x64, forward-only and
fwd+bwd), C5 (8192x16 per GPU = 1M poses over 8 GPUs), and the module-API (separate ops) step.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = 'poses/sec FK+projection+loss fwd+bwd'
B = 26
SEED = 22742  # reference default seed (modeling.py:120)
HIPS, NECK = 1, 8


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=400)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--clips', type=int, default=256, help='clips per GPU (BASELINE configs[1]: 256)')
    ap.add_argument('--clip-len', type=int, default=16)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-sweep', action='store_true', help='skip the other BASELINE configs (C3 / C4 / C5 sizes)')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-exchange', action='store_true', help='debug: N > 1 without the in-kernel loss-sum exchange')
    ap.add_argument('--cpu-seconds', type=float, default=12.0, help='budget of the bounded CPU sample')
    return ap.parse_args()


def algorithmic_bytes_per_pose(kind='fused', clip_len=16, world=False):
    """DESIGN.md 'Algorithmic bytes' (fp32, 26 bones; the per-clip skeleton row is a 4-row table: amortised to ~0).
    fused   : one pass, 6-D entry fused, module outputs materialised: read d6 624 + gt abs loc 312 + gt 2-D 208 (the two
              channels the loss reads) [+48 world]; write rel_rot 936 + abs_rot 936 + abs_loc 312 + projection 312 +
              grad 624                                                                          = 4264 B/pose
    forward : module API forward: read changes 936 + world 48 + 1248/L; write 2496  = 3480 + 1248/L
    loc2d   : autoencoder flow (C3): read pred 208 + gt 208, write grad 208                      = 624 B/pose"""
    if kind == 'fused':
        return 624 + 312 + 208 + (48 if world else 0) + 2496 + 624
    if kind == 'forward':
        return 936 + 48 + 1248.0 / clip_len + 2496
    if kind == 'loc2d':
        return 624
    raise ValueError(kind)


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler(object):
    """Polls NVML (sm clock, throttle reasons) from a thread while the timed region runs."""

    BITS = (('hw_slowdown', 0x8), ('sw_thermal_slowdown', 0x20), ('hw_thermal_slowdown', 0x40),
            ('sw_power_cap', 0x4), ('hw_power_brake_slowdown', 0x80))

    def __init__(self, index, period=0.002):
        self.samples, self.reasons, self.period, self.ok = [], set(), period, False
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # NVML missing: report it instead of inventing clocks
            self.err = repr(e)
        self.thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for name, bit in self.BITS:
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self.ok:
            self.thread.start()

    def stop(self):
        self._stop.set()
        if self.ok and self.thread.is_alive():
            self.thread.join()

    def summary(self, note):
        if not self.ok:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'error': self.err}
        s = sorted(self.samples)
        return {'sm_mhz': (s[len(s) // 2] if s else None), 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons),
                'samples': len(s), 'sampled': note}


# ------------------------------------------------------------------------------------------ the b200 arm
class GraphedSteps(object):
    """Steps over `n_sets` rotating buffer sets (their footprint together exceeds L2, so every timed step streams from
    HBM), replayed as CUDA graphs.  ANY step count runs the same way: one graph holds the steps back to back (cycling
    through the buffer sets; long runs are cut into graphs of CHUNK steps); inside a graph consecutive steps chain
    through programmatic dependent launch where the kernel supports it.  Nothing but the step's own kernels is ever
    on the stream: no collective, no host work between steps."""

    n_sets = 1
    launches_per_step = 1

    def launch(self, i):   # enqueue step i (buffer set i % n_sets) on the current stream
        raise NotImplementedError

    def _graph(self, count):
        cache = self.__dict__.setdefault('_graphs', {})
        if count not in cache:
            st = torch.cuda.Stream(device=self.dev)
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):   # warm the launch path (kernel attributes, lazy module load) outside capture
                for i in range(min(count, 2)):
                    self.launch(i)
            torch.cuda.current_stream().wait_stream(st)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for i in range(count):
                    self.launch(i)
            torch.cuda.synchronize()
            cache[count] = g
        return cache[count]

    CHUNK = 240   # steps per graph for long runs (a multiple of every n_sets used here: 2, 3, 4, 12, 16 -> 240)

    def _plan(self, count):
        """Graph sizes whose replays make `count` steps: chunks of CHUNK steps, then ONE graph with the rest, so any
        step count (the driver's 20 as well as 400) is a handful of graph launches of back-to-back chained steps."""
        return [self.CHUNK] * (count // self.CHUNK) + ([count % self.CHUNK] if count % self.CHUNK else [])

    def prepare(self, count):
        """Capture (untimed) every graph run_steps(count) replays."""
        for c in set(self._plan(count)):
            self._graph(c)

    def run_steps(self, count):
        for c in self._plan(count):
            self._graph(c).replay()

    def drop_graphs(self):
        self.__dict__.pop('_graphs', None)


def n_sets_for(bytes_per_set, want_bytes=192 << 20, max_sets=16):
    return int(max(2, min(max_sets, -(-want_bytes // max(int(bytes_per_set), 1)))))


def skeleton_tables(dev):
    """Reference skeleton table (4 x 26 bones), bone tree and training camera from the PACKAGE's own loaders
    (data/carla/reference.py, walker_control/p3d_pose_projection.py): the product arm never touches oracle/."""
    from pedestrians_video_2_carla_b200.data.carla import reference as R
    from pedestrians_video_2_carla_b200.walker_control.p3d_pose_projection import P3dPoseProjection
    pose = R.get_poses(dev)[0]
    loc, rot = R.get_relative_tensors(dev)
    cam = P3dPoseProjection(device=dev, pedestrian=pose).camera.struct
    return pose.tree, cam, loc.contiguous(), rot.contiguous()


def synthetic_inputs(n_clips, clip_len, g):
    """Seeded CPU draws shared by BOTH arms: skeleton ids, 6-D head output near identity, the independent
    6-D draw the targets are made from, and the joints to drop from the 2-D target (p = 0.1, never the hips)."""
    base = torch.tensor([1., 0, 0, 0, 1, 0])
    sid = torch.randint(0, 4, (n_clips,), generator=g)
    d6 = torch.randn((n_clips, clip_len, B, 6), generator=g) * 0.3 + base
    d6_gt = torch.randn((n_clips, clip_len, B, 6), generator=g) * 0.3 + base
    drop = torch.rand((n_clips, clip_len, B), generator=g) < 0.1
    drop[..., HIPS] = False
    return sid, d6, d6_gt, drop


class FusedShard(GraphedSteps):
    """One rank's shard of the headline workload: p2c_fused_step per buffer set, direct C-ABI calls."""

    def __init__(self, dev, n_clips, clip_len, seed, want_bytes=192 << 20, max_sets=16, exchange=None):
        from pedestrians_video_2_carla_b200 import _lib, ops
        self.lib, self.cdll, self.ops, self.dev = _lib, _lib.load(), ops, dev
        self.N, self.L = n_clips, clip_len
        self.poses = n_clips * clip_len
        self.tree, self.cam, self.loc, self.rot = skeleton_tables(dev)
        self.exchange = exchange
        self.n_sets = n_sets_for(self.poses * (algorithmic_bytes_per_pose('fused') + 4 * B), want_bytes, max_sets)
        self.loss_block = torch.zeros((self.n_sets, 8), device=dev)    # row i = the losses[8] of buffer set i
        g = torch.Generator(device='cpu').manual_seed(seed)
        self.sets = []
        for _ in range(self.n_sets):
            s = {}
            sid, d6, d6_gt, drop = synthetic_inputs(n_clips, clip_len, g)
            s['sid'], s['d6'] = sid.int().to(dev), d6.to(dev)
            zeros = torch.zeros((n_clips, clip_len, B, 3), device=dev)
            t = ops.fused_step(d6_gt.to(dev), self.tree, self.cam, self.loc, self.rot, zeros, zeros, ref_index=s['sid'])
            gt2 = t.projection_2d_transformed.detach().clone()
            gt2[drop.to(dev)] = 0
            s['gt2'], s['gt3'] = gt2, t.absolute_pose_loc.detach().clone()
            s['rel_rot'] = torch.empty((n_clips, clip_len, B, 3, 3), device=dev)
            s['abs_rot'] = torch.empty_like(s['rel_rot'])
            s['abs_loc'] = torch.empty((n_clips, clip_len, B, 3), device=dev)
            s['proj'] = torch.empty_like(s['abs_loc'])
            s['g_d6'] = torch.empty_like(s['d6'])
            s['losses'] = self.loss_block[len(self.sets)]
            s['args'] = self._args(s)
            self.sets.append(s)
        nws = self.cdll.p2c_fused_step_workspace_floats(self.poses, B, 1)
        ws = torch.zeros(max(int(nws), 1), device=dev)   # scratch, shared by all sets (steps are stream-ordered)
        for s in self.sets:
            s['ws'] = ws
            s['args'].workspace = _lib.ptr(ws)
        self.launches_per_step = int(self.cdll.p2c_fused_step_launches(self.sets[0]['args']))

    def _args(self, s):
        L = self.lib
        a = L.FusedStep()
        a.tree, a.hips, a.neck = self.tree, HIPS, NECK
        a.n_clips, a.clip_len, a.mode, a.input_6d, a.normalize, a.mask_missing_joints = self.N, self.L, 0, 1, 1, 1
        a.pose_in, a.ref_loc, a.ref_rot, a.ref_index = L.ptr(s['d6']), L.ptr(self.loc), L.ptr(self.rot), L.ptr(s['sid'])
        a.world_loc, a.world_rot, a.camera = None, None, self.cam       # ZeroTrajectory: identity world transform
        a.gt_2d, a.gt_2d_channels, a.gt_3d = L.ptr(s['gt2']), 3, L.ptr(s['gt3'])
        a.weight_2d, a.weight_3d = 1.0, 1.0
        a.rel_rot, a.abs_rot, a.abs_loc, a.proj, a.proj_norm = (L.ptr(s['rel_rot']), L.ptr(s['abs_rot']),
                                                                L.ptr(s['abs_loc']), L.ptr(s['proj']), None)
        a.g_pose_in, a.losses = L.ptr(s['g_d6']), L.ptr(s['losses'])
        if self.exchange is not None:
            self.exchange.attach(a)
        return a

    def launch(self, i):
        """Enqueue step i on the current stream (direct C-ABI call; no torch ops)."""
        self.lib.check(self.cdll.p2c_fused_step(self.sets[i % self.n_sets]['args'], self.lib.stream()), 'fused_step')


class Loc2dShard(GraphedSteps):
    """BASELINE configs[2] (autoencoder flow): the model's `pose_2d` output (N,L,26,2) against the 2-D targets,
    Loc2DPoseLoss forward + gradient = ONE p2c_loc2d_step launch per step (624 B/pose)."""

    def __init__(self, dev, n_clips, clip_len, seed, want_bytes=192 << 20, max_sets=64, exchange=None):
        import ctypes as C
        from pedestrians_video_2_carla_b200 import _lib
        self.lib, self.cdll, self.dev, self.C = _lib, _lib.load(), dev, C
        self.N, self.L, self.poses = n_clips, clip_len, n_clips * clip_len
        self.exchange = exchange
        self.n_sets = n_sets_for(self.poses * 624, want_bytes, max_sets)
        g = torch.Generator(device='cpu').manual_seed(seed)
        self.ws = torch.zeros(int(self.cdll.p2c_loc2d_step_workspace_floats()), device=dev)
        self.out = torch.zeros((self.n_sets, 4), device=dev)
        self.sets = []
        for _ in range(self.n_sets):
            gt = torch.randn((n_clips, clip_len, B, 2), generator=g)
            drop = torch.rand((n_clips, clip_len, B), generator=g) < 0.1
            drop[..., HIPS] = False
            gt[drop] = 0
            pred = gt + 0.1 * torch.randn((n_clips, clip_len, B, 2), generator=g)
            self.sets.append({'pred': pred.to(dev), 'gt': gt.to(dev), 'grad': torch.empty((n_clips, clip_len, B, 2), device=dev)})
        if not self.cdll.p2c_loc2d_step_supported(self.poses, B, 2):
            raise RuntimeError('p2c_loc2d_step does not take %d poses' % self.poses)

    def launch(self, i):
        s, L, x = self.sets[i % self.n_sets], self.lib, self.exchange
        L.check(self.cdll.p2c_loc2d_step(L.ptr(s['pred']), L.ptr(s['gt']), self.poses, B, 2, HIPS, 1.0, L.ptr(s['grad']),
                                         L.ptr(self.ws), L.ptr(self.out[i % self.n_sets]),
                                         self.C.c_void_p(x.block) if x is not None else None,
                                         x.rank if x is not None else 0, x.world if x is not None else 0, L.stream()),
                'loc2d_step')


class ForwardShard(GraphedSteps):
    """BASELINE configs[3]: FK + projection FORWARD through the module API's entry (p2c_fk_fwd: matrix input, the four
    module tensors materialised), mixed adult/child male/female skeletons per clip."""

    def __init__(self, dev, n_clips, clip_len, seed, want_bytes=192 << 20, max_sets=4):
        from pedestrians_video_2_carla_b200 import _lib, ops
        self.lib, self.cdll, self.dev = _lib, _lib.load(), dev
        self.N, self.L, self.poses = n_clips, clip_len, n_clips * clip_len
        self.tree, self.cam, self.loc, self.rot = skeleton_tables(dev)
        self.n_sets = n_sets_for(self.poses * algorithmic_bytes_per_pose('forward', clip_len), want_bytes, max_sets)
        g = torch.Generator(device='cpu').manual_seed(seed)
        self.sets = []
        eye = torch.eye(3, device=dev)
        for _ in range(self.n_sets):
            s = {}
            sid, d6, _, _ = synthetic_inputs(n_clips, clip_len, g)
            s['sid'] = sid.int().to(dev)
            s['mats'] = ops.rotation_6d_to_matrix(d6.to(dev)).detach().contiguous()
            s['wl'] = torch.zeros((n_clips, clip_len, 3), device=dev)
            s['wr'] = eye.expand(n_clips, clip_len, 3, 3).contiguous()
            s['rel_rot'] = torch.empty((n_clips, clip_len, B, 3, 3), device=dev)
            s['abs_rot'] = torch.empty_like(s['rel_rot'])
            s['abs_loc'] = torch.empty((n_clips, clip_len, B, 3), device=dev)
            s['proj'] = torch.empty_like(s['abs_loc'])
            a = _lib.FkFwd()
            a.tree, a.n_clips, a.clip_len, a.mode = self.tree, n_clips, clip_len, 0
            P = _lib.ptr
            a.pose_in, a.ref_loc, a.ref_rot, a.ref_index = P(s['mats']), P(self.loc), P(self.rot), P(s['sid'])
            a.world_loc, a.world_rot, a.camera = P(s['wl']), P(s['wr']), self.cam
            a.rel_rot, a.abs_rot, a.abs_loc, a.proj = P(s['rel_rot']), P(s['abs_rot']), P(s['abs_loc']), P(s['proj'])
            s['args'] = a
            self.sets.append(s)

    def launch(self, i):
        self.lib.check(self.cdll.p2c_fk_fwd(self.sets[i % self.n_sets]['args'], self.lib.stream()), 'fk_fwd')


def dist_setup(args):
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29533')
        backend = 'nccl' if (args.impl == 'b200' and torch.cuda.is_available()) else 'gloo'
        if backend == 'nccl':
            torch.cuda.set_device(local)
        dist.init_process_group(backend, rank=rank, world_size=world)
    return world, rank, local


def timed_region(shard, steps, warmup, world):
    """W untimed steps, then exactly K steps between CUDA events on the launching stream, barrier + synchronize on
    both sides, max over ranks.  The timed stream carries nothing but the steps' own kernels (for N > 1 the step kernel
    itself stores the four loss sums into every rank's block: no collective, no host work between steps).
    A fixed-length busy kernel in front of the start event keeps the host's launch latency out of the device time.
    Returns milliseconds for the K steps."""
    import torch.distributed as dist
    shard.prepare(warmup)
    shard.prepare(steps)
    shard.run_steps(warmup)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # A fixed-length busy kernel (~0.2 ms) goes first, so that the K steps are already queued behind it when the start
    # event fires: the events bracket the DEVICE time of exactly K steps, not the host's graph-launch latency (tens of
    # microseconds, i.e. several 14 us steps, which an idle stream would otherwise spend waiting for the host).
    # (long enough for the host to get through the FIRST launch of the K-step graphs -- their upload grows with the
    # node count: with a fixed 0.2 ms gate a 240-step graph's first replay read 10.5 us per step against 9.8)
    torch.cuda._sleep(int(400000 + 6000 * steps * max(int(getattr(shard, 'launches_per_step', 1)), 1)))
    e0.record()
    shard.run_steps(steps)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=shard.dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def e2e_region(shard, steps, warmup, world, exchange):
    """The same metric through the public API (ops.fused_step + backward) with HOST inputs: every step copies that
    step's d6 / targets / skeleton ids from pinned host memory, runs forward + backward, and reads the step's loss
    block (loc_2d, loc_3d, loc_2d_3d and the four sums: 8 floats) back."""
    import torch.distributed as dist
    ops, dev = shard.ops, shard.dev
    # Host side of a static-shape training loop: every step's inputs (6-D output stand-in, 2-D / 3-D targets, skeleton
    # ids) sit in ONE pinned staging image and cross PCIe as one copy into one of two device images; the copy of step
    # i+1 runs on a copy stream while step i computes and its loss is read back (a prefetching loader).  Outputs are
    # preallocated (ops.fused_step_buffers), so the call reuses its argument block.
    keys = ('d6', 'gt2', 'gt3', 'sid')
    ref = shard.sets[0]
    offs, total = {}, 0
    for k in keys:
        offs[k] = total
        total += (ref[k].numel() * ref[k].element_size() + 255) // 256 * 256

    def views(buf):
        return {k: buf[offs[k]:offs[k] + ref[k].numel() * ref[k].element_size()].view(ref[k].dtype).view(ref[k].shape)
                for k in keys}
    host = []
    for s in shard.sets[:4]:
        hb = torch.empty(total, dtype=torch.uint8).pin_memory()
        hv = views(hb)
        for k in keys:
            hv[k].copy_(s[k])
        host.append(hb)
    dimg = [torch.empty(total, dtype=torch.uint8, device=dev) for _ in range(2)]
    dbufs = [views(d) for d in dimg]
    h2d = sum(ref[k].numel() * ref[k].element_size() for k in keys)
    n, l, b = ref['d6'].shape[:3]
    pre = ops.fused_step_buffers(n, l, b, dev)
    loss_host = torch.zeros(8).pin_memory()
    copy_stream = torch.cuda.Stream(dev)
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    for e in consumed:
        e.record()

    def prefetch(i):
        k = i % 2
        copy_stream.wait_event(consumed[k])
        with torch.cuda.stream(copy_stream):
            dimg[k].copy_(host[i % len(host)], non_blocking=True)
        ready[k].record(copy_stream)

    def one(i, last):
        k = i % 2
        if not last:
            prefetch(i + 1)
        torch.cuda.current_stream().wait_event(ready[k])
        dbuf = dbufs[k]
        x = dbuf['d6'].requires_grad_(True)
        r = ops.fused_step(x, shard.tree, shard.cam, shard.loc, shard.rot, dbuf['gt2'], dbuf['gt3'], ref_index=dbuf['sid'],
                           want_transformed=False, out=pre, exchange=exchange)
        r.backward()            # = r.loc_2d_3d.backward(): the gradient is already in the step's output block
        consumed[k].record()
        loss_host.copy_(r.losses, non_blocking=False)   # device -> host read of the step's result
        dbuf['d6'] = x.detach()
        return float(loss_host[2])
    prefetch(0)
    for i in range(warmup):
        one(i, i == warmup - 1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    prefetch(0)
    for i in range(steps):
        one(i, i == steps - 1)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item()), h2d, 32


# ------------------------------------------------------------------------------------------ CPU arm (oracle port)
def cpu_reference(n_clips, clip_len, steps, warmup, seconds=None, threads=None):
    """The reference's algorithm (oracle/: per-bone recursion, per-frame Python loops, torch CPU ops, autograd
    backward) on this box's host cores, on the SAME synthetic inputs as the b200 arm's first buffer set.
    `steps` timed steps after `warmup`; with `seconds` set the loop also stops once that budget is spent (>= 3 steps).
    Returns (poses_per_sec, threads, steps timed, poses per step)."""
    import oracle as O
    if threads:
        torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(SEED)
    loc, rot, _, par = O.reference_tensors(torch.float32)
    cam = O.training_camera()
    sid, d6, d6_gt, drop = synthetic_inputs(n_clips, clip_len, g)
    d6.requires_grad_(True)
    zeros = torch.zeros(n_clips, clip_len, B, 3)
    with torch.no_grad():
        t = O.full_step(O.rotation_6d_to_matrix(d6_gt), loc[sid], rot[sid], par, cam, zeros, zeros)
    gt2, gt3 = t['projection_2d_transformed'].clone(), t['absolute_pose_loc'].clone()
    gt2[drop] = 0
    ref_loc, ref_rot = loc[sid], rot[sid]

    def one():
        d6.grad = None
        out = O.full_step(O.rotation_6d_to_matrix(d6), ref_loc, ref_rot, par, cam, gt2, gt3)
        out['loc_2d_3d'].backward()
        return float(out['loc_2d_3d'].detach())
    for _ in range(max(1, warmup)):
        one()
    times = []
    t_end = time.perf_counter() + (seconds if seconds else 1e9)
    while len(times) < steps and (time.perf_counter() < t_end or len(times) < 3):
        t0 = time.perf_counter()
        one()
        times.append(time.perf_counter() - t0)
    return n_clips * clip_len * len(times) / sum(times), torch.get_num_threads(), len(times), n_clips * clip_len


def config_of(args, world):
    """The workload description: IDENTICAL in both arms (the driver compares it)."""
    poses = args.clips * args.clip_len
    per_set = poses * (algorithmic_bytes_per_pose('fused') + 4 * B)
    n_sets = n_sets_for(per_set)
    l2 = '%d rotating buffer sets per GPU (%.0f MB > 126 MB L2), no flush kernel' % (n_sets, n_sets * per_set / 1e6)
    return {'workload': 'pose_lifting step hot path (6-D output -> scan -> FK -> projection -> hips/neck '
                        'normalise -> loc_2d_3d, fwd+bwd), %dx%d CARLA_SKELETON poses per GPU'
                        % (args.clips, args.clip_len),
            'clips_per_gpu': args.clips, 'clip_len': args.clip_len, 'bones': B, 'mode': 'pose_changes',
            'mask_missing_joints': True, 'transform': 'hips_neck', 'l2': l2,
            'parallelism': 'clips sharded by batch, %d rank(s); exchange of 4 loss sums' % world}


def measured_peak():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'MEASURED_PEAKS.json hbm_gbs (of measured)'
    except Exception:
        return 6650.0, 'B200_PROFILING.md fallback 6.65 TB/s (of fallback)'


# dram__bytes_read.sum + dram__bytes_write.sum of the step kernel, one launch, from the committed `ncu --set full`
# captures (profiles/*_tile_<clips>x<len>.txt); writes still in L2 when the kernel ends are not in it
NCU_DRAM_BYTES = {('fused', 256, 16): 5.192e6, ('fused', 8192, 16): 164.17e6 + 352.72e6,
                  ('fused', 4096, 64): 327.82e6 + 760.78e6}   # profiles/r02f_tile_*.txt (final build of round 2)


def roofline_of(kind, poses, clip_len, ms_per_step, note):
    peak, src = measured_peak()
    bpp = algorithmic_bytes_per_pose(kind, clip_len)
    achieved = poses * bpp / (ms_per_step * 1e-3) / 1e9
    return {'bound': 'hbm', 'achieved': round(achieved, 1), 'peak': peak, 'unit': 'GB/s',
            'frac': round(achieved / peak, 4), 'traffic': NCU_DRAM_BYTES.get((kind, poses // clip_len, clip_len)),
            'traffic_unit': 'bytes per launch (ncu, profiles/)', 'bytes_per_pose': round(bpp, 1), 'peak_source': src,
            'kernel': note}


def main_reference(args, world, rank):
    """--impl reference: the reference's CPU algorithm (oracle port; pytorch3d / pedestrians_scenarios are not
    installable here) on all host threads, same workload, same --steps / --warmup.  Rank 0 only."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    # every step is one full batch of the workload; the run is bounded by the step count the driver asks for (a
    # 256x16 step is ~0.15 s on 16 cores) and, for very long requests, by a wall-clock budget
    pps, used, iters, per = cpu_reference(args.clips, args.clip_len, args.steps, args.warmup, seconds=150.0)
    line = {'impl': 'reference', 'metric': METRIC, 'value': round(pps, 1), 'unit': 'poses/s', 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': round(per / pps * 1e3, 3),
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': config_of(args, args.gpus),
            'cpu_baseline': {'value': round(pps, 1), 'unit': 'poses/s', 'cores': used, 'kind': 'port',
                             'sample': '%d full %dx%d steps of oracle.full_step + autograd backward%s'
                                       % (iters, args.clips, args.clip_len,
                                          '' if iters == args.steps else ' (150 s budget reached before %d)' % args.steps)},
            'e2e': {'value': round(pps, 1), 'unit': 'poses/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    world, rank, local = dist_setup(args)
    if args.impl == 'reference':
        main_reference(args, world, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference '
                         'for the CPU arm)')
    import __graft_entry__
    from pedestrians_video_2_carla_b200 import _build
    if not os.path.exists(_build.LIB_PATH):
        if rank == 0:
            __graft_entry__.build()
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
    dev = torch.device('cuda', local)
    torch.cuda.set_device(dev)
    exchange = None
    if world > 1 and not args.no_exchange:
        from pedestrians_video_2_carla_b200.exchange import LossExchange
        exchange = LossExchange.distributed()
    shard = FusedShard(dev, args.clips, args.clip_len, SEED + rank, exchange=exchange)

    sampler = ClockSampler(local)
    sampler.start()
    ms = timed_region(shard, args.steps, args.warmup, world)
    note = 'timed region'
    if ms < 400.0:   # microsecond steps: keep the same load running until NVML has been polled enough (the number
        # of extra rounds comes from the max-reduced time, so every rank does the same)
        shard.run_steps((int(400.0 / max(ms / args.steps, 1e-4)) // shard.CHUNK + 1) * shard.CHUNK)
        torch.cuda.synchronize()
        note = 'timed region + the same step loop continued for ~0.4 s (region shorter than the NVML poll)'
    sampler.stop()
    ms_per_step = ms / args.steps
    poses_total = world * shard.poses
    value = poses_total * args.steps / (ms * 1e-3)

    line = {'metric': METRIC, 'value': round(value, 1), 'unit': 'poses/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': round(ms_per_step, 6), 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': config_of(args, args.gpus),
            'clocks': sampler.summary(note), 'gpu_launches': shard.launches_per_step * args.steps,
            'timing': 'CUDA events on the launching stream around K graph-replayed steps, queued behind a busy '
                      'kernel (0.2 ms + 3 us per launch: the host finishes launching first) so the events see device time only; barrier + synchronize on both sides; max over ranks'}
    if exchange is not None:
        # every rank's steps are complete (timed_region ends with synchronize + barrier): read what the peers' kernels
        # stored into THIS rank's block for the last step
        import torch.distributed as dist
        exchange.flush()              # (a step's sums leave at the start of the next step's kernel: send the last ones)
        torch.cuda.synchronize()
        dist.barrier()
        got = exchange.read().cpu().tolist()
        line['exchange'] = {'kind': 'in-kernel peer stores over NVLink (CUDA IPC mapped blocks), no collective on the step path',
                            'ranks_arrived': int(got[4]), 'world': world, 'global_loc_2d': got[5], 'global_loc_3d': got[6],
                            'local_loc_2d': float(shard.loss_block[0, 0]), 'steps_published': int(got[7])}
    else:
        line['exchange'] = {'kind': 'none (1 rank)'}
    if not args.no_e2e:
        e2e_steps = max(10, min(args.steps, 200))
        e2e_s, h2d, d2h = e2e_region(shard, e2e_steps, max(3, min(args.warmup, 10)), world, exchange)
        line['e2e'] = {'value': round(poses_total * e2e_steps / e2e_s, 1), 'unit': 'poses/s', 'h2d_bytes_per_step': h2d,
                       'd2h_bytes_per_step': d2h, 'steps': e2e_steps,
                       'api': 'r = ops.fused_step(..., out=fused_step_buffers); r.backward(); inputs from one pinned '
                              'staging image per step, copied on a prefetch stream; the 8-float loss block read back',
                       'note': 'PCIe-bound: in the real flow the 6-D tensor (2.56 of the 5.1 MB) is produced on the '
                               'device by the model and only the targets cross PCIe'}
    line['roofline'] = roofline_of('fused', shard.poses, args.clip_len, ms_per_step,
                                   'p2c_fused_step (%d launch(es)/step; duration = whole step, CUDA events)'
                                   % shard.launches_per_step)
    shard.drop_graphs()
    if not args.no_sweep:
        line['sweep'] = sweep(dev, world, rank, exchange, args)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        pps, used, iters, per = cpu_reference(args.clips, args.clip_len, 50, 1, seconds=args.cpu_seconds)
        line['cpu_baseline'] = {'value': round(pps, 1), 'unit': 'poses/s', 'cores': used, 'kind': 'port',
                                'sample': '%d full %dx%d steps of oracle.full_step + autograd backward (%.1f s)'
                                          % (iters, args.clips, args.clip_len, iters * per / pps)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier()
        if exchange is not None:
            exchange.close()
        dist.destroy_process_group()


def sweep(dev, world, rank, exchange, args):
    """The other BASELINE configs, each device-timed like the headline (graph replays, CUDA events, max over ranks):
    C5 = 8192x16 per GPU (1M poses over 8 GPUs) fwd+bwd; C4 = 4096x64 mixed skeletons, forward-only (module API entry)
    and fwd+bwd; C3 = autoencoder flow, 2-D masked MSE fwd+bwd on 256x16 IN TOTAL, batch-sharded over the ranks
    (strong scaling), plus a bandwidth-bound size.  Values are whole-job poses/s."""
    out = []
    steps, warm = 20, 4

    def entry(name, kind, make, total_poses, scaling):
        try:
            sh = make()
            ms = timed_region(sh, steps, warm, world)
            r = roofline_of(kind, sh.poses, sh.L, ms / steps, name)
            out.append({'config': name, 'clips_per_gpu': sh.N, 'clip_len': sh.L, 'n_gpus': world, 'scaling': scaling,
                        'poses_per_s': round(total_poses * steps / (ms * 1e-3), 1), 'ms_per_step': round(ms / steps, 5),
                        'roofline_frac': r['frac'], 'achieved_gbs': r['achieved'], 'bytes_per_pose': r['bytes_per_pose'],
                        'launches_per_step': sh.launches_per_step})
            sh.drop_graphs()
            del sh
            torch.cuda.empty_cache()
        except Exception as e:  # report, never hide
            out.append({'config': name, 'error': repr(e)})
    entry('C5: FK+projection+loc_2d_3d fwd+bwd, 8192x16 per GPU', 'fused',
          lambda: FusedShard(dev, 8192, 16, SEED + 1000 + rank, want_bytes=1, max_sets=2, exchange=exchange),
          world * 8192 * 16, 'weak')
    clips3 = max(1, 256 // world)
    entry('C3: autoencoder flow, Loc2DPoseLoss fwd+bwd, 256x16 in total', 'loc2d',
          lambda: Loc2dShard(dev, clips3, 16, SEED + 3000 + rank, exchange=exchange), world * clips3 * 16, 'strong')
    if world == 1:
        entry('C4: FK+projection+loc_2d_3d fwd+bwd, 4096x64', 'fused',
              lambda: FusedShard(dev, 4096, 64, SEED + 2000, want_bytes=1, max_sets=2), 4096 * 64, 'weak')
        entry('C4: FK+projection forward-only (module API entry, matrix input), 4096x64 mixed skeletons', 'forward',
              lambda: ForwardShard(dev, 4096, 64, SEED + 2500, want_bytes=1, max_sets=2), 4096 * 64, 'weak')
        entry('C3 at a bandwidth-bound size: Loc2DPoseLoss fwd+bwd, 8192x16', 'loc2d',
              lambda: Loc2dShard(dev, 8192, 16, SEED + 3500, want_bytes=1, max_sets=3), 8192 * 16, 'weak')
        try:
            out.append(separate_ops_step(dev, args))
        except Exception as e:
            out.append({'config': 'module-API step (separate ops)', 'error': repr(e)})
        if rank == 0 and not args.no_cpu_baseline:
            # C5's CPU arm: the oracle on a 1/64 slice of one GPU's shard (128 clips x 16), scaled linearly
            pps, used, iters, per = cpu_reference(128, 16, 10, 1, seconds=6.0)
            out.append({'config': 'C5 CPU reference path: 1/64 slice of a shard (128x16), %d steps' % iters,
                        'poses_per_s': round(pps, 1), 'cores': used, 'kind': 'port',
                        'seconds_for_1M_poses_scaled': round(1048576 / pps, 1)})
    return out


def separate_ops_step(dev, args, steps=20, warm=5):
    """What the UNMODIFIED reference flows call (modules/movements/movements.py:105-118 ->
    modules/layers/projection.py:73-123 -> modules/flow/pose_lifting.py:167-170 -> modules/flow/base.py:440-469):
    rotation_6d_to_matrix, ProjectionModule.forward on 3x3 matrices, the normaliser, loc_2d and loc_3d as separate
    calls, then loss.backward() through torch autograd -- every node one of this library's kernels.  256x16, eager
    launches (the module API allocates its outputs), CUDA events."""
    from pedestrians_video_2_carla_b200 import ops
    from pedestrians_video_2_carla_b200.modules.layers.projection import ProjectionModule
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.normalizer import Normalizer
    n, l = args.clips, args.clip_len
    g = torch.Generator().manual_seed(SEED + 77)
    sid, d6, d6_gt, drop = synthetic_inputs(n, l, g)
    tree, cam, loc, rot = skeleton_tables(dev)
    zeros = torch.zeros((n, l, B, 3), device=dev)
    t = ops.fused_step(d6_gt.to(dev), tree, cam, loc, rot, zeros, zeros, ref_index=sid.int().to(dev))
    gt2 = t.projection_2d_transformed.detach().clone()
    gt2[drop.to(dev)] = 0
    gt3 = t.absolute_pose_loc.detach().clone()
    pm = ProjectionModule()
    ages = [('adult', 'adult', 'child', 'child')[i] for i in sid.tolist()]
    genders = [('female', 'male', 'female', 'male')[i] for i in sid.tolist()]
    x = d6.to(dev).requires_grad_(True)
    pm.on_batch_start((x, None, {'age': ages, 'gender': genders}), 0)
    norm = Normalizer()

    def one():
        x.grad = None
        mats = ops.rotation_6d_to_matrix(x)
        proj, outs = pm(mats, None, None)
        pt = norm(proj)
        loss = ops.masked_mse(pt, gt2, 2, HIPS) + ops.masked_mse(outs['absolute_pose_loc'], gt3, 3, -1)
        loss.backward()
        return loss
    for _ in range(warm):
        one()
    torch.cuda.synchronize()

    def timed(fn):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps
    eager_ms = timed(one)
    out = {'config': 'module-API step (separate ops: rot6d -> ProjectionModule.forward -> Normalizer -> loc_2d + loc_3d '
                     '-> autograd backward), %dx%d' % (n, l), 'eager_ms_per_step': round(eager_ms, 5),
           'eager_note': 'host-bound: ~25 autograd nodes / allocator calls per step'}
    try:   # the same step as ONE CUDA graph (forward + autograd backward captured): the device time of its kernels
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                one()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        x.grad = None
        with torch.cuda.graph(graph):
            one()
        torch.cuda.synchronize()
        for _ in range(warm):
            graph.replay()
        torch.cuda.synchronize()
        ms = timed(graph.replay)
        out.update(ms_per_step=round(ms, 5), poses_per_s=round(n * l / (ms * 1e-3), 1),
                   timing='one CUDA graph per step (device time of the step\'s kernels), CUDA events')
    except Exception as e:   # noqa: BLE001
        out.update(ms_per_step=round(eager_ms, 5), poses_per_s=round(n * l / (eager_ms * 1e-3), 1),
                   timing='eager (graph capture failed: %r)' % (e,))
    return out


if __name__ == '__main__':
    main()
