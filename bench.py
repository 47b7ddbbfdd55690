#!/usr/bin/env python
"""bench.py -- poses/sec of the FK + projection + loc_2d_3d hot path, forward + backward (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm (oracle port) on host cores

A "step" is one pass of the hot path over one batch: rotation_6d model output (N,L,B,6) -> relative-rotation scan
-> FK over the CARLA_SKELETON tree -> pinhole projection -> hips/neck normalisation -> loc_2d + loc_3d + loc_2d_3d
-> gradient w.r.t. the 6-D input.  Default workload = BASELINE.json configs[1]: 256 clips x 16 frames per GPU
(weak scaling: every rank runs its own 256x16 shard; the only collective is the all-reduce of the loss sums).
One JSON line is printed by rank 0.
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=400)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--clips', type=int, default=256, help='clips per GPU (BASELINE configs[1]: 256)')
    ap.add_argument('--clip-len', type=int, default=16)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-sweep', action='store_true', help='skip the large-batch roofline sweep (C4/C5 shard sizes)')
    ap.add_argument('--cpu-seconds', type=float, default=12.0, help='budget of the bounded CPU sample')
    return ap.parse_args()


def algorithmic_bytes_per_pose(clip_len, world=False, gt2_channels=2):
    """DESIGN.md 'Algorithmic bytes': one pass, 6-D entry fused, module outputs materialised.
    read : d6 26*6*4 = 624, gt abs loc 312, gt 2-D 208 (the two channels the loss reads), world loc+rot 48 if given
    write: rel_rot 936 + abs_rot 936 + abs_loc 312 + projection 312 = 2496, grad wrt d6 624
    (+ the per-clip skeleton row 1248/L is a 4-row table shared by all clips: amortised to ~0)"""
    return 624 + 312 + 4 * B * gt2_channels + (48 if world else 0) + 2496 + 624


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
class Shard(object):
    """One rank's shard of the workload: `n_sets` rotating buffer sets (their footprint together exceeds L2, so every
    timed step streams its inputs from HBM), each with a prebuilt C-ABI argument struct and a captured CUDA graph."""

    def __init__(self, dev, n_clips, clip_len, seed, want_bytes=192 << 20, max_sets=16, capture=True, world=1):
        import oracle as O  # checker-side data only: reference skeleton table + camera constants (init, untimed)
        from pedestrians_video_2_carla_b200 import _lib, ops
        self.lib, self.cdll, self.ops, self.dev = _lib, _lib.load(), ops, dev
        self.N, self.L = n_clips, clip_len
        self.poses = n_clips * clip_len
        loc, rot, _, par = O.reference_tensors(torch.float32)
        cam = O.training_camera()
        self.tree = _lib.make_tree(par)
        self.cam = _lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py)
        self.loc, self.rot = loc.to(dev), rot.to(dev)
        self.n_sets = n_sets_for(self.poses, clip_len, want_bytes, max_sets)
        self.loss_block = torch.zeros((self.n_sets, 8), device=dev)    # row i = the losses[8] of buffer set i
        self.packed = torch.zeros((self.n_sets, 4), device=dev)        # the four sums of a round's steps, to be reduced
        self.global_sums = torch.zeros((self.n_sets, 4), device=dev)   # ... and reduced over the ranks
        g = torch.Generator(device='cpu').manual_seed(seed)
        base = torch.tensor([1., 0, 0, 0, 1, 0])
        self.sets = []
        for _ in range(self.n_sets):
            s = {}
            s['sid'] = torch.randint(0, 4, (n_clips,), generator=g).int().to(dev)
            s['d6'] = (torch.randn((n_clips, clip_len, B, 6), generator=g) * 0.3 + base).to(dev)
            d6_gt = (torch.randn((n_clips, clip_len, B, 6), generator=g) * 0.3 + base).to(dev)
            zeros = torch.zeros((n_clips, clip_len, B, 3), device=dev)
            t = ops.fused_step(d6_gt, self.tree, self.cam, self.loc, self.rot, zeros, zeros, ref_index=s['sid'])
            gt2 = t.projection_2d_transformed.clone()
            drop = (torch.rand((n_clips, clip_len, B), generator=g) < 0.1).to(dev)
            drop[..., 1] = False
            gt2[drop] = 0
            s['gt2'], s['gt3'] = gt2, t.absolute_pose_loc.clone()
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
        self.graphs = None
        self.world = world
        self.allreduce = 'none (1 rank)' if world == 1 else 'eager, after every step graph'
        if capture:
            self._capture()

    def _args(self, s):
        L = self.lib
        a = L.FusedStep()
        a.tree, a.hips, a.neck = self.tree, 1, 8
        a.n_clips, a.clip_len, a.mode, a.input_6d, a.normalize, a.mask_missing_joints = self.N, self.L, 0, 1, 1, 1
        a.pose_in, a.ref_loc, a.ref_rot, a.ref_index = L.ptr(s['d6']), L.ptr(self.loc), L.ptr(self.rot), L.ptr(s['sid'])
        a.world_loc, a.world_rot, a.camera = None, None, self.cam       # ZeroTrajectory: identity world transform
        a.gt_2d, a.gt_2d_channels, a.gt_3d = L.ptr(s['gt2']), 3, L.ptr(s['gt3'])
        a.weight_2d, a.weight_3d = 1.0, 1.0
        a.rel_rot, a.abs_rot, a.abs_loc, a.proj, a.proj_norm = (L.ptr(s['rel_rot']), L.ptr(s['abs_rot']),
                                                                L.ptr(s['abs_loc']), L.ptr(s['proj']), None)
        a.g_pose_in, a.losses = L.ptr(s['g_d6']), L.ptr(s['losses'])
        return a

    def launch(self, i):
        """Enqueue step i on the current stream (direct C-ABI call; no torch ops)."""
        self.lib.check(self.cdll.p2c_fused_step(self.sets[i % self.n_sets]['args'], self.lib.stream()), 'fused_step')

    def _reduce(self, i):
        """The path's only exchange: the four loss sums of step i, summed over the ranks (NCCL)."""
        import torch.distributed as dist
        return dist.all_reduce(self.sets[i % self.n_sets]['losses'][3:7], async_op=True)

    def _capture(self):
        st = torch.cuda.Stream(device=self.dev)
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            for i in range(self.n_sets):
                self.launch(i)
                if self.world > 1:
                    self._reduce(i).wait()
        torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        self.graphs = []
        for i in range(self.n_sets):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                self.launch(i)
            self.graphs.append(gr)
        self.round = torch.cuda.CUDAGraph()   # one replay = n_sets consecutive steps (one per buffer set)
        if self.world == 1:
            with torch.cuda.graph(self.round):
                for i in range(self.n_sets):
                    self.launch(i)
        else:
            # N > 1: the exchange is captured INTO the round graph.  The four loss sums of the round's steps are packed
            # into one (n_sets, 4) block and summed over the ranks by ONE NCCL call, forked onto the process group's
            # stream at the top of the NEXT replay, where it runs beside that round's kernels (a few hundred bytes over
            # NVLink: pure latency, ~25 us at 8 ranks) and joins before the replay ends; flush_exchange() reduces the
            # last round.  Every step's sums are reduced, one round late; gradients are local (DDP semantics).  If
            # this torch / NCCL pair cannot capture collectives, the steps fall back to single-step graphs followed
            # by an eager all-reduce.
            import torch.distributed as dist
            try:
                with torch.cuda.graph(self.round):
                    w = dist.all_reduce(self.packed, async_op=True)
                    for i in range(self.n_sets):
                        self.launch(i)
                    w.wait()
                    self.global_sums.copy_(self.packed)
                    self.packed.copy_(self.loss_block[:, 3:7])
                self.allreduce = ('captured in the round graph: one NCCL all-reduce per %d steps (their 4 loss sums '
                                  'each), beside the next round\'s kernels' % self.n_sets)
            except Exception as e:   # noqa: BLE001 - any capture failure takes the eager route
                sys.stderr.write('bench: collective capture failed (%s); eager all-reduce per step\n' % (e,))
                self.round = None
        torch.cuda.synchronize()

    def flush_exchange(self):
        """Reduce the sums of the last replayed round (the round graph reduces those of the round before)."""
        if self.world > 1 and self.round is not None:
            import torch.distributed as dist
            dist.all_reduce(self.packed)
            self.global_sums.copy_(self.packed)
            self.packed.zero_()

    def run_steps(self, first, count):
        """Steps first .. first+count-1, as whole-round graph replays where aligned, single-step graphs otherwise."""
        i, end = first, first + count
        while i < end:
            if self.round is not None and i % self.n_sets == 0 and end - i >= self.n_sets:
                self.round.replay()
                i += self.n_sets
            else:
                self.graphs[i % self.n_sets].replay()
                if self.world > 1:
                    self._reduce(i).wait()
                i += 1


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


def timed_region(shard, steps, warmup, world, sampler_note=None):
    """W untimed steps, then exactly K steps between CUDA events on the launching stream, barrier + synchronize on
    both sides, max over ranks.  For N > 1 every step ends with the NCCL all-reduce of the four loss sums."""
    import torch.distributed as dist
    dev = shard.dev

    def step_range(first, count):
        shard.run_steps(first, count)
    step_range(0, warmup)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    first = -(-warmup // shard.n_sets) * shard.n_sets   # start the timed steps on a round boundary
    e0.record()
    step_range(first, steps)
    shard.flush_exchange()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def e2e_region(shard, steps, warmup, world):
    """The same metric through the public API (ops.fused_step + backward) with HOST inputs: every step copies that
    step's d6 / targets / skeleton ids from pinned host memory, runs forward + backward, and reads the loss back."""
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
    loss_host = torch.zeros(1).pin_memory()
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
                           want_transformed=False, out=pre)
        r.backward()            # = r.loc_2d_3d.backward(): the gradient is already in the step's output block
        consumed[k].record()
        if world > 1:
            dist.all_reduce(r.losses[3:7])
        loss_host.copy_(r.loc_2d_3d.detach().reshape(1), non_blocking=False)   # device -> host read of the result
        dbuf['d6'] = x.detach()
        return float(loss_host[0])
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
    return float(dt.item()), h2d, 4


# ------------------------------------------------------------------------------------------ CPU arm (oracle port)
def cpu_reference(n_clips, clip_len, seconds, max_iters=50, threads=None):
    """The reference's algorithm (oracle/: per-bone recursion, per-frame Python loops, torch CPU ops, autograd
    backward) on this box's host cores.  Returns (poses_per_sec, threads, iterations, poses per iteration)."""
    import oracle as O
    if threads:
        torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(SEED)
    loc, rot, _, par = O.reference_tensors(torch.float32)
    cam = O.training_camera()
    sid = torch.randint(0, 4, (n_clips,), generator=g)
    base = torch.tensor([1., 0, 0, 0, 1, 0])
    d6 = (torch.randn((n_clips, clip_len, B, 6), generator=g) * 0.3 + base).requires_grad_(True)
    with torch.no_grad():
        t = O.full_step(O.rotation_6d_to_matrix(torch.randn((n_clips, clip_len, B, 6), generator=g) * 0.3 + base),
                        loc[sid], rot[sid], par, cam, torch.zeros(n_clips, clip_len, B, 3), torch.zeros(n_clips, clip_len, B, 3))
    gt2, gt3 = t['projection_2d_transformed'].clone(), t['absolute_pose_loc'].clone()
    gt2[torch.rand((n_clips, clip_len, B), generator=g) < 0.1] = 0
    ref_loc, ref_rot = loc[sid], rot[sid]

    def one():
        d6.grad = None
        out = O.full_step(O.rotation_6d_to_matrix(d6), ref_loc, ref_rot, par, cam, gt2, gt3)
        out['loc_2d_3d'].backward()
        return float(out['loc_2d_3d'].detach())
    one()  # warm-up
    times = []
    t_end = time.perf_counter() + seconds
    while len(times) < max_iters and (time.perf_counter() < t_end or len(times) < 3):
        t0 = time.perf_counter()
        one()
        times.append(time.perf_counter() - t0)
    return n_clips * clip_len * len(times) / sum(times), torch.get_num_threads(), len(times), n_clips * clip_len


def n_sets_for(poses, clip_len, want_bytes=192 << 20, max_sets=16):
    per_set = poses * (algorithmic_bytes_per_pose(clip_len) + 4 * B * 1)  # gt_2d is stored with 3 channels
    return int(max(2, min(max_sets, -(-want_bytes // max(per_set, 1)))))


def config_of(args, world):
    poses = args.clips * args.clip_len
    n_sets = n_sets_for(poses, args.clip_len)
    l2 = '%d rotating buffer sets per GPU (%.0f MB > 126 MB L2), no flush kernel' % (
        n_sets, n_sets * poses * (algorithmic_bytes_per_pose(args.clip_len) + 104) / 1e6)
    return {'workload': 'pose_lifting step hot path (6-D output -> scan -> FK -> projection -> hips/neck '
                        'normalise -> loc_2d_3d, fwd+bwd), %dx%d CARLA_SKELETON poses per GPU'
                        % (args.clips, args.clip_len),
            'clips_per_gpu': args.clips, 'clip_len': args.clip_len, 'bones': B, 'mode': 'pose_changes',
            'mask_missing_joints': True, 'transform': 'hips_neck', 'l2': l2,
            'parallelism': 'clips sharded by batch, %d rank(s); all-reduce of 4 loss sums' % world}


def measured_peak():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'MEASURED_PEAKS.json hbm_gbs (of measured)'
    except Exception:
        return 6650.0, 'B200_PROFILING.md fallback 6.65 TB/s (of fallback)'


# dram__bytes_read.sum + dram__bytes_write.sum of the step kernel, one launch, from the committed `ncu --set full`
# captures (profiles/r01_tile_<clips>x<len>.txt); writes still in L2 when the kernel ends are not in it
NCU_DRAM_BYTES = {(256, 16): 5.184e6, (8192, 16): 164.11e6 + 353.60e6, (4096, 64): 327.81e6 + 759.60e6}


def roofline_of(poses, clip_len, ms_per_step, note):
    peak, src = measured_peak()
    bpp = algorithmic_bytes_per_pose(clip_len)
    achieved = poses * bpp / (ms_per_step * 1e-3) / 1e9
    return {'bound': 'hbm', 'achieved': round(achieved, 1), 'peak': peak, 'unit': 'GB/s',
            'frac': round(achieved / peak, 4), 'traffic': NCU_DRAM_BYTES.get((poses // clip_len, clip_len)),
            'traffic_unit': 'bytes per launch (ncu, profiles/r01_tile_*)', 'bytes_per_pose': bpp, 'peak_source': src,
            'kernel': note}


def main_reference(args, world, rank):
    """--impl reference: the reference's CPU algorithm (oracle port; pytorch3d / pedestrians_scenarios are not
    installable here) on all host threads, same workload.  Rank 0 only."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    budget = max(5.0, min(120.0, 0.25 * (args.steps + args.warmup)))
    pps, used, iters, per = cpu_reference(args.clips, args.clip_len, budget, max_iters=max(3, args.steps))
    line = {'impl': 'reference', 'metric': METRIC, 'value': round(pps, 1), 'unit': 'poses/s', 'n_gpus': args.gpus,
            'steps': iters, 'warmup': 1, 'ms_per_step': round(per / pps * 1e3, 3), 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': config_of(args, args.gpus),
            'cpu_baseline': {'value': round(pps, 1), 'unit': 'poses/s', 'cores': used, 'kind': 'port',
                             'sample': '%d full %dx%d steps of oracle.full_step + autograd backward'
                                       % (iters, args.clips, args.clip_len)},
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
    shard = Shard(dev, args.clips, args.clip_len, SEED + rank, world=world)

    sampler = ClockSampler(local)
    sampler.start()
    ms = timed_region(shard, args.steps, args.warmup, world)
    note = 'timed region'
    if ms < 400.0:   # microsecond steps: keep the same load running until NVML has been polled enough.  The decision
        # and the number of extra rounds come from the max-reduced time, so every rank replays the same number of
        # graphs (they hold collectives when N > 1)
        rounds = int(400.0 / max(ms / args.steps * shard.n_sets, 1e-3)) + 1
        for r in range(rounds):
            shard.run_steps(r * shard.n_sets, shard.n_sets)
        torch.cuda.synchronize()
        note = 'timed region + the same step loop continued for ~0.4 s (region shorter than the NVML poll)'
    sampler.stop()
    ms_per_step = ms / args.steps
    poses_total = world * shard.poses
    value = poses_total * args.steps / (ms * 1e-3)

    e2e_steps = max(10, min(args.steps, 200))
    e2e_s, h2d, d2h = e2e_region(shard, e2e_steps, max(3, min(args.warmup, 10)), world)
    e2e_value = poses_total * e2e_steps / e2e_s

    line = {'metric': METRIC, 'value': round(value, 1), 'unit': 'poses/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': round(ms_per_step, 6), 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': dict(config_of(args, world), allreduce=shard.allreduce),
            'clocks': sampler.summary(note),
            'e2e': {'value': round(e2e_value, 1), 'unit': 'poses/s', 'h2d_bytes_per_step': h2d,
                    'd2h_bytes_per_step': d2h, 'steps': e2e_steps,
                    'api': 'r = ops.fused_step(..., out=fused_step_buffers); r.backward(); inputs from one pinned staging image '
                           'per step, copied on a prefetch stream'},
            'gpu_launches': shard.launches_per_step * args.steps}
    line['roofline'] = roofline_of(shard.poses, args.clip_len, ms_per_step,
                                   'p2c_fused_step (%d launch(es)/step; duration = whole step, CUDA events)'
                                   % shard.launches_per_step)
    if rank == 0 and world == 1 and not args.no_sweep:
        line['sweep'] = sweep(dev)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        pps, used, iters, per = cpu_reference(args.clips, args.clip_len, args.cpu_seconds)
        line['cpu_baseline'] = {'value': round(pps, 1), 'unit': 'poses/s', 'cores': used, 'kind': 'port',
                                'sample': '%d full %dx%d steps of oracle.full_step + autograd backward (%.1f s)'
                                          % (iters, args.clips, args.clip_len, iters * per / pps)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        # graphs that hold captured NCCL kernels go first; then leave without running the process group's
        # destructors (a communicator torn down at interpreter exit, with graph-captured collectives behind it, can
        # wait forever on this torch / NCCL pair)
        import torch.distributed as dist
        shard.round = None
        shard.graphs = None
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def sweep(dev):
    """Large-batch points (per-GPU shard sizes of BASELINE configs C4 / C5) where the path is bandwidth- rather than
    latency-bound: device-timed poses/s and fraction of the measured HBM peak.  Not the headline."""
    out = []
    for n, l in ((8192, 16), (4096, 64)):
        try:
            sh = Shard(dev, n, l, SEED + 1000 + n, want_bytes=1, max_sets=2)
            ms = timed_region(sh, 20, 4, 1)
            r = roofline_of(sh.poses, l, ms / 20, 'p2c_fused_step')
            out.append({'clips': n, 'clip_len': l, 'poses_per_s': round(sh.poses * 20 / (ms * 1e-3), 1),
                        'ms_per_step': round(ms / 20, 4), 'roofline_frac': r['frac'], 'achieved_gbs': r['achieved']})
            del sh
            torch.cuda.empty_cache()
        except Exception as e:  # report, never hide
            out.append({'clips': n, 'clip_len': l, 'error': repr(e)})
    return out


if __name__ == '__main__':
    main()
