"""Where the end-to-end step (host inputs -> fused step + backward -> loss on the host) spends its time, and which
host-side arrangement is fastest: four copies vs one packed staging buffer, serial vs prefetched on a copy stream.
Run on the GPU box: python tools/e2e_variants.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

dev = torch.device('cuda', 0)
sh = bench.FusedShard(dev, 256, 16, 1, want_bytes=1, max_sets=2)
ops = sh.ops
s = sh.sets[0]
keys = ('d6', 'gt2', 'gt3', 'sid')
host = {k: s[k].cpu().pin_memory() for k in keys}
loss_host = torch.zeros(1).pin_memory()


def packed(n):
    """n device sets + one pinned host image: every tensor is a 16-byte aligned view of ONE byte buffer."""
    offs, total = {}, 0
    for k in keys:
        offs[k] = total
        total += (s[k].numel() * s[k].element_size() + 255) // 256 * 256
    hbuf = torch.empty(total, dtype=torch.uint8).pin_memory()
    for k in keys:
        nb = s[k].numel() * s[k].element_size()
        hbuf[offs[k]:offs[k] + nb].view(s[k].dtype).view(s[k].shape).copy_(host[k])
    sets = []
    for _ in range(n):
        dbuf = torch.empty(total, dtype=torch.uint8, device=dev)
        views = {k: dbuf[offs[k]:offs[k] + s[k].numel() * s[k].element_size()].view(s[k].dtype).view(s[k].shape)
                 for k in keys}
        sets.append((dbuf, views))
    return hbuf, sets


def timeit(f, n=300):
    for i in range(20):
        f(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        f(i)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


def step(v):
    x = v['d6'].requires_grad_(True)
    r = ops.fused_step(x, sh.tree, sh.cam, sh.loc, sh.rot, v['gt2'], v['gt3'], ref_index=v['sid'],
                       want_transformed=False)
    r.loc_2d_3d.backward()
    v['d6'] = x.detach()
    return r


def read(r):
    loss_host.copy_(r.loc_2d_3d.detach().reshape(1))
    return float(loss_host[0])


dsep = [{k: torch.empty_like(s[k]) for k in keys} for _ in range(2)]
hbuf, psets = packed(2)


def copies4(i):
    for k in keys:
        dsep[0][k].copy_(host[k], non_blocking=True)


def copies1(i):
    psets[0][0].copy_(hbuf, non_blocking=True)


def serial4(i):
    copies4(i)
    return read(step(dsep[0]))


def serial1(i):
    copies1(i)
    return read(step(psets[0][1]))


cs = torch.cuda.Stream(dev)
ready = [torch.cuda.Event() for _ in range(2)]
done = [torch.cuda.Event() for _ in range(2)]
for e in done:
    e.record()
for e in ready:
    e.record()


def prefetch1(i):
    b = i % 2
    cs.wait_event(done[b])
    with torch.cuda.stream(cs):
        psets[b][0].copy_(hbuf, non_blocking=True)
    ready[b].record(cs)


def pipelined1(i):
    b = i % 2
    prefetch1(i + 1)
    torch.cuda.current_stream().wait_event(ready[b])
    r = step(psets[b][1])
    done[b].record()
    return read(r)


def step_only(i):
    return read(step(dsep[0]))


print('copies only, 4 tensors   %.1f us' % timeit(copies4))
print('copies only, 1 packed    %.1f us' % timeit(copies1))
print('step + loss read         %.1f us' % timeit(step_only))
print('serial, 4 copies         %.1f us' % timeit(serial4))
print('serial, 1 packed copy    %.1f us' % timeit(serial1))
prefetch1(0)
print('prefetched, packed       %.1f us' % timeit(pipelined1))
