"""CPU-side cost of each segment of the end-to-end step (no copies): where do the ~180 us around a 14 us kernel go?
Run on the GPU box: python tools/e2e_segments.py"""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

dev = torch.device('cuda', 0)
sh = bench.FusedShard(dev, 256, 16, 1, want_bytes=1, max_sets=2)
ops = sh.ops
s = sh.sets[0]
v = {k: s[k].clone() for k in ('d6', 'gt2', 'gt3', 'sid')}
loss_host = torch.zeros(1).pin_memory()
N, L, B = v['d6'].shape[:3]
pre = dict(rel_rot=torch.empty((N, L, B, 3, 3), device=dev), abs_rot=torch.empty((N, L, B, 3, 3), device=dev),
           abs_loc=torch.empty((N, L, B, 3), device=dev), proj=torch.empty((N, L, B, 3), device=dev),
           g_pose_in=torch.empty_like(v['d6']), losses=torch.empty(8, device=dev))
acc = {}


def seg(name, t0):
    t1 = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t1 - t0)
    return t1


def one(out=None, sync=True):
    t = time.perf_counter()
    x = v['d6'].requires_grad_(True)
    t = seg('requires_grad', t)
    r = ops.fused_step(x, sh.tree, sh.cam, sh.loc, sh.rot, v['gt2'], v['gt3'], ref_index=v['sid'],
                       want_transformed=False, out=out)
    t = seg('fused_step', t)
    r.loc_2d_3d.backward()
    t = seg('backward', t)
    v['d6'] = x.detach()
    if sync:
        loss_host.copy_(r.loc_2d_3d.detach().reshape(1))
        t = seg('loss read (sync)', t)
        return float(loss_host[0])


for label, out in (('fresh outputs', None), ('preallocated outputs', pre)):
    for _ in range(30):
        one(out)
    acc.clear()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(300):
        one(out)
    torch.cuda.synchronize()
    tot = (time.perf_counter() - t0) / 300 * 1e6
    print('%s: %.1f us/step; ' % (label, tot) + ', '.join('%s %.1f' % (k, a / 300 * 1e6) for k, a in acc.items()))

pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    one(None)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('tottime').print_stats(18)
