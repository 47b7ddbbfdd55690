"""A few launches of the module-API kernels (for ncu): python tools/prof_module.py fwd|bwd|loc2d N L [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from pedestrians_video_2_carla_b200 import _lib, ops  # noqa: E402

kind, n, l = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
dev = torch.device('cuda', 0)
torch.cuda.set_device(dev)
if kind == 'loc2d':
    sh = bench.Loc2dShard(dev, n, l, 1, want_bytes=1, max_sets=2)
    run = lambda i: sh.launch(i)   # noqa: E731
elif kind == 'fwd':
    sh = bench.ForwardShard(dev, n, l, 1, want_bytes=1, max_sets=2)
    run = lambda i: sh.launch(i)   # noqa: E731
else:
    sh = bench.ForwardShard(dev, n, l, 1, want_bytes=1, max_sets=2)
    s = sh.sets[0]
    x = s['mats'].clone().requires_grad_(True)
    rr, ar, al, pj = ops.fk_clip(x, sh.loc, sh.rot, s['sid'], s['wl'], s['wr'], sh.tree, sh.cam, _lib.MODE_POSE_CHANGES, True)
    loss = (al ** 2).sum() + (pj[..., :2] ** 2).sum() * 1e-6

    def run(i):
        x.grad = None
        loss.backward(retain_graph=True)
for i in range(iters):
    run(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(iters):
    run(i)
e1.record()
torch.cuda.synchronize()
print('%s %dx%d: %.2f us per call' % (kind, n, l, e0.elapsed_time(e1) / iters * 1e3))
