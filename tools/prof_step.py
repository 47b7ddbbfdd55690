"""Runs the fused step a few times on one buffer set (for ncu): python tools/prof_step.py N L [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

n, l = int(sys.argv[1]), int(sys.argv[2])
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = torch.device('cuda', 0)
sh = bench.FusedShard(dev, n, l, 1, want_bytes=1, max_sets=2)
for k in os.environ.get('P2C_PROF_DROP', '').split(','):   # experiment: drop outputs (rel_rot,abs_rot,abs_loc,proj)
    if k:
        for st in sh.sets:
            setattr(st['args'], k, None)
for i in range(iters):
    sh.launch(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(iters):
    sh.launch(i)
e1.record()
torch.cuda.synchronize()
print('%dx%d: %.2f us/step, %.1f M poses/s' % (n, l, e0.elapsed_time(e1) / iters * 1e3, n * l * iters / e0.elapsed_time(e1) / 1e3))
