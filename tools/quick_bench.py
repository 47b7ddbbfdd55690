"""Device time of the fused step at the three benchmarked sizes for the library named by P2C_LIB (A/B runs on ONE box).
  P2C_LIB=... python tools/quick_bench.py [label]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

dev = torch.device('cuda', 0)
torch.cuda.set_device(dev)
out = []
for n, l, steps in ((256, 16, 240), (8192, 16, 20), (4096, 64, 20)):
    sh = bench.FusedShard(dev, n, l, bench.SEED, **({} if n == 256 else dict(want_bytes=1, max_sets=2)))
    best = min(bench.timed_region(sh, steps, 10, 1) / steps for _ in range(3))
    out.append('%dx%d %.2f us' % (n, l, best * 1e3))
    sh.drop_graphs()
    del sh
    torch.cuda.empty_cache()
print('%-8s %s' % (sys.argv[1] if len(sys.argv) > 1 else os.path.basename(os.environ.get('P2C_LIB', 'default')), ' | '.join(out)))
