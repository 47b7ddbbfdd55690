"""Launches each SURVEY 8f kernel a few times at 8192 x 16 (for `ncu --set full -k regex:...`; no timing here).
Run on the GPU box: python tools/widen_once.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pedestrians_video_2_carla_b200 import _lib  # noqa: E402
from pedestrians_video_2_carla_b200._lib import ptr, stream  # noqa: E402

dev = torch.device('cuda', 0)
N, L, B = 8192, 16, 26
lib = _lib.load()
poses = N * L
c, o, g = (torch.randn((N, L, B, 3, 3), device=dev) for _ in range(3))
x, y, gy = (torch.randn((N, L, B, 3), device=dev) for _ in range(3))
sh, sc = torch.randn((4, 3), device=dev), torch.rand(4, device=dev) + 0.5
idx = torch.randint(0, 4, (N,), device=dev, dtype=torch.int32)
ws = torch.empty(lib.p2c_metric_workspace_floats(), device=dev)
st = torch.zeros(2, dtype=torch.float64, device=dev)
for _ in range(3):
    lib.p2c_cum_changes_fwd(ptr(c), ptr(o), N, L, B, stream())
    lib.p2c_cum_changes_bwd(ptr(c), ptr(o), ptr(g), ptr(c), N, L, B, stream())
    lib.p2c_renormalize_fwd(ptr(x), ptr(y), poses, B, 3, 3, 1, 8, 1, 1e-5, ptr(sh), ptr(sc), ptr(idx), L, stream())
    lib.p2c_renormalize_bwd(ptr(x), ptr(gy), ptr(y), poses, B, 3, 3, 1, 8, 1, 1e-5, ptr(sc), ptr(idx), L, stream())
    lib.p2c_mpjpe_sum(ptr(x), ptr(y), poses, B, 3, ptr(ws), ptr(st), 1, stream())
    lib.p2c_pck_count(ptr(x), ptr(y), ptr(y), poses, B, 3, 3, 1, 1, 8, 1, 0.05, 1e-5, ptr(ws), ptr(st), 1, stream())
torch.cuda.synchronize()
print('ok')
