"""Device-timed throughput of the SURVEY 8f kernels (csrc/widen_kernels.cu, metrics_kernels.cu) against the HBM peak:
algorithmic bytes / CUDA-event time over rotating buffer sets larger than L2.  Run on the GPU box:
python tools/widen_bench.py [n_clips clip_len]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pedestrians_video_2_carla_b200 import _lib, ops  # noqa: E402
from pedestrians_video_2_carla_b200._lib import ptr, stream  # noqa: E402

dev = torch.device('cuda', 0)
N, L = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (8192, 16)
B = 26
peak = 6548.8
try:
    peak = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs'])
except Exception:
    pass
lib = _lib.load()
poses = N * L
sets = max(2, min(12, int((200 << 20) / (poses * 936 * 2)) + 1))   # > 126 MB L2 over a round of the sets


def bufs(shape, dtype=torch.float32):
    return [torch.randn(shape, device=dev).to(dtype) if dtype.is_floating_point else
            torch.randint(0, 4, shape, device=dev, dtype=dtype) for _ in range(sets)]


def timed(fn, nbytes, name, reps=5, rounds=4):
    """`rounds` passes over the buffer sets captured into one CUDA graph (no per-launch host time in the number),
    replayed after a warm-up long enough for the clocks to settle; best of `reps` replays."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(sets):
            fn(i)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(rounds):
            for i in range(sets):
                fn(i)
    for _ in range(20):
        graph.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        graph.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / (sets * rounds))
    gbs = nbytes / best / 1e6
    print('%-34s %9.1f us  %8.1f GB/s  %5.1f %% of %.0f GB/s' % (name, best * 1e3, gbs, 100 * gbs / peak, peak))


print('%d clips x %d frames x %d bones, %d rotating sets' % (N, L, B, sets))
c, o, g = bufs((N, L, B, 3, 3)), bufs((N, L, B, 3, 3)), bufs((N, L, B, 3, 3))
timed(lambda i: lib.p2c_cum_changes_fwd(ptr(c[i]), ptr(o[i]), N, L, B, stream()), poses * 936 * 2, 'cum_changes_fwd')
timed(lambda i: lib.p2c_cum_changes_bwd(ptr(c[i]), ptr(o[i]), ptr(g[i]), ptr(o[(i + 1) % sets]), N, L, B, stream()),
      poses * 936 * 4, 'cum_changes_bwd')
x, y, gy = bufs((N, L, B, 3)), bufs((N, L, B, 3)), bufs((N, L, B, 3))
sh, sc = torch.randn((4, 3), device=dev), torch.rand(4, device=dev) + 0.5
idx = torch.randint(0, 4, (N,), device=dev, dtype=torch.int32)
timed(lambda i: lib.p2c_renormalize_fwd(ptr(x[i]), ptr(y[i]), poses, B, 3, 3, 1, 8, 1, 1e-5, ptr(sh), ptr(sc), ptr(idx),
                                        L, stream()), poses * 312 * 2, 'renormalize_fwd (from_abs)')
timed(lambda i: lib.p2c_renormalize_bwd(ptr(x[i]), ptr(gy[i]), ptr(y[i]), poses, B, 3, 3, 1, 8, 1, 1e-5, ptr(sc),
                                        ptr(idx), L, stream()), poses * 312 * 3, 'renormalize_bwd (from_abs)')
q, m = bufs((N * L * B, 4)), bufs((N * L * B, 3, 3))
timed(lambda i: lib.p2c_quaternion_to_matrix_fwd(ptr(q[i]), ptr(m[i]), N * L * B, stream()), poses * B * 52,
      'quaternion_to_matrix_fwd')
timed(lambda i: lib.p2c_quaternion_to_matrix_bwd(ptr(q[i]), ptr(m[i]), ptr(q[(i + 1) % sets]), N * L * B, stream()),
      poses * B * 68, 'quaternion_to_matrix_bwd')
ws = torch.empty(lib.p2c_metric_workspace_floats(), device=dev)
st = torch.zeros(2, dtype=torch.float64, device=dev)
timed(lambda i: lib.p2c_mpjpe_sum(ptr(x[i]), ptr(y[i]), poses, B, 3, ptr(ws), ptr(st), 1, stream()), poses * 312 * 2,
      'mpjpe_sum')
timed(lambda i: lib.p2c_pck_count(ptr(x[i]), ptr(y[i]), ptr(y[i]), poses, B, 3, 3, 1, 1, 8, 1, 0.05, 1e-5, ptr(ws),
                                  ptr(st), 1, stream()), poses * 312 * 2, 'pck_count (bbox, masked)')
ws2 = torch.empty(lib.p2c_mse_workspace_floats(poses), device=dev)
o3 = torch.empty(3, device=dev)
timed(lambda i: lib.p2c_masked_mse_fwd(ptr(c[i]), ptr(o[i]), poses, B, 9, 9, 9, -1, ptr(ws2), ptr(o3), stream()),
      poses * 936 * 2, 'masked_mse_fwd 9ch (rot_3d)')
