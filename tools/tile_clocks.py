"""Phase timeline of the tile kernel (CTA 0, second tile): P2C_TILE_CLOCKS=1 python tools/tile_clocks.py N L"""
import ctypes, os, sys
os.environ['P2C_TILE_CLOCKS'] = '1'
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from pedestrians_video_2_carla_b200 import _lib
n, l = int(sys.argv[1]), int(sys.argv[2])
sh = bench.Shard(torch.device('cuda', 0), n, l, 1, want_bytes=1, max_sets=2, capture=False)
for i in range(3):
    sh.launch(i)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 64)()
lib = _lib.load()
lib.p2c_debug_tile_clocks.argtypes = [ctypes.c_void_p]
assert lib.p2c_debug_tile_clocks(buf) == 0
names = ['top', 'd6 landed', 'bar1', 'A done(bar2)', 'B done(bar3)', 'FK done(bar4)', 'targets landed', 'loss done(bar5)',
         'torque done(bar6)', 'D done', 'd6 again landed', 'bar7', 'E done(bar8)']
v = list(buf)[:13]
for i in range(1, 13):
    print('%-20s +%6d cycles' % (names[i], v[i] - v[i - 1]))
print('tile total %d cycles' % (v[12] - v[0]))

w = list(buf)
print('FK levels (warp 0): task, prefetch, barrier wait')
prev = v[4]
for lv in range(8):
    a, b, c = w[16 + 3 * lv], w[17 + 3 * lv], w[18 + 3 * lv]
    print('  level %d: task %5d  prefetch %5d  barrier %5d' % (lv, a - prev, b - a, c - b))
    prev = c
