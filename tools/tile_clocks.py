"""Phase timeline of the tile kernel (CTA 0, second tile): P2C_TILE_CLOCKS=1 python tools/tile_clocks.py N L"""
import ctypes, os, sys
os.environ.setdefault('P2C_TILE_CLOCKS', '2')
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
names = ['top', 'd6 landed', 'bar1', 'A done(bar2)', 'B done(bar3)', 'FK done(bar4)', 'shift/scale + targets(bar4b)',
         'loss done(bar5)', 'norm bwd + subtree sums(bar6)', 'y done(bar7)', 'E done(bar8)']
v = list(buf)
v = v[:10] + [v[12]]
for i in range(1, 11):
    print('%-30s +%6d cycles' % (names[i], v[i] - v[i - 1]))
print('tile total %d cycles' % (v[10] - v[0]))
w = list(buf)
if w[16]:
    print('CTA 0: entry -> tile loop %d, tile loop %d, store drain %d, loss fold + ticket %d cycles' % (
        w[17] - w[16], w[18] - w[17], w[19] - w[18], w[20] - w[19]))
if w[21]:
    print('forward walk done %d, count poll done %d cycles after barrier 3' % (w[21] - w[4], w[22] - w[4]))
