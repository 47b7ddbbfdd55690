"""Phase timeline of the tile kernel (CTA P2C_TILE_CTA, n-th tile): needs the debug build of the library,
  P2C_BUILD_CLOCKS=1 python -c "import __graft_entry__ as g; g.build()"   (touch csrc/fused_tile.cu first)
  P2C_TILE_CLOCKS=1 python tools/tile_clocks.py N L"""
import ctypes, os, sys
os.environ.setdefault('P2C_TILE_CLOCKS', '2')
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from pedestrians_video_2_carla_b200 import _lib
n, l = int(sys.argv[1]), int(sys.argv[2])
sh = bench.FusedShard(torch.device('cuda', 0), n, l, 1, want_bytes=1, max_sets=2)
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
if buf[10] and buf[11]:
    print('  inside norm bwd + subtree sums: sums over bones(bar5a) +%d, own parts(bar5b) +%d, subtree walks +%d' % (buf[10] - buf[7], buf[11] - buf[10], buf[8] - buf[11]))
w = list(buf)
if w[16]:
    print('CTA 0: entry -> tile loop %d, tile loop %d, store drain %d, loss fold + ticket %d cycles' % (
        w[17] - w[16], w[18] - w[17], w[19] - w[18], w[20] - w[19]))
if w[23]:
    print('entry: top -> dependency wait over %d, -> tile loop %d, -> first 6-D records landed %d cycles' % (w[23] - w[16], w[17] - w[23], v[1] - w[17]))
if w[21]:
    print('forward walk done %d, count poll done %d cycles after barrier 3' % (w[21] - w[4], w[22] - w[4]))
if w[32]:
    print('trunk walk, cycles after barrier 3 per bone:', [int(w[32 + b] - w[4]) for b in range(8)])
if w[40]:
    print('scan (thread 0), cycles after barrier 2 per two-frame trip:', [int(w[40 + k] - w[3]) for k in range(16) if w[40 + k]])
if os.environ.get('P2C_WARP_CLOCKS'):
    wc = (ctypes.c_longlong * 512)()
    lib.p2c_debug_warp_clocks.argtypes = [ctypes.c_void_p]
    assert lib.p2c_debug_warp_clocks(wc) == 0
    tags = ['', '#1', '#2', '#3', '#4', '#4b', '#5', '#5a', '#5b', '#6', '#7', '#8']
    arr = [[wc[wp * 16 + k] for k in range(16)] for wp in range(32)]
    nw = max(wp for wp in range(32) if arr[wp][2]) + 1
    prev = None
    for k in range(1, 12):
        t = [arr[wp][k] for wp in range(nw)]
        if not all(t):
            continue
        rel = max(t)
        late = sorted(range(nw), key=lambda wp: -t[wp])[:4]
        early = sorted(t)[nw // 2]
        print('barrier %-3s released %6d after the previous one; median warp arrived %5d before the release; last warps: %s' % (
            tags[k], rel - prev if prev else 0, rel - early, ', '.join('w%d (-%d)' % (wp, rel - t[wp]) for wp in late)))
        prev = rel
if w[50] and w[54]:
    print('fold (service warp of the last CTA): partial stored + ticket known +%d, fence + partials read +%d, shuffles +%d, totals stored +%d => %d cycles after its barrier 5' % (
        w[51] - w[50], w[52] - w[51], w[53] - w[52], w[54] - w[53], w[54] - w[50]))
ct = (ctypes.c_ulonglong * 2048)()
lib.p2c_debug_cta_times.argtypes = [ctypes.c_void_p]
assert lib.p2c_debug_cta_times(ct) == 0
import numpy as np
t = np.array(list(ct), dtype=np.int64).reshape(-1, 2)
t = t[t[:, 0] > 0]
t0 = t[:, 0].min()
st, en = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
tiles = None
print('%d CTAs: start us min %.1f med %.1f max %.1f | end us min %.1f med %.1f max %.1f | life us min %.1f med %.1f max %.1f' % (
    len(t), st.min(), np.median(st), st.max(), en.min(), np.median(en), en.max(), (en - st).min(), np.median(en - st), (en - st).max()))
life = en - st
if os.environ.get('P2C_CLOCKS_ALL'):
    print('life by CTA (us):', ' '.join('%.1f' % x for x in life))
    print('start by CTA (us):', ' '.join('%.1f' % x for x in st))
order = np.argsort(life)
print('shortest-lived CTAs', order[:6].tolist(), 'longest-lived', order[-6:].tolist())
tt = np.array(list(ct), dtype=np.int64).reshape(-1, 2)
k = int(os.environ.get('P2C_TILE_CTA', '0'))
print('CTA %d: %d cycles in %.1f us => SM clock %.3f GHz' % (k, w[20] - w[16], (tt[k, 1] - tt[k, 0]) / 1e3,
                                                           (w[20] - w[16]) / float(tt[k, 1] - tt[k, 0])))
