import sys, time, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
dev = torch.device('cuda', 0)
sh = bench.Shard(dev, 256, 16, 1, want_bytes=1, max_sets=2, capture=False)
ops = sh.ops
s = sh.sets[0]
host = {k: s[k].cpu().pin_memory() for k in ('d6', 'gt2', 'gt3', 'sid')}
dbuf = {k: torch.empty_like(s[k]) for k in host}
loss_host = torch.zeros(1).pin_memory()
def t(f, n=200):
    for _ in range(10): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
def copies():
    for k in dbuf: dbuf[k].copy_(host[k], non_blocking=True)
def step():
    x = dbuf['d6'].requires_grad_(True)
    r = ops.fused_step(x, sh.tree, sh.cam, sh.loc, sh.rot, dbuf['gt2'], dbuf['gt3'], ref_index=dbuf['sid'], want_transformed=False)
    r.loc_2d_3d.backward()
    dbuf['d6'] = x.detach()
    return r
def step_sync():
    r = step(); loss_host.copy_(r.loc_2d_3d.detach().reshape(1)); return float(loss_host[0])
def full():
    copies(); return step_sync()
print('copies only %.1f us' % t(copies))
print('step (no sync) %.1f us' % t(step))
print('step + D2H sync %.1f us' % t(step_sync))
print('full %.1f us' % t(full))
def fwd_only():
    with torch.no_grad():
        ops.fused_step(dbuf['d6'], sh.tree, sh.cam, sh.loc, sh.rot, dbuf['gt2'], dbuf['gt3'], ref_index=dbuf['sid'], want_transformed=False)
print('fused_step no-grad call %.1f us' % t(fwd_only))
def raw():
    sh.launch(0)
print('raw C-ABI launch %.1f us' % t(raw))
