"""Debug driver for the tile kernel: one small case against the oracle, per-output errors."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
import torch
import oracle as O
from _util import make_case, make_targets, oracle_step, scaled_err
from pedestrians_video_2_carla_b200 import ops, _lib
_lib.load()
n, l = int(sys.argv[1]), int(sys.argv[2])
world = len(sys.argv) > 3 and sys.argv[3] == 'w'
c = make_case(n, l, 1234, world=world)
gt2, gt3 = make_targets(c)
d6 = c['d6'].clone().requires_grad_(True)
want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
want['loc_2d_3d'].backward()
DEV = 'cuda:0'
cu = lambda t: None if t is None else t.float().to(DEV)
cam = c['cam']
wl = wr = None
if world:
    wl, wr = O.calculate_world_from_changes((c['n'], c['l']), c['wl'], c['wr'], dtype=torch.float64)
x = cu(c['d6']).requires_grad_(True)
res = ops.fused_step(x, _lib.make_tree(c['par']), _lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py),
                     cu(c['loc']), cu(c['rot']), cu(gt2), cu(gt3), ref_index=c['sid'].int().to(DEV), world_loc=cu(wl),
                     world_rot=cu(wr), hips=c['hips'], neck=c['neck'])
torch.cuda.synchronize()
print('forward ok')
res.loc_2d_3d.backward()
torch.cuda.synchronize()
for k in ('relative_pose_rot', 'absolute_pose_rot', 'absolute_pose_loc', 'projection_2d', 'projection_2d_transformed',
          'loc_2d', 'loc_3d', 'loc_2d_3d'):
    print('%-28s %.3e' % (k, scaled_err(getattr(res, k), want[k])))
print('%-28s %.3e' % ('grad', scaled_err(x.grad, d6.grad)))
