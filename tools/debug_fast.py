"""Debug helper: fast vs generic fused path vs the float64 oracle, error broken down by bone / frame."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
import torch
import oracle as O
from _util import make_case, make_targets, oracle_step
from pedestrians_video_2_carla_b200 import _lib, ops

n, l = int(sys.argv[1]), int(sys.argv[2])
c = make_case(n, l, 7 * n + l, world=(n % 2 == 1))
gt2, gt3 = make_targets(c)
d6 = c['d6'].clone().requires_grad_(True)
want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
want['loc_2d_3d'].backward()
cam = c['cam']
dev = 'cuda:0'
cu = lambda t: None if t is None else t.float().to(dev)
wl = wr = None
if c['wl'] is not None:
    wl, wr = O.calculate_world_from_changes((n, l), c['wl'], c['wr'], dtype=torch.float64)
outs = {}
for name, env in (('fast', '0'), ('generic', '1')):
    os.environ['P2C_DISABLE_FAST'] = env
    x = cu(c['d6']).requires_grad_(True)
    res = ops.fused_step(x, _lib.make_tree(c['par']), _lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py),
                         cu(c['loc']), cu(c['rot']), cu(gt2), cu(gt3), ref_index=c['sid'].int().to(dev), world_loc=cu(wl), world_rot=cu(wr))
    res.loc_2d_3d.backward()
    outs[name] = (res, x.grad.double().cpu())
ref = d6.grad
scale = ref.abs().max()
for name in outs:
    e = (outs[name][1] - ref).abs() / scale
    print(name, 'max scaled err', float(e.max()), 'loss', float(outs[name][0].loc_2d_3d), float(want['loc_2d_3d']))
    print('  by bone ', ['%.1e' % v for v in e.amax(dim=(0, 1, 3)).tolist()])
    print('  by frame', ['%.1e' % v for v in e.amax(dim=(0, 2, 3)).tolist()])
    for k in ('absolute_pose_loc', 'projection_2d', 'relative_pose_rot'):
        got = getattr(outs[name][0], k).double().cpu()
        print('  ', k, float((got - want[k]).abs().max() / want[k].abs().max()))
