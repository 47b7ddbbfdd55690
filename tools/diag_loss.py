"""Is the step's loc_2d the masked MSE of its own projection_2d_transformed?  python tools/diag_loss.py [clips len seed]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402

import oracle as O  # noqa: E402
from _util import make_case, make_targets, oracle_step  # noqa: E402
from pedestrians_video_2_carla_b200 import _lib, ops  # noqa: E402

n, l, seed = (int(v) for v in (sys.argv[1:4] + ['256', '16', '7'])[:3])
c = make_case(n, l, seed, world=False)
gt2, gt3 = make_targets(c)
dev = 'cuda:0'
cam = c['cam']
want = oracle_step(c, O.rotation_6d_to_matrix(c['d6']), gt=(gt2, gt3))
x = c['d6'].float().to(dev)
res = ops.fused_step(x, _lib.make_tree(c['par']), _lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py),
                     c['loc'].float().to(dev), c['rot'].float().to(dev), gt2.float().to(dev), gt3.float().to(dev),
                     ref_index=c['sid'].int().to(dev))
pn = res.projection_2d_transformed.detach().double().cpu()
g32 = gt2.float().double()
mask = O.missing_joints_mask(g32[..., :2], c['hips'])
e = ((pn[..., :2] - g32[..., :2]) ** 2)
from_out = e[mask].mean()
print(os.path.basename(_lib.lib_path()), 'kernel loc_2d %.9e | mse of its own outputs (f64) %.9e | oracle f64 %.9e' % (
    float(res.loc_2d), float(from_out), float(want['loc_2d'])))
print('  rel: kernel vs own outputs %.2e, own outputs vs oracle %.2e, kernel vs oracle %.2e' % (
    abs(float(res.loc_2d) - float(from_out)) / float(from_out), abs(float(from_out) - float(want['loc_2d'])) / float(want['loc_2d']),
    abs(float(res.loc_2d) - float(want['loc_2d'])) / float(want['loc_2d'])))
# per-pose contribution: where do the kernel's outputs deviate from the oracle's in loss terms
eo = ((want['projection_2d_transformed'][..., :2] - gt2[..., :2]) ** 2)
d = ((e - eo) * mask.unsqueeze(-1)).sum(-1)
top = d.abs().flatten().topk(5)
tot = float(eo[mask].sum())
for v, i in zip(top.values.tolist(), top.indices.tolist()):
    nn, ll, bb = i // (l * 26), (i // 26) % l, i % 26
    print('  clip %d frame %d bone %d: term differs by %.3e (%.2e of the whole sum); pred cuda %s oracle %s gt %s 1/z %.4e' % (
        nn, ll, bb, v, v / tot, pn[nn, ll, bb, :2].tolist(), want['projection_2d_transformed'][nn, ll, bb, :2].tolist(),
        gt2[nn, ll, bb, :2].tolist(), float(pn[nn, ll, bb, 2])))
print('  sums (kernel):', res.losses.cpu().tolist())
