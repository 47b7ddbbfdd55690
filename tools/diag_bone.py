"""Where does one bone's gradient error come from?  python tools/diag_bone.py [bone clips clip_len seed]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402

import oracle as O  # noqa: E402
from _util import make_case, make_targets, oracle_step  # noqa: E402
from pedestrians_video_2_carla_b200 import _lib, ops  # noqa: E402

bone, n, l, seed = (int(v) for v in (sys.argv[1:5] + ['8', '256', '16', '7'])[:4])
c = make_case(n, l, seed, world=False)
gt2, gt3 = make_targets(c)
dev = 'cuda:0'
cam = c['cam']
tree, cams = _lib.make_tree(c['par']), _lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py)
for name, kw in (('both', {}), ('2d only', dict(w3=0.0)), ('3d only', dict(w2=0.0)), ('no normalise', dict(normalize=False)),
                 ('no mask', dict(mask=False)), ('relative_rot', dict(mode='relative_rot'))):
    w2, w3 = kw.get('w2', 1.0), kw.get('w3', 1.0)
    normalize, mask, mode = kw.get('normalize', True), kw.get('mask', True), kw.get('mode', 'pose_changes')
    g2, g3 = make_targets(c, mode, normalize) if (not normalize or mode != 'pose_changes') else (gt2, gt3)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), mode, normalize, mask, gt=(g2, g3))
    (w2 * want['loc_2d'] + w3 * want['loc_3d']).backward()
    x = c['d6'].float().to(dev).requires_grad_(True)
    res = ops.fused_step(x, tree, cams, c['loc'].float().to(dev), c['rot'].float().to(dev), g2.float().to(dev),
                         g3.float().to(dev), ref_index=c['sid'].int().to(dev), normalize=normalize, mask_missing_joints=mask,
                         weight_2d=w2, weight_3d=w3, mode=_lib.MODE_POSE_CHANGES if mode == 'pose_changes' else _lib.MODE_RELATIVE_ROT)
    res.loc_2d_3d.backward()
    got, ref = x.grad.double().cpu(), d6.grad
    err = (got - ref).abs()
    sc_b = ref.abs().amax(dim=(0, 1, 3))
    e_b = err.amax(dim=(0, 1, 3))
    print('== %s: bone %d scale %.3e (global %.3e) err %.3e rel %.2e | worst other bone rel %.2e' % (
        name, bone, sc_b[bone], ref.abs().max(), e_b[bone], e_b[bone] / sc_b[bone],
        max(float(e_b[b] / sc_b[b]) for b in range(26) if b != bone and sc_b[b] > 0)))
    eb = err[:, :, bone]
    flat = eb.flatten().topk(4)
    for v, i in zip(flat.values.tolist(), flat.indices.tolist()):
        nn, ll, ch = i // (l * 6), (i // 6) % l, i % 6
        print('   n %d frame %d ch %d: got % .6e ref % .6e' % (nn, ll, ch, got[nn, ll, bone, ch], ref[nn, ll, bone, ch]))
    print('   per-frame max err:', ' '.join('%.1e' % v for v in eb.amax(dim=(0, 2)).tolist()))
    print('   per-channel max err:', ' '.join('%.1e' % v for v in eb.amax(dim=(0, 1)).tolist()),
          ' scale per channel:', ' '.join('%.1e' % v for v in ref[:, :, bone].abs().amax(dim=(0, 1)).tolist()))
