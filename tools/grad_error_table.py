"""Per-bone error of the fused step's gradient (and per-channel error of its outputs) against the float64 oracle, next
to the error of the oracle evaluated in float32 (= the reference PyTorch path's own precision).  Run once per library
variant (P2C_LIB, tools/variants.py) to see what the MUFU approximations / fp32 subtree sums cost.

  python tools/grad_error_table.py [clips clip_len seed]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402

import oracle as O  # noqa: E402
from _util import cast_case, make_case, make_targets, oracle_step, group_err  # noqa: E402
from pedestrians_video_2_carla_b200 import _lib, ops  # noqa: E402


def main():
    n, l, seed = (int(v) for v in (sys.argv[1:4] + ['64', '16', '5'])[:3])
    c = make_case(n, l, seed, world=False)
    gt2, gt3 = make_targets(c)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    c32 = cast_case(c, torch.float32)
    x32 = c32['d6'].clone().requires_grad_(True)
    o32 = oracle_step(c32, O.rotation_6d_to_matrix(x32), gt=(gt2.float(), gt3.float()))
    o32['loc_2d_3d'].backward()
    dev = 'cuda:0'
    cam = c['cam']
    x = c['d6'].float().to(dev).requires_grad_(True)
    res = ops.fused_step(x, _lib.make_tree(c['par']), _lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py),
                         c['loc'].float().to(dev), c['rot'].float().to(dev), gt2.float().to(dev), gt3.float().to(dev),
                         ref_index=c['sid'].int().to(dev))
    res.loc_2d_3d.backward()
    torch.cuda.synchronize()
    out = {'lib': os.path.basename(_lib.lib_path()), 'case': [n, l, seed]}
    g_cuda = group_err(x.grad, d6.grad, (2,))          # per bone
    g_o32 = group_err(x32.grad, d6.grad, (2,))
    out['grad_per_bone_cuda'] = ['%.1e' % v for v in g_cuda.tolist()]
    out['grad_per_bone_fp32_oracle'] = ['%.1e' % v for v in g_o32.tolist()]
    out['grad_worst'] = {'cuda': float(g_cuda.max()), 'fp32_oracle': float(g_o32.max())}
    for k, got in (('absolute_pose_loc', res.absolute_pose_loc), ('projection_2d', res.projection_2d),
                   ('projection_2d_transformed', res.projection_2d_transformed),
                   ('absolute_pose_rot', res.absolute_pose_rot), ('relative_pose_rot', res.relative_pose_rot)):
        axes = (2, 3) if want[k].dim() == 4 else (2, 3, 4)
        out[k] = {'cuda': float(group_err(got, want[k], axes).max()), 'fp32_oracle': float(group_err(o32[k], want[k], axes).max())}
    for k, got in (('loc_2d', res.loc_2d), ('loc_3d', res.loc_3d)):
        out[k] = {'cuda': abs(float(got) - float(want[k])) / abs(float(want[k])),
                  'fp32_oracle': abs(float(o32[k]) - float(want[k])) / abs(float(want[k]))}
    print(json.dumps(out))


if __name__ == '__main__':
    main()
