"""Golden vectors for the SURVEY 8f rows (extra losses, Normalizer dim=3, ReferenceSkeletonsDeNormalizer, the
absolute_loc movements output) from the REFERENCE'S OWN Python code, with the same third-party stand-ins as
make_golden.py (run in the build container only: `python tests/golden/make_golden_widen.py`).
A separate file and generator so the fixtures of make_golden.py stay byte-identical.

Outputs: tests/golden/reference_widen.npz (+ .json manifest)."""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as G  # noqa: E402

SEED = G.SEED + 100


def main():
    names = G.install_stubs()
    import warnings
    warnings.simplefilter('ignore')
    from pedestrians_video_2_carla.data.carla import reference as ref
    from pedestrians_video_2_carla.data.carla.skeleton import CARLA_SKELETON
    from pedestrians_video_2_carla.loss import (cum_pose_changes, loc_2d, loc_2d_loc_rot_3d, loc_3d, loc_rot_3d,
                                                pose_changes, rot_3d, weighted_loc_2d_loc_rot_3d)
    from pedestrians_video_2_carla.modules.flow.output_types import MovementsModelOutputType
    from pedestrians_video_2_carla.modules.layers.projection import ProjectionModule
    from pedestrians_video_2_carla.transforms.pose.normalization.hips_neck_extractor import HipsNeckExtractor
    from pedestrians_video_2_carla.transforms.pose.normalization.normalizer import Normalizer
    from pedestrians_video_2_carla.transforms.pose.normalization.reference_skeletons_denormalizer import \
        ReferenceSkeletonsDeNormalizer

    g = torch.Generator().manual_seed(SEED)
    out = {}
    ages = ['adult', 'child', 'adult', 'child', 'adult']
    genders = ['female', 'male', 'male', 'female', 'male']
    meta = {'age': ages, 'gender': genders}
    N, L = len(ages), 6
    mse = torch.nn.MSELoss(reduction='mean')
    ident6 = torch.tensor([1., 0, 0, 0, 1, 0])

    # -- 1. the remaining pose_lifting losses on the pose_changes path, fwd + bwd through the reference graph
    d6 = (torch.randn((N, L, 26, 6), generator=g) * 0.15 + ident6).requires_grad_(True)
    mats = G.rotation_6d_to_matrix(d6)
    mats.retain_grad()
    pm = ProjectionModule(movements_output_type=MovementsModelOutputType.pose_changes)
    pm.on_batch_start((torch.zeros((N, L, 26, 2)), None, meta), 0)
    proj, d = pm(mats, None, None)
    norm = Normalizer(HipsNeckExtractor(CARLA_SKELETON))
    proj_t = norm(proj)
    with torch.no_grad():
        t6 = torch.randn((N, L, 26, 6), generator=g) * 0.15 + ident6
        tmats = G.rotation_6d_to_matrix(t6)
        tproj, td = pm(tmats, None, None)
        tproj_t = Normalizer(HipsNeckExtractor(CARLA_SKELETON))(tproj)
        drop = torch.rand((N, L, 26), generator=g) < 0.1
        drop[..., CARLA_SKELETON.crl_hips__C.value] = False
        tproj_t[drop] = 0.0
    targets = {'projection_2d_transformed': tproj_t, 'absolute_pose_loc': td['absolute_pose_loc'],
               'absolute_pose_rot': td['absolute_pose_rot'], 'pose_changes': tmats}
    kw = dict(criterion=mse, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON)
    l2d = loc_2d.Loc2DPoseLoss(mask_missing_joints=True, **kw)(targets=targets, projection_2d=proj,
                                                               projection_2d_transformed=proj_t)
    l3d = loc_3d.calculate_loss_loc_3d(absolute_pose_loc=d['absolute_pose_loc'], targets=targets, **kw)
    r3d = rot_3d.calculate_loss_rot_3d(absolute_pose_rot=d['absolute_pose_rot'], targets=targets, **kw)
    req = {'loc_2d': l2d, 'loc_3d': l3d, 'rot_3d': r3d}
    weights = {'loc_2d': 0.5, 'loc_3d': 2.0, 'rot_3d': 3.0}
    lw = weighted_loc_2d_loc_rot_3d.calculate_loss_weighted_loc_2d_loc_rot_3d(requirements=req, loss_weights=weights)
    cum = cum_pose_changes.calculate_loss_cum_pose_changes(criterion=mse, pose_inputs=mats, targets=targets)
    pc = pose_changes.calculate_loss_pose_changes(criterion=torch.nn.MSELoss(reduction='sum'), pose_inputs=mats,
                                                  targets=targets)
    out['loss/d6'], out['loss/mats'], out['loss/gt_pose_changes'] = d6, mats, tmats
    out['loss/gt_projection_2d_transformed'] = tproj_t
    out['loss/gt_absolute_pose_loc'], out['loss/gt_absolute_pose_rot'] = td['absolute_pose_loc'], td['absolute_pose_rot']
    out['loss/absolute_pose_rot'] = d['absolute_pose_rot']
    out['loss/loc_2d'], out['loss/loc_3d'], out['loss/rot_3d'] = l2d, l3d, r3d
    out['loss/loc_rot_3d'] = loc_rot_3d.calculate_loss_loc_rot_3d(requirements=req)
    out['loss/loc_2d_loc_rot_3d'] = loc_2d_loc_rot_3d.calculate_loss_loc_2d_loc_rot_3d(requirements=req)
    out['loss/weighted'], out['loss/cum_pose_changes'], out['loss/pose_changes'] = lw, cum, pc
    out['loss/weights'] = torch.tensor([weights['loc_2d'], weights['loc_3d'], weights['rot_3d']])
    for name, val in (('weighted', lw), ('cum_pose_changes', cum), ('pose_changes', pc), ('rot_3d', r3d)):
        d6.grad = None
        mats.grad = None
        val.backward(retain_graph=True)
        out['loss/grad_mats_' + name] = mats.grad.clone()
        out['loss/grad_d6_' + name] = d6.grad.clone()

    # -- 2. Normalizer(dim=3), ReferenceSkeletonsDeNormalizer.from_abs / from_projection, fwd + bwd
    den = ReferenceSkeletonsDeNormalizer()
    x3 = (td['absolute_pose_loc'] * 1.7 + torch.randn((N, L, 26, 3), generator=g) * 0.05 + 0.3).requires_grad_(True)
    n3 = Normalizer(HipsNeckExtractor(CARLA_SKELETON))(x3, dim=3)
    a3 = den.from_abs(x3, meta, autonormalize=True)
    w3 = torch.randn((N, L, 26, 3), generator=g)
    (a3 * w3).sum().backward()
    out['den/x3'], out['den/norm3'], out['den/from_abs'], out['den/w3'], out['den/grad_x3'] = x3, n3, a3, w3, x3.grad
    x3b = x3.detach().clone().requires_grad_(True)
    (Normalizer(HipsNeckExtractor(CARLA_SKELETON))(x3b, dim=3) * w3).sum().backward()
    out['den/grad_x3_norm3'] = x3b.grad
    out['den/from_abs_plain'] = den.from_abs(n3.detach(), meta, autonormalize=False)
    x2 = tproj.detach().clone()                      # (N, L, 26, 3): pixels + inverse depth
    x2[..., :2] += torch.randn((N, L, 26, 2), generator=g) * 3.0
    x2.requires_grad_(True)
    p2 = den.from_projection(x2, meta, autonormalize=True)
    w2 = torch.randn((N, L, 26, 3), generator=g)
    (p2 * w2).sum().backward()
    out['den/x2'], out['den/from_projection'], out['den/w2'], out['den/grad_x2'] = x2, p2, w2, x2.grad
    out['den/from_projection_2ch'] = den.from_projection(x2.detach()[..., :2].contiguous(), meta, autonormalize=True)

    # -- 3. ProjectionModule with the absolute_loc / absolute_loc_rot movements output, fwd + bwd
    pa = ProjectionModule(movements_output_type=MovementsModelOutputType.absolute_loc)
    pa.on_batch_start((torch.zeros((N, L, 26, 2)), None, meta), 0)
    xa = x3.detach().clone().requires_grad_(True)
    aproj, ad = pa(xa, None, None)
    wp = torch.randn((N, L, 26, 3), generator=g) * torch.tensor([1e-2, 1e-2, 1.0])
    (aproj * wp).sum().backward()
    out['absloc/x'], out['absloc/projection_2d'], out['absloc/absolute_pose_loc'] = xa, aproj, ad['absolute_pose_loc']
    out['absloc/w'], out['absloc/grad_x'] = wp, xa.grad
    assert ad['absolute_pose_rot'] is None and ad['relative_pose_rot'] is None and ad['relative_pose_loc'] is None

    # -- 4. metrics (metrics/mpjpe.py, mrpe.py, pck.py) with a stand-in for torchmetrics.Metric (state = attributes)
    class Metric(object):
        def __init__(self, dist_sync_on_step=False, **kw):
            pass

        def add_state(self, name, default, dist_reduce_fx=None):
            setattr(self, name, default.clone())
    G._mod('torchmetrics', Metric=Metric)
    from pedestrians_video_2_carla.metrics.mpjpe import MPJPE
    from pedestrians_video_2_carla.metrics.mrpe import MRPE
    from pedestrians_video_2_carla.metrics.pck import PCK
    wlc_p = torch.randn((N, L, 3), generator=g) * 0.05
    wlc_t = torch.randn((N, L, 3), generator=g) * 0.05
    noisy = tproj.detach().clone()
    noisy[..., :2] += torch.randn((N, L, 26, 2), generator=g) * 12.0
    tp = tproj.detach().clone()
    tp[drop] = 0.0                                   # missing joints are perfect zeros in the targets
    tp[1, 2] = 0.0                                   # a frame with no detection at all
    noisy_t = proj_t.detach() + torch.randn((N, L, 26, 3), generator=g) * torch.tensor([0.05, 0.05, 0.0])
    preds = {'absolute_pose_loc': d['absolute_pose_loc'].detach(), 'world_loc_changes': wlc_p,
             'world_loc': None, 'projection_2d': noisy, 'projection_2d_transformed': noisy_t}
    tgts = {'absolute_pose_loc': td['absolute_pose_loc'], 'world_loc_changes': wlc_t, 'projection_2d': tp,
            'projection_2d_transformed': tproj_t}
    out['metric/pred_abs_loc'], out['metric/gt_abs_loc'] = preds['absolute_pose_loc'], tgts['absolute_pose_loc']
    out['metric/pred_wlc'], out['metric/gt_wlc'] = wlc_p, wlc_t
    out['metric/pred_projection_2d'], out['metric/gt_projection_2d'] = noisy, tp
    out['metric/pred_projection_2d_transformed'], out['metric/gt_projection_2d_transformed'] = noisy_t, tproj_t
    m = MPJPE()
    m.update(preds, tgts)
    m.update(preds, tgts)                            # running state over two updates
    out['metric/mpjpe'] = m.compute()
    m = MRPE()
    m.update(preds, tgts)
    out['metric/mrpe'] = m.compute()
    m = MRPE()
    m.update(dict(preds, world_loc_changes=None, world_loc=torch.cumsum(wlc_p, dim=1) * 1.5), tgts)
    out['metric/mrpe_world_loc'] = m.compute()
    for name, kw in (('bbox', {}), ('hn', {'get_normalization_tensor': 'hn'}),
                     ('bbox_nomask', {'mask_missing_joints': False}),
                     ('hn_transformed', {'get_normalization_tensor': 'hn', 'key': 'projection_2d_transformed',
                                         'threshold': 0.1})):
        m = PCK(**kw)
        m.update(preds, tgts)
        out['metric/pck_' + name] = torch.stack((m.correct.double(), m.total.double(), m.compute().double()))

    # -- 5. Normalizer with the bbox / hips_neck_bbox extractors (dataset-input transforms, base_datamodule.py:61,276)
    from pedestrians_video_2_carla.transforms.pose.normalization.bbox_extractor import BBoxExtractor
    from pedestrians_video_2_carla.transforms.pose.normalization.hips_neck_bbox_fallback_extractor import \
        HipsNeckBBoxFallbackExtractor
    xin = tproj.detach().clone()
    xin[..., 2] = torch.rand((N, L, 26), generator=g)              # confidences
    undetected = torch.rand((N, L, 26), generator=g) < 0.15
    xin[undetected] = 0.0
    xin[0, 0, CARLA_SKELETON.crl_hips__C.value] = 0.0               # undetected hips -> fallback scale
    xin[0, 1, CARLA_SKELETON.crl_neck__C.value] = 0.0               # undetected neck -> fallback scale
    xin[2, 3] = 0.0                                                 # nothing detected at all
    out['ext/x'] = xin
    out['ext/bbox'] = Normalizer(BBoxExtractor(CARLA_SKELETON))(xin)
    out['ext/hips_neck_bbox'] = Normalizer(HipsNeckBBoxFallbackExtractor(CARLA_SKELETON))(xin)
    out['ext/hips_neck_bbox_2ch'] = Normalizer(HipsNeckBBoxFallbackExtractor(CARLA_SKELETON))(xin[..., :2].contiguous())

    skel_id = [ref.CARLA_REFERENCE_SKELETON_TYPES.index((a, s)) for a, s in zip(ages, genders)]
    arrays = {k: v.detach().numpy() for k, v in out.items()}
    arrays['skeleton_id'] = np.asarray(skel_id, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, 'reference_widen.npz'), **arrays)
    manifest = {'seed': SEED, 'ages': ages, 'genders': genders, 'keys': {k: list(v.shape) for k, v in arrays.items()},
                'reference_commit': '410bba4', 'torch': torch.__version__,
                'note': 'generated by tests/golden/make_golden_widen.py from the reference code with third-party '
                        'stubs'}
    json.dump(manifest, open(os.path.join(HERE, 'reference_widen.json'), 'w'), indent=1)
    print('wrote', len(arrays), 'arrays;', os.path.getsize(os.path.join(HERE, 'reference_widen.npz')), 'bytes')
    for k in ('loss/rot_3d', 'loss/weighted', 'loss/cum_pose_changes', 'loss/pose_changes'):
        print(k, float(arrays[k]))


if __name__ == '__main__':
    main()
