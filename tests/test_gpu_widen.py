"""SURVEY 8f rows 1-2 on the GPU: the remaining LossModes of the pose_lifting flow, Normalizer(dim=3),
ReferenceSkeletonsDeNormalizer and the absolute_loc(_rot) movements output of ProjectionModule, through the drop-in
classes (-> ctypes -> C ABI -> kernels), against (a) the golden vectors produced by the reference's own code
(tests/golden/reference_widen.npz) and (b) the float64 oracle on larger seeded cases.  Tolerance 1e-4 scale-relative."""
import os

import numpy as np
import pytest
import torch

import oracle as O
from _util import assert_close, assert_close_fp32_aware

pytestmark = pytest.mark.gpu
DEV = torch.device('cuda', 0)
HERE = os.path.dirname(os.path.abspath(__file__))
TYPES = (('adult', 'female'), ('adult', 'male'), ('child', 'female'), ('child', 'male'))


@pytest.fixture(scope='module')
def gold():
    z = np.load(os.path.join(HERE, 'golden', 'reference_widen.npz'))
    return {k: torch.from_numpy(z[k]) for k in z.files}


@pytest.fixture(scope='module')
def api():
    from pedestrians_video_2_carla_b200 import ops
    from pedestrians_video_2_carla_b200.loss import LossModes
    from pedestrians_video_2_carla_b200.modules.flow_output_types import MovementsModelOutputType
    from pedestrians_video_2_carla_b200.modules.layers.projection import ProjectionModule
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.denormalizer import DeNormalizer
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.normalizer import Normalizer
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.reference_skeletons_denormalizer import \
        ReferenceSkeletonsDeNormalizer
    return dict(locals())


def _meta(sid):
    return {'age': [TYPES[int(i)][0] for i in sid], 'gender': [TYPES[int(i)][1] for i in sid]}


def _losses(api, mats, proj_t, d, targets, weights=None):
    """modules/flow/base.py:440-469: requirements first, then the dependent modes."""
    L = api['LossModes']
    kw = dict(input_nodes=None, output_nodes=None, mask_missing_joints=True)
    sliced = dict(pose_inputs=mats, targets=targets, projection_2d_transformed=proj_t, **d)
    res = {}
    for mode in (L.loc_2d, L.loc_3d, L.rot_3d, L.cum_pose_changes, L.pose_changes):
        fn, crit = mode.value[0], mode.value[1]
        if isinstance(fn, type):
            from pedestrians_video_2_carla_b200.data.carla.skeleton import CARLA_SKELETON
            res[mode.name] = fn(criterion=crit, input_nodes=CARLA_SKELETON, output_nodes=CARLA_SKELETON,
                                mask_missing_joints=True)(**sliced)
        else:
            res[mode.name] = fn(criterion=crit, **kw, **sliced)
    for mode in (L.loc_2d_3d, L.loc_rot_3d, L.loc_2d_loc_rot_3d, L.weighted_loc_2d_loc_rot_3d):
        res[mode.name] = mode.value[0](requirements=res, loss_weights=weights, **kw)
    return res


def test_extra_losses_against_reference_golden(api, gold):
    sid = gold['skeleton_id']
    n, l = gold['loss/d6'].shape[:2]
    pm = api['ProjectionModule'](api['MovementsModelOutputType'].pose_changes)
    pm.on_batch_start((torch.zeros((n, l, 26, 2), device=DEV), None, _meta(sid)), 0)
    d6 = gold['loss/d6'].to(DEV).requires_grad_(True)
    mats = api['ops'].rotation_6d_to_matrix(d6)
    mats.retain_grad()
    proj, d = pm(mats, None, None)
    proj_t = api['Normalizer']()(proj)
    targets = {'projection_2d_transformed': gold['loss/gt_projection_2d_transformed'].to(DEV),
               'absolute_pose_loc': gold['loss/gt_absolute_pose_loc'].to(DEV),
               'absolute_pose_rot': gold['loss/gt_absolute_pose_rot'].to(DEV),
               'pose_changes': gold['loss/gt_pose_changes'].to(DEV)}
    w = dict(zip(('loc_2d', 'loc_3d', 'rot_3d'), gold['loss/weights'].tolist()))
    res = _losses(api, mats, proj_t, d, targets, w)
    for ours, theirs in (('loc_2d', 'loc_2d'), ('loc_3d', 'loc_3d'), ('rot_3d', 'rot_3d'), ('loc_rot_3d', 'loc_rot_3d'),
                         ('loc_2d_loc_rot_3d', 'loc_2d_loc_rot_3d'), ('weighted_loc_2d_loc_rot_3d', 'weighted'),
                         ('cum_pose_changes', 'cum_pose_changes'), ('pose_changes', 'pose_changes')):
        assert_close(ours, res[ours], gold['loss/' + theirs])
    for ours, theirs in (('weighted_loc_2d_loc_rot_3d', 'weighted'), ('cum_pose_changes', 'cum_pose_changes'),
                         ('pose_changes', 'pose_changes'), ('rot_3d', 'rot_3d')):
        d6.grad = None
        mats.grad = None
        res[ours].backward(retain_graph=True)
        assert_close('grad_mats_' + ours, mats.grad, gold['loss/grad_mats_' + theirs], 2e-4)
        assert_close('grad_d6_' + ours, d6.grad, gold['loss/grad_d6_' + theirs], 2e-4)


def test_rotation_losses_fused_in_the_step_against_reference_golden(api, gold):
    """rot_3d / loc_rot_3d / loc_2d_loc_rot_3d / weighted_loc_2d_loc_rot_3d through ProjectionModule.fused_step (ONE
    launch: rot_3d is formed where abs_rot sits in shared memory, its gradient joins the torque pass) against the values
    and d/d(6-D) the reference's own code produced (tests/golden/reference_widen.npz)."""
    sid = gold['skeleton_id']
    n, l = gold['loss/d6'].shape[:2]
    pm = api['ProjectionModule'](api['MovementsModelOutputType'].pose_changes)
    pm.on_batch_start((torch.zeros((n, l, 26, 2), device=DEV), None, _meta(sid)), 0)
    targets = {'projection_2d_transformed': gold['loss/gt_projection_2d_transformed'].to(DEV),
               'absolute_pose_loc': gold['loss/gt_absolute_pose_loc'].to(DEV),
               'absolute_pose_rot': gold['loss/gt_absolute_pose_rot'].to(DEV)}
    w = dict(zip(('loc_2d', 'loc_3d', 'rot_3d'), gold['loss/weights'].tolist()))
    for mode, theirs in (('rot_3d', 'rot_3d'), ('loc_rot_3d', 'loc_rot_3d'), ('loc_2d_loc_rot_3d', 'loc_2d_loc_rot_3d'),
                         ('weighted_loc_2d_loc_rot_3d', 'weighted')):
        d6 = gold['loss/d6'].to(DEV).requires_grad_(True)
        proj, out = pm.fused_step(d6, targets, loss_weights=w, loss_mode=mode)
        assert_close(mode, out['loss'][mode], gold['loss/' + theirs])
        assert_close('rot_3d', out['loss']['rot_3d'], gold['loss/rot_3d'])
        assert_close('absolute_pose_rot', out['absolute_pose_rot'], gold['loss/absolute_pose_rot'])
        if 'loss/grad_d6_' + theirs in gold:
            out['loss'][mode].backward()
            assert_close('grad_d6_' + mode, d6.grad, gold['loss/grad_d6_' + theirs], 2e-4)
    with pytest.raises(NotImplementedError):
        pm.fused_step(gold['loss/d6'].to(DEV), targets, loss_mode='cum_pose_changes')


def test_missing_targets_give_none(api):
    L = api['LossModes']
    x = torch.zeros((1, 2, 26, 3, 3), device=DEV)
    assert L.rot_3d.value[0](criterion=None, input_nodes=None, output_nodes=None, absolute_pose_rot=x,
                             targets={}) is None
    assert L.rot_3d.value[0](criterion=None, input_nodes=None, output_nodes=None, absolute_pose_rot=None,
                             targets={'absolute_pose_rot': x}) is None
    assert L.loc_rot_3d.value[0](requirements={'loc_3d': x.sum()}) is None
    assert L.loc_2d_loc_rot_3d.value[0](requirements={'loc_3d': x.sum(), 'rot_3d': x.sum()}) is None
    assert L.weighted_loc_2d_loc_rot_3d.value[0](requirements={'loc_2d': x.sum()}, loss_weights={}) is None


@pytest.mark.parametrize('n,l,b', [(64, 16, 26), (3, 1, 26), (5, 37, 24), (1, 64, 1)])
def test_cumulative_changes_against_oracle(api, n, l, b):
    g = torch.Generator().manual_seed(22742 + n + l)
    # non-orthonormal changes are accepted, as everywhere on the path (tests/walker_control/test_p3d_pose.py:93-95)
    c = torch.eye(3, dtype=torch.float64) + torch.randn((n, l, b, 3, 3), generator=g, dtype=torch.float64) * 0.08
    t = torch.eye(3, dtype=torch.float64) + torch.randn((n, l, b, 3, 3), generator=g, dtype=torch.float64) * 0.08
    c64 = c.clone().requires_grad_(True)
    want = O.loss_cum_pose_changes(c64, t)
    want.backward()
    c32 = c.float().requires_grad_(True)
    want32 = O.loss_cum_pose_changes(c32, t.float())
    want32.backward()
    cg = c.float().to(DEV).requires_grad_(True)
    cum = api['ops'].cumulative_changes(cg)
    assert_close('cumulative', cum, O.cumulative_changes(c))
    got = api['LossModes'].cum_pose_changes.value[0](criterion=None, pose_inputs=cg,
                                                     targets={'pose_changes': t.float().to(DEV)})
    assert_close('cum_pose_changes', got, want)
    got.backward()
    assert_close_fp32_aware('grad', cg.grad, c64.grad, c32.grad)


def test_cumulative_changes_empty_and_bad_shapes(api):
    out = api['ops'].cumulative_changes(torch.zeros((0, 4, 26, 3, 3), device=DEV))
    assert out.shape == (0, 4, 26, 3, 3)
    with pytest.raises(RuntimeError):
        api['ops'].cumulative_changes(torch.zeros((2, 4, 26, 9), device=DEV))
    with pytest.raises(RuntimeError):
        api['ops'].cumulative_changes(torch.zeros((2, 4, 26, 3, 3)))      # CPU tensor: no fallback


def test_denormalizer_against_reference_golden(api, gold):
    sid = gold['skeleton_id']
    meta = _meta(sid)
    den = api['ReferenceSkeletonsDeNormalizer']()
    x3 = gold['den/x3'].to(DEV).requires_grad_(True)
    assert_close('norm3', api['Normalizer']()(x3, dim=3), gold['den/norm3'])
    a3 = den.from_abs(x3, meta, autonormalize=True)
    assert_close('from_abs', a3, gold['den/from_abs'])
    (a3 * gold['den/w3'].to(DEV)).sum().backward()
    assert_close('grad_x3', x3.grad, gold['den/grad_x3'], 2e-4)
    x3b = gold['den/x3'].to(DEV).requires_grad_(True)
    (api['Normalizer']()(x3b, dim=3) * gold['den/w3'].to(DEV)).sum().backward()
    assert_close('grad_x3_norm3', x3b.grad, gold['den/grad_x3_norm3'], 2e-4)
    assert_close('from_abs_plain', den.from_abs(gold['den/norm3'].to(DEV), meta), gold['den/from_abs_plain'])
    x2 = gold['den/x2'].to(DEV).requires_grad_(True)
    p2 = den.from_projection(x2, meta, autonormalize=True)
    assert_close('from_projection', p2, gold['den/from_projection'])
    (p2 * gold['den/w2'].to(DEV)).sum().backward()
    assert_close('grad_x2', x2.grad, gold['den/grad_x2'], 2e-4)
    p2c = den.from_projection(gold['den/x2'][..., :2].contiguous().to(DEV), meta, autonormalize=True)
    assert_close('from_projection_2ch', p2c, gold['den/from_projection_2ch'])
    # age / gender substitutions (reference_skeletons_denormalizer.py:11-32)
    sub = den.from_abs(gold['den/x3'][:2].to(DEV), {'age': ['senior', float('nan')], 'gender': ['neutral', 'male']},
                       autonormalize=True)
    plain = den.from_abs(gold['den/x3'][:2].to(DEV), {'age': ['adult', 'adult'], 'gender': ['female', 'male']},
                         autonormalize=True)
    assert torch.equal(sub, plain)


def test_denormalizer_from_reference_matches_oracle(api):
    g = torch.Generator().manual_seed(5)
    ref = torch.randn((3, 26, 3), generator=g, dtype=torch.float64)
    x = torch.randn((3, 4, 26, 3), generator=g, dtype=torch.float64)
    shift, scale = O.get_shift_scale(ref)
    want = O.denormalize(x, scale, shift, 3)
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.hips_neck_extractor import HipsNeckExtractor
    got = api['DeNormalizer'].from_reference(HipsNeckExtractor(), ref.float().to(DEV))(x.float().to(DEV), dim=3)
    assert_close('denormalize', got, want)


def test_degenerate_scale_is_zeroed_like_the_reference(api):
    """hips == neck position: the reference divides by 0 and nan_to_zero()s the result (normalizer.py:27-29)."""
    x = torch.randn((2, 3, 26, 3), generator=torch.Generator().manual_seed(1))
    x[0, 1, O.NECK] = x[0, 1, O.HIPS]
    want = O.normalize(x.double(), 3)
    xg = x.to(DEV).requires_grad_(True)
    got = api['Normalizer']()(xg, dim=3)
    assert_close('normalize dim 3', got, want)
    assert float(got.detach()[0, 1].abs().max()) == 0.0
    got.sum().backward()
    assert torch.isfinite(xg.grad).all() and float(xg.grad[0, 1].abs().max()) == 0.0


def test_absolute_loc_movements_against_reference_golden(api, gold):
    sid = gold['skeleton_id']
    n, l = gold['absloc/x'].shape[:2]
    T = api['MovementsModelOutputType']
    pm = api['ProjectionModule'](T.absolute_loc)
    pm.on_batch_start((torch.zeros((n, l, 26, 2), device=DEV), None, _meta(sid)), 0)
    x = gold['absloc/x'].to(DEV).requires_grad_(True)
    proj, d = pm(x, None, None)
    assert d['relative_pose_loc'] is None and d['relative_pose_rot'] is None and d['absolute_pose_rot'] is None
    assert_close('absolute_pose_loc', d['absolute_pose_loc'], gold['absloc/absolute_pose_loc'])
    assert_close('projection_2d', proj, gold['absloc/projection_2d'])
    (proj * gold['absloc/w'].to(DEV)).sum().backward()
    assert_close('grad_x', x.grad, gold['absloc/grad_x'], 2e-4)
    # absolute_loc_rot: same locations, the rotations are handed through (projection.py:138-142)
    pr = api['ProjectionModule'](T.absolute_loc_rot)
    pr.on_batch_start((torch.zeros((n, l, 26, 2), device=DEV), None, _meta(sid)), 0)
    rot = torch.eye(3, device=DEV).expand(n, l, 26, 3, 3).contiguous()
    proj2, d2 = pr((gold['absloc/x'].to(DEV), rot), None, None)
    assert torch.equal(proj2, proj.detach()) and d2['absolute_pose_rot'] is rot
    with pytest.raises(RuntimeError):
        pr(gold['absloc/x'].to(DEV), None, None)


def test_absolute_loc_large_batch_against_oracle(api):
    """BASELINE config 2 size (256 x 16) through the absolute_loc entry vs the float64 oracle."""
    n, l = 256, 16
    g = torch.Generator().manual_seed(22742 + 7)
    loc, rot, _, par = O.reference_tensors(torch.float64)
    ident = torch.eye(3, dtype=torch.float64).reshape(1, 1, 3, 3).repeat(4, 26, 1, 1)
    ref_abs, _, _ = O.p3d_pose_forward(ident, loc, rot, par)
    sid = torch.randint(0, 4, (n,), generator=g)
    x = (ref_abs[sid].unsqueeze(1) * 0.6 + torch.randn((n, l, 26, 3), generator=g, dtype=torch.float64) * 0.05)
    x64 = x.clone().requires_grad_(True)
    abs_loc = O.abs_from_abs_loc(x64, ref_abs, sid)
    cam = O.training_camera()
    zl, zr = torch.zeros((n * l, 3), dtype=torch.float64), torch.eye(3, dtype=torch.float64).repeat(n * l, 1, 1)
    proj = O.project(abs_loc.reshape(n * l, 26, 3), zl, zr, cam).reshape(n, l, 26, 3)
    w = torch.randn((n, l, 26, 3), generator=g, dtype=torch.float64) * torch.tensor([1e-2, 1e-2, 1.0])
    (proj * w).sum().backward()
    pm = api['ProjectionModule'](api['MovementsModelOutputType'].absolute_loc)
    pm.on_batch_start((torch.zeros((n, l, 26, 2), device=DEV), None, _meta(sid)), 0)
    xg = x.float().to(DEV).requires_grad_(True)
    got, d = pm(xg, None, None)
    assert_close('absolute_pose_loc', d['absolute_pose_loc'], abs_loc)
    assert_close('projection_2d', got, proj)
    (got * w.float().to(DEV)).sum().backward()
    assert_close('grad', xg.grad, x64.grad, 2e-4)


@pytest.mark.parametrize('shape', [(1000,), (7, 5, 26), (1,), (0,)])
def test_quaternion_to_matrix_against_oracle(api, shape):
    """The north star's quaternion entry (pytorch3d quaternion_to_matrix, real part first, not normalised first)."""
    g = torch.Generator().manual_seed(22742 + len(shape))
    q = torch.randn(tuple(shape) + (4,), generator=g, dtype=torch.float64)
    w = torch.randn(tuple(shape) + (3, 3), generator=g, dtype=torch.float64)
    q64 = q.clone().requires_grad_(True)
    want = O.quaternion_to_matrix(q64)
    (want * w).sum().backward()
    qg = q.float().to(DEV).requires_grad_(True)
    got = api['ops'].quaternion_to_matrix(qg)
    assert got.shape == tuple(shape) + (3, 3)
    if q.numel() == 0:
        return
    assert_close('matrix', got, want)
    (got * w.float().to(DEV)).sum().backward()
    assert_close('grad', qg.grad, q64.grad, 2e-4)


def test_unit_quaternions_match_scipy_and_feed_the_fk(api):
    from scipy.spatial.transform import Rotation
    g = torch.Generator().manual_seed(3)
    q = torch.randn((64, 4), generator=g, dtype=torch.float64)
    q = q / q.norm(dim=-1, keepdim=True)
    want = Rotation.from_quat(q[:, [1, 2, 3, 0]].numpy()).as_matrix()      # scipy is scalar-last
    got = api['ops'].quaternion_to_matrix(q.float().to(DEV))
    assert_close('vs scipy', got, torch.from_numpy(want))
    with pytest.raises(RuntimeError):
        api['ops'].quaternion_to_matrix(torch.zeros((3, 3), device=DEV))


def test_metrics_against_reference_golden(gold):
    """MPJPE / MRPE / PCK classes (kernel accumulators) vs the reference's torchmetrics classes run on the same
    inputs; counts are exact, distances within 1e-4."""
    from pedestrians_video_2_carla_b200.metrics import MPJPE, MRPE, PCK
    d = {k: v.to(DEV) for k, v in gold.items() if k.startswith('metric/')}
    preds = {'absolute_pose_loc': d['metric/pred_abs_loc'], 'world_loc_changes': d['metric/pred_wlc'],
             'world_loc': None, 'projection_2d': d['metric/pred_projection_2d'],
             'projection_2d_transformed': d['metric/pred_projection_2d_transformed']}
    tgts = {'absolute_pose_loc': d['metric/gt_abs_loc'], 'world_loc_changes': d['metric/gt_wlc'],
            'projection_2d': d['metric/gt_projection_2d'],
            'projection_2d_transformed': d['metric/gt_projection_2d_transformed']}
    m = MPJPE()
    m.update(preds, tgts)
    m.update(preds, tgts)
    assert_close('mpjpe', m.compute(), gold['metric/mpjpe'])
    m.reset()
    m.update({}, tgts)                                # missing key: silently skipped like the reference (KeyError)
    assert float(m.state[1]) == 0.0
    r = MRPE()
    r.update(preds, tgts)
    assert_close('mrpe', r.compute(), gold['metric/mrpe'])
    r = MRPE()
    r.update(dict(preds, world_loc_changes=None, world_loc=torch.cumsum(d['metric/pred_wlc'], dim=1) * 1.5), tgts)
    assert_close('mrpe world_loc', r.compute(), gold['metric/mrpe_world_loc'])
    for name, kw in (('bbox', {}), ('hn', {'get_normalization_tensor': 'hn'}),
                     ('bbox_nomask', {'mask_missing_joints': False}),
                     ('hn_transformed', {'get_normalization_tensor': 'hn', 'key': 'projection_2d_transformed',
                                         'threshold': 0.1})):
        p = PCK(**kw)
        p.update(preds, tgts)
        want = gold['metric/pck_' + name]
        assert (float(p.state[0]), float(p.state[1])) == (float(want[0]), float(want[1])), name
        assert_close('pck ' + name, p.compute(), want[2].float())


def test_metrics_large_batch_against_oracle():
    """256 x 16 (BASELINE config 2 size): kernel accumulators vs the float64 oracle, run twice (determinism)."""
    from pedestrians_video_2_carla_b200.metrics import MPJPE, PCK
    g = torch.Generator().manual_seed(22742 + 9)
    n, l = 256, 16
    gt = torch.randn((n, l, 26, 3), generator=g, dtype=torch.float64)
    pred = gt + torch.randn((n, l, 26, 3), generator=g, dtype=torch.float64) * 0.05
    want = O.metric_mpjpe(pred, gt)
    vals = []
    for _ in range(2):
        m = MPJPE()
        m.update({'absolute_pose_loc': pred.float().to(DEV)}, {'absolute_pose_loc': gt.float().to(DEV)})
        vals.append(m.compute())
    assert torch.equal(vals[0], vals[1])
    assert_close('mpjpe', vals[0], want)
    g2 = gt.clone()
    g2[..., :2] = g2[..., :2] * 40 + 300
    g2[torch.rand((n, l, 26), generator=g) < 0.1] = 0
    p2 = g2 + torch.randn((n, l, 26, 3), generator=g, dtype=torch.float64) * torch.tensor([2.0, 2.0, 0.0])
    # decisions within 1e-3 of the threshold can flip in fp32: allow that many
    correct, total = O.metric_pck(p2.float().double(), g2.float().double(), g2.float().double(), 'bbox', 0.05)
    p = PCK()
    p.update({'projection_2d': p2.float().to(DEV)}, {'projection_2d': g2.float().to(DEV)})
    assert float(p.state[1]) == total
    assert abs(float(p.state[0]) - correct) <= 3


def test_bbox_extractors_against_reference_golden(api, gold):
    """Normalizer(BBoxExtractor) and Normalizer(HipsNeckBBoxFallbackExtractor) (the reference's default
    `hips_neck_bbox` transform) incl. undetected hips / neck / whole frames."""
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.bbox_extractor import BBoxExtractor
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.hips_neck_bbox_fallback_extractor import \
        HipsNeckBBoxFallbackExtractor
    x = gold['ext/x'].to(DEV)
    N = api['Normalizer']
    assert_close('bbox', N(BBoxExtractor())(x), gold['ext/bbox'])
    assert_close('hips_neck_bbox', N(HipsNeckBBoxFallbackExtractor())(x), gold['ext/hips_neck_bbox'])
    assert_close('hips_neck_bbox 2ch', N(HipsNeckBBoxFallbackExtractor())(x[..., :2].contiguous()),
                 gold['ext/hips_neck_bbox_2ch'])
    # with hips and neck detected everywhere it equals the plain hips/neck normaliser
    y = gold['den/x2'].to(DEV)
    assert_close('== hips_neck', N(HipsNeckBBoxFallbackExtractor())(y), N()(y), 1e-6)


def test_normalizer_reports_shift_and_scale(api, gold):
    """Normalizer.shift / .scale after a call (reference normalizer.py:39-49, read by projection_2d_mixin.py:229-230)
    for the three extractors, incl. undetected hips / neck / whole frames (NaN / inf scales compare as such)."""
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.bbox_extractor import BBoxExtractor
    from pedestrians_video_2_carla_b200.transforms.pose.normalization.hips_neck_bbox_fallback_extractor import \
        HipsNeckBBoxFallbackExtractor
    N = api['Normalizer']
    x = gold['ext/x']
    for name, ext in (('hips_neck', None), ('bbox', BBoxExtractor()), ('hips_neck_bbox', HipsNeckBBoxFallbackExtractor())):
        n = N(ext) if ext is not None else N()
        with pytest.raises(RuntimeError):
            n.scale
        n(x.to(DEV))
        shift, scale = O.extractor_shift_scale(x.double(), name)
        got_shift, got_scale = n.shift.double().cpu(), n.scale.double().cpu()
        assert got_shift.shape == shift.shape and got_scale.shape == scale.shape
        fin = torch.isfinite(scale)
        assert torch.equal(torch.isfinite(got_scale), fin), name
        assert_close(name + ' scale', got_scale[fin], scale[fin])
        fs = torch.isfinite(shift).all(-1)
        assert torch.equal(torch.isfinite(got_shift).all(-1), fs), name
        assert_close(name + ' shift', got_shift[fs], shift[fs])
    # dim = 3 (absolute locations): hips / neck over three channels
    x3 = gold['den/x3'].to(DEV)
    n3 = N()
    n3(x3, dim=3)
    s3, c3 = O.get_shift_scale(gold['den/x3'].double())
    assert_close('shift dim3', n3.shift, s3)
    assert_close('scale dim3', n3.scale, c3.reshape(n3.scale.shape))


def test_device_target_generator(api):
    """Carla2D3DIterableDataset.generate_batch on the device: the distribution's invariants and, given the drawn Euler
    changes, parity of every target with the oracle's ProjectionModule / Normalizer."""
    from pedestrians_video_2_carla_b200.data.carla.datasets.carla_2d3d_dataset import Carla2D3DIterableDataset
    ds = Carla2D3DIterableDataset(batch_size=12, clip_length=9, random_changes_each_frame=3, max_change_in_deg=5,
                                  max_world_rot_change_in_deg=2, max_initial_world_rot_change_in_deg=30,
                                  transform='hips_neck', device=DEV, seed=22742)
    x, t, meta = ds.generate_batch()
    assert x.shape == (12, 9, 26, 2) and len(meta['age']) == 12
    e = t['euler_pose_changes']
    # the draws themselves: ONE kernel, restated bit for bit by the oracle (counter-based Philox stream)
    we, ww, ws = O.draw_sparse_changes(22742, 0, 12, 9, 26, 3, np.float32(np.deg2rad(5)), np.float32(np.deg2rad(2)),
                                       np.float32(np.deg2rad(30)))
    assert np.array_equal(e.cpu().numpy(), we)
    assert np.array_equal(meta['skeleton_index'].cpu().numpy(), ws)
    yaw = torch.atan2(t['world_rot_changes'][..., 1, 0], t['world_rot_changes'][..., 0, 0]).cpu().numpy()
    assert np.allclose(yaw, ww[..., 2], atol=1e-6)
    assert [TYPES[i] for i in ws.tolist()] == list(zip(meta['age'], meta['gender']))
    changed = (e.abs().sum(-1) > 0).sum(-1)
    assert int(changed.min()) == 3 and int(changed.max()) == 3           # exactly 3 distinct bones per frame
    assert float(e.abs().max()) <= np.deg2rad(5) + 1e-7
    assert float(t['world_loc_changes'].abs().max()) == 0.0
    # the deterministic part against the oracle, from the generator's own draws
    loc, rot, _, par = O.reference_tensors(torch.float64)
    sid = torch.tensor([TYPES.index((a, g)) for a, g in zip(meta['age'], meta['gender'])])
    mats = O.euler_angles_to_matrix(e.double().cpu())
    assert_close('pose_changes', t['pose_changes'], mats)
    wr = O.euler_angles_to_matrix(torch.stack((torch.zeros(12, 9), torch.zeros(12, 9),
                                               torch.atan2(t['world_rot_changes'][..., 1, 0],
                                                           t['world_rot_changes'][..., 0, 0]).cpu()), -1).double())
    assert_close('world yaw only', t['world_rot_changes'], wr)
    proj, out = O.projection_module_forward(mats, loc[sid], rot[sid], par, O.training_camera(),
                                            torch.zeros((12, 9, 3), dtype=torch.float64), wr)
    assert_close('absolute_pose_loc', t['absolute_pose_loc'], out['absolute_pose_loc'])
    assert_close('projection_2d', t['projection_2d'], proj[..., :2])
    assert_close('projection_2d_transformed', t['projection_2d_transformed'], O.hips_neck_normalize(proj)[..., :2])
    assert torch.equal(x, t['projection_2d_transformed'])
    # same seed, same batch; and the iterator yields per-clip items
    ds2 = Carla2D3DIterableDataset(batch_size=12, clip_length=9, max_world_rot_change_in_deg=2,
                                   max_initial_world_rot_change_in_deg=30, transform='hips_neck', device=DEV, seed=22742)
    assert torch.equal(ds2.generate_batch()[0], x)
    second = ds2.generate_batch()
    assert not torch.equal(second[0], x)      # the batch counter advances the stream
    we2, _, _ = O.draw_sparse_changes(22742, 1, 12, 9, 26, 3, np.float32(np.deg2rad(5)), np.float32(np.deg2rad(2)),
                                      np.float32(np.deg2rad(30)))
    assert np.array_equal(second[1]['euler_pose_changes'].cpu().numpy(), we2)
    item = next(iter(ds2))
    assert item[0].shape == (9, 26, 2) and item[1]['absolute_pose_loc'].shape == (9, 26, 3)
