"""GPU parity of p2c_fused_step (ProjectionModule -> Normalizer -> loc_2d/loc_3d/loc_2d_3d fwd + bwd) against the
oracle, the committed golden fixture (outputs of the reference's own code), and size-independent properties at
BASELINE.json's full sizes."""
import numpy as np
import pytest
import torch

import oracle as O
from _util import (REL_TOL, assert_close, assert_close_fp32_aware, cast_case, make_case, make_targets, oracle_step,
                   smpl_parents)

pytestmark = pytest.mark.gpu

DEV = 'cuda:0'
# the golden fixture holds the reference's own FLOAT32 run: two fp32 evaluations are compared, each ~1e-4 of a
# (bone, channel) group from the exact value on the worst group
GOLDEN_TOL, GOLDEN_GRAD_TOL = 2e-4, 4e-4


@pytest.fixture(scope='module')
def ops():
    from pedestrians_video_2_carla_b200 import ops as _ops
    return _ops


@pytest.fixture(scope='module')
def lib():
    from pedestrians_video_2_carla_b200 import _lib
    _lib.load()
    return _lib


def cu(t):
    return None if t is None else t.float().to(DEV)


def run_cuda(ops, lib, c, pose, gt2, gt3, mode, normalize, mask, w2=1.0, w3=1.0):
    cam = c['cam']
    wl = wr = None
    if c['wl'] is not None:
        wl, wr = O.calculate_world_from_changes((c['n'], c['l']), c['wl'], c['wr'], dtype=torch.float64)
    x = cu(pose).requires_grad_(True)
    res = ops.fused_step(x, lib.make_tree(c['par']), lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py),
                         cu(c['loc']), cu(c['rot']), cu(gt2), cu(gt3), ref_index=c['sid'].int().to(DEV),
                         world_loc=cu(wl), world_rot=cu(wr),
                         mode=lib.MODE_POSE_CHANGES if mode == 'pose_changes' else lib.MODE_RELATIVE_ROT,
                         hips=c['hips'], neck=c['neck'], normalize=normalize, mask_missing_joints=mask,
                         weight_2d=w2, weight_3d=w3)
    res.loc_2d_3d.backward()
    return res, x.grad


class Ref32(object):
    """The step evaluated by the oracle in float32: its gradient and its outputs (what the reference PyTorch path itself
    delivers in fp32), the yardstick of assert_close_fp32_aware."""

    def __init__(self, grad, out):
        self.grad, self.out = grad, {k: v.detach() for k, v in out.items() if torch.is_tensor(v)}


def oracle_grad32(c, gt2, gt3, mode='pose_changes', normalize=True, mask=True, pose=None, w2=1.0, w3=1.0):
    """The same step evaluated by the oracle in float32 (= the reference PyTorch path's own precision)."""
    c32 = cast_case(c, torch.float32)
    x = (c32['d6'] if pose is None else pose.float()).clone().requires_grad_(True)
    mats = O.rotation_6d_to_matrix(x) if x.shape[-1] == 6 and x.dim() == 4 else x
    out = oracle_step(c32, mats, mode, normalize, mask, gt=(gt2.float(), gt3.float()))
    (w2 * out['loc_2d'] + w3 * out['loc_3d']).backward()
    return Ref32(x.grad, out)


def compare(res, grad, want, ref_grad, normalize=True, ref_grad32=None, grad_tol=REL_TOL):
    assert_close('relative_pose_rot', res.relative_pose_rot, want['relative_pose_rot'])
    assert_close('absolute_pose_rot', res.absolute_pose_rot, want['absolute_pose_rot'])
    assert_close('absolute_pose_loc', res.absolute_pose_loc, want['absolute_pose_loc'])
    assert_close('projection_2d', res.projection_2d, want['projection_2d'])
    if normalize and ref_grad32 is not None:
        # (p - p_hips) / scale of pixel coordinates ~400: joints next to the hips are small differences of large
        # numbers, where the reference's own fp32 evaluation is 1e-4 of the joint's scale from the exact value
        assert_close_fp32_aware('projection_2d_transformed', res.projection_2d_transformed,
                                want['projection_2d_transformed'], ref_grad32.out['projection_2d_transformed'])
    elif normalize:
        assert_close('projection_2d_transformed', res.projection_2d_transformed, want['projection_2d_transformed'])
    assert_close('loc_2d', res.loc_2d, want['loc_2d'])
    assert_close('loc_3d', res.loc_3d, want['loc_3d'])
    assert_close('loc_2d_3d', res.loc_2d_3d, want['loc_2d_3d'])
    if ref_grad32 is None:
        assert_close('grad', grad, ref_grad)
    else:
        assert_close_fp32_aware('grad', grad, ref_grad, ref_grad32.grad, tol=grad_tol)


@pytest.mark.parametrize('mode', ['pose_changes', 'relative_rot'])
@pytest.mark.parametrize('six_d', [True, False])
@pytest.mark.parametrize('normalize,mask', [(True, True), (False, False), (True, False)])
def test_fused_step_carla(ops, lib, mode, six_d, normalize, mask):
    c = make_case(12, 16, 100 + 7 * six_d + normalize + 2 * mask, mode=mode)
    gt2, gt3 = make_targets(c, mode, normalize)
    d6 = c['d6'].clone().requires_grad_(True)
    mats = O.rotation_6d_to_matrix(d6)
    mats.retain_grad()
    want = oracle_step(c, mats, mode, normalize, mask, gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    if six_d:
        res, grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, mode, normalize, mask)
        compare(res, grad, want, d6.grad, normalize, oracle_grad32(c, gt2, gt3, mode, normalize, mask))
    else:
        res, grad = run_cuda(ops, lib, c, mats.detach(), gt2, gt3, mode, normalize, mask)
        compare(res, grad, want, mats.grad, normalize,
                oracle_grad32(c, gt2, gt3, mode, normalize, mask, pose=mats.detach()))


@pytest.mark.parametrize('n,l', [(1, 1), (3, 1), (2, 64), (33, 5), (64, 16)])
def test_fused_step_shapes(ops, lib, n, l):
    """Ragged / extreme clip shapes incl. C1 (64x16, reference tests/walker_control batch)."""
    c = make_case(n, l, 7 * n + l, world=(n % 2 == 1))
    gt2, gt3 = make_targets(c)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    res, grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, 'pose_changes', True, True)
    compare(res, grad, want, d6.grad, ref_grad32=oracle_grad32(c, gt2, gt3))


@pytest.mark.parametrize('variant', ['P2C_DISABLE_TILE', 'P2C_FORCE_SLIM', 'P2C_DISABLE_TMA', 'P2C_DISABLE_FAST'])
@pytest.mark.parametrize('mode', ['pose_changes', 'relative_rot'])
def test_fused_step_kernel_variants(ops, lib, monkeypatch, request, variant, mode):
    """Every older kernel variant behind p2c_fused_step (bone-per-lane double-buffered TMA kernel, its 4-CTA/SM slim
    form, the register-staged kernel, the exact generic composite) against the oracle on the same case; the default
    (pose-per-lane tile kernel, fused_tile.cu) is what every other test in this file runs."""
    monkeypatch.setenv(variant, '1')
    lib.load().p2c_debug_reload_env()    # the switches are read once per process: re-read them for this test ...
    request.addfinalizer(lambda: (monkeypatch.undo(), lib.load().p2c_debug_reload_env()))   # ... and after it
    c = make_case(20, 16, 4242, mode=mode)
    gt2, gt3 = make_targets(c, mode)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), mode, gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    res, grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, mode, True, True)
    # the bone-per-lane fallbacks form subtree sums as DIFFERENCES of one warp prefix scan over all bones: bones with a
    # short lever arm (toes, fingers) carry the rounding of the whole prefix -- 2e-4 of their (small) scale
    compare(res, grad, want, d6.grad, ref_grad32=oracle_grad32(c, gt2, gt3, mode),
            grad_tol=REL_TOL if variant == 'P2C_DISABLE_FAST' else 2e-4)


def test_fused_step_weights(ops, lib):
    c = make_case(6, 4, 5)
    gt2, gt3 = make_targets(c)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
    (0.25 * want['loc_2d'] + 3.0 * want['loc_3d']).backward()
    res, grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, 'pose_changes', True, True, w2=0.25, w3=3.0)
    assert_close('weighted loss', res.loc_2d_3d, 0.25 * want['loc_2d'] + 3.0 * want['loc_3d'])
    assert_close_fp32_aware('weighted grad', grad, d6.grad, oracle_grad32(c, gt2, gt3, w2=0.25, w3=3.0).grad)


def test_fused_step_smpl_tree(ops, lib):
    """data/smpl/utils.py:106 drives P3dPose with the 22-bone SMPL tree: the path must accept arbitrary trees."""
    c = make_case(5, 8, 77, par=smpl_parents())
    gt2, gt3 = make_targets(c)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    res, grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, 'pose_changes', True, True)
    compare(res, grad, want, d6.grad, ref_grad32=oracle_grad32(c, gt2, gt3))


def test_fused_step_non_orthonormal_matrices(ops, lib):
    """Matrix input is not assumed to be a rotation (reference tests/walker_control/test_p3d_pose.py:93-95)."""
    c = make_case(4, 6, 31)
    gt2, gt3 = make_targets(c)
    mats = (O.rotation_6d_to_matrix(c['d6']) + 0.1 * torch.randn((4, 6, 26, 3, 3), generator=c['gen'],
                                                                dtype=torch.float64)).requires_grad_(True)
    want = oracle_step(c, mats, gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    res, grad = run_cuda(ops, lib, c, mats.detach(), gt2, gt3, 'pose_changes', True, True)
    compare(res, grad, want, mats.grad, ref_grad32=oracle_grad32(c, gt2, gt3, pose=mats.detach()))


@pytest.mark.parametrize('mode', ['pose_changes', 'relative_rot'])
def test_fused_step_golden_fixture(ops, lib, golden, mode):
    """tests/golden/reference_run.npz: outputs of the reference's OWN code for the full step."""
    k = mode + '/'
    loc, rot, _, par = O.reference_tensors(torch.float64)
    n, l = golden[k + 'd6'].shape[:2]
    c = dict(n=n, l=l, b=26, par=par, loc=loc, rot=rot, sid=torch.as_tensor(golden['skeleton_id']).long(),
             wl=torch.as_tensor(golden[k + 'world_loc_change']).double(),
             wr=torch.as_tensor(golden[k + 'world_rot_change']).double(), cam=O.training_camera(), hips=O.HIPS,
             neck=O.NECK)
    res, grad = run_cuda(ops, lib, c, torch.as_tensor(golden[k + 'd6']).double(),
                         torch.as_tensor(golden[k + 'gt_projection_2d_transformed']).double(),
                         torch.as_tensor(golden[k + 'gt_absolute_pose_loc']).double(), mode, True, True)
    for key, got in (('relative_pose_rot', res.relative_pose_rot), ('absolute_pose_loc', res.absolute_pose_loc),
                     ('absolute_pose_rot', res.absolute_pose_rot), ('projection_2d', res.projection_2d),
                     ('projection_2d_transformed', res.projection_2d_transformed), ('loc_2d', res.loc_2d),
                     ('loc_3d', res.loc_3d), ('loc_2d_3d', res.loc_2d_3d)):
        assert_close('golden ' + key, got, torch.as_tensor(np.asarray(golden[k + key])), tol=GOLDEN_TOL)
    assert_close('golden grad_d6', grad, torch.as_tensor(golden[k + 'grad_d6']), tol=GOLDEN_GRAD_TOL)


@pytest.mark.parametrize('n,l', [(256, 16), (4096, 64)])
def test_full_size_properties(ops, lib, n, l):
    """BASELINE configs C2 (256x16) and C4 (4096x64): properties that need no oracle run at full size.
    (1) identity changes => rel_rot == reference rot, every clip of one skeleton type equal
        (reference tests/walker_control/test_p3d_pose.py:58-91);
    (2) clip independence: a slice of the batch run alone gives bit-identical outputs;
    (3) targets == own outputs => loss 0 and gradient 0."""
    g = torch.Generator().manual_seed(n + l)
    loc, rot, _, par = O.reference_tensors(torch.float32)
    cam = O.training_camera()
    tree, cams = lib.make_tree(par), lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py)
    sid = torch.randint(0, 4, (n,), generator=g).int().to(DEV)
    loc, rot = loc.to(DEV), rot.to(DEV)
    ident = torch.tensor([1., 0, 0, 0, 1, 0]).expand(n, l, 26, 6).contiguous().to(DEV)
    zeros = torch.zeros((n, l, 26, 3), device=DEV)
    r = ops.fused_step(ident, tree, cams, loc, rot, zeros, zeros, ref_index=sid)
    assert torch.equal(r.relative_pose_rot, rot[sid.long()].unsqueeze(1).expand(n, l, 26, 3, 3))
    first = [int((sid == k).nonzero()[0]) for k in range(4)]
    for k in range(4):
        same = r.absolute_pose_loc[sid == k]
        assert torch.equal(same, same[:1].expand_as(same))
        assert torch.equal(r.projection_2d[first[k], 0], r.projection_2d[first[k], l - 1])

    d6 = (torch.randn((n, l, 26, 6), generator=g) * 0.3 + torch.tensor([1., 0, 0, 0, 1, 0])).to(DEV)
    full = ops.fused_step(d6, tree, cams, loc, rot, zeros, zeros, ref_index=sid)
    lo, hi = n // 3, n // 3 + 17
    part = ops.fused_step(d6[lo:hi].contiguous(), tree, cams, loc, rot, zeros[lo:hi].contiguous(),
                          zeros[lo:hi].contiguous(), ref_index=sid[lo:hi].contiguous())
    assert torch.equal(part.absolute_pose_loc, full.absolute_pose_loc[lo:hi])
    assert torch.equal(part.projection_2d, full.projection_2d[lo:hi])
    assert torch.equal(part.relative_pose_rot, full.relative_pose_rot[lo:hi])

    x = d6.clone().requires_grad_(True)
    again = ops.fused_step(x, tree, cams, loc, rot, full.projection_2d_transformed, full.absolute_pose_loc,
                           ref_index=sid, mask_missing_joints=False)
    again.loc_2d_3d.backward()
    assert float(again.loc_2d_3d.detach()) == 0.0
    assert float(x.grad.abs().max()) == 0.0


@pytest.mark.parametrize('n,l', [(2048, 16), (640, 64)])
def test_many_tiles_deterministic_and_chunk_consistent(ops, lib, n, l):
    """More tiles than CTAs (tiles handed out by the grid-wide counter, mask count in the pre-pass, one loss partial per
    tile): (1) two runs are bitwise equal although the tile -> CTA assignment differs from run to run; (2) with the
    mask off, the same clips run in chunks small enough for one tile per CTA (the other kernel instance) give the same
    outputs bit for bit, gradients equal up to the 1/N of the mean, and loss sums that add up."""
    g = torch.Generator().manual_seed(7 * n + l)
    loc, rot, _, par = O.reference_tensors(torch.float32)
    cam = O.training_camera()
    tree, cams = lib.make_tree(par), lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py)
    sid = torch.randint(0, 4, (n,), generator=g).int().to(DEV)
    loc, rot = loc.to(DEV), rot.to(DEV)
    base = torch.tensor([1., 0, 0, 0, 1, 0])
    d6 = (torch.randn((n, l, 26, 6), generator=g) * 0.3 + base).to(DEV)
    d6_gt = (torch.randn((n, l, 26, 6), generator=g) * 0.3 + base).to(DEV)
    zeros = torch.zeros((n, l, 26, 3), device=DEV)
    t = ops.fused_step(d6_gt, tree, cams, loc, rot, zeros, zeros, ref_index=sid)
    gt2, gt3 = t.projection_2d_transformed.clone(), t.absolute_pose_loc.clone()
    drop = (torch.rand((n, l, 26), generator=g) < 0.1).to(DEV)
    gt2[drop] = 0

    def run(lo, hi, mask):
        x = d6[lo:hi].clone().requires_grad_(True)
        r = ops.fused_step(x, tree, cams, loc, rot, gt2[lo:hi].contiguous(), gt3[lo:hi].contiguous(),
                           ref_index=sid[lo:hi].contiguous(), mask_missing_joints=mask)
        r.loc_2d_3d.backward()
        return r, x.grad

    r1, g1 = run(0, n, True)
    r2, g2 = run(0, n, True)
    assert torch.equal(r1.losses, r2.losses)
    assert torch.equal(g1, g2)
    kept = (gt2[..., 0] != 0) & (gt2[..., 1] != 0)
    kept[..., 1] = True   # hips: always kept (utils/tensors.py:29-40)
    assert float(r1.losses[4]) == float(2 * kept.sum())   # kept (joint, channel) pairs

    full, gf = run(0, n, False)
    chunk = 256 if l == 16 else 128   # clips per chunk: 128 tiles <= SMs (one tile per CTA)
    sums = torch.zeros(4, dtype=torch.float64)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        part, gp = run(lo, hi, False)
        assert torch.equal(part.absolute_pose_loc, full.absolute_pose_loc[lo:hi])
        assert torch.equal(part.projection_2d, full.projection_2d[lo:hi])
        assert torch.equal(part.absolute_pose_rot, full.absolute_pose_rot[lo:hi])
        # (the same arithmetic up to the 1/N of the mean: a property of the whole tensor, floor = 1)
        assert_close('chunk grad', gp * ((hi - lo) / n), gf[lo:hi], tol=2e-6, floor=1.0)
        sums += part.losses[3:7].double().cpu()
    assert_close('loss sums', full.losses[3:7].double().cpu(), sums, tol=1e-6)


def test_preallocated_buffers_reuse_the_argument_block(ops, lib):
    """fused_step(out=fused_step_buffers(...)) over several steps with DIFFERENT inputs (the end-to-end loop of
    bench.py): the cached argument block must pick up each step's input pointers, and the results must equal the
    fresh-output call bit for bit."""
    cam = None
    pre = None
    tree = None
    for seed in (101, 102, 103):
        c = make_case(8, 16, seed, world=False)
        gt2, gt3 = make_targets(c)
        if pre is None:
            pre = ops.fused_step_buffers(c['n'], c['l'], c['b'], DEV, transformed=True)
            tree = lib.make_tree(c['par'])
            k = c['cam']
            cam = lib.make_camera(k.R.tolist(), k.T.tolist(), k.f, k.px, k.py)
            loc, rot = cu(c['loc']), cu(c['rot'])
        args = dict(ref_index=c['sid'].int().to(DEV), hips=c['hips'], neck=c['neck'])
        x = cu(c['d6']).requires_grad_(True)
        r = ops.fused_step(x, tree, cam, loc, rot, cu(gt2), cu(gt3), out=pre, **args)
        r.loc_2d_3d.backward()
        assert '_plan' in pre
        y = cu(c['d6']).requires_grad_(True)
        f = ops.fused_step(y, tree, cam, loc, rot, cu(gt2), cu(gt3), **args)
        f.loc_2d_3d.backward()
        assert torch.equal(r.loc_2d_3d, f.loc_2d_3d) and torch.equal(x.grad, y.grad)
        for name in ('relative_pose_rot', 'absolute_pose_rot', 'absolute_pose_loc', 'projection_2d',
                     'projection_2d_transformed'):
            assert torch.equal(getattr(r, name), getattr(f, name)), name
        want = oracle_step(c, O.rotation_6d_to_matrix(c['d6']), gt=(gt2, gt3))
        assert_close('loc_2d_3d', r.loc_2d_3d, want['loc_2d_3d'])


def test_result_backward_equals_autograd_backward(ops, lib):
    """FusedStepResult.backward() == loc_2d_3d.backward() for a leaf input and for a model output (non-leaf)."""
    c = make_case(6, 16, 77, world=False)
    gt2, gt3 = make_targets(c)
    tree = lib.make_tree(c['par'])
    k = c['cam']
    cam = lib.make_camera(k.R.tolist(), k.T.tolist(), k.f, k.px, k.py)
    args = dict(ref_index=c['sid'].int().to(DEV), hips=c['hips'], neck=c['neck'])

    def run(direct, through_model):
        w = torch.full((6,), 1.25, device=DEV, requires_grad=True)
        leaf = cu(c['d6']).requires_grad_(True)
        x = leaf * w if through_model else leaf
        r = ops.fused_step(x, tree, cam, cu(c['loc']), cu(c['rot']), cu(gt2), cu(gt3), **args)
        if direct:
            r.backward()
        else:
            r.loc_2d_3d.backward()
        return leaf.grad, w.grad
    for through_model in (False, True):
        a, aw = run(True, through_model)
        b, bw = run(False, through_model)
        assert torch.equal(a, b)
        if through_model:
            assert torch.equal(aw, bw)


@pytest.mark.parametrize('n,l', [(256, 16), (2048, 16), (640, 64)])
def test_bench_sizes_against_the_oracle(ops, lib, n, l):
    """VALUES at the sizes and kernel instances bench.py times, mask ON: 256x16 = the headline step (one tile per CTA,
    tile_kernel<26,1,0,1>, in-kernel joint count); 2048x16 / 640x64 = many tiles per CTA (tile_kernel<26,1,0,0> /
    <26,2,0,0> + tile_count_kernel, the instances of the 8192x16 and 4096x64 sweep points).  Outputs, the three losses
    and the gradient against the float64 oracle, per bone and channel."""
    c = make_case(n, l, 9000 + n + l, world=False)
    gt2, gt3 = make_targets(c)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    res, grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, 'pose_changes', True, True)
    compare(res, grad, want, d6.grad, ref_grad32=oracle_grad32(c, gt2, gt3))
    kept = (gt2[..., 0] != 0) & (gt2[..., 1] != 0)
    kept[..., c['hips']] = True
    assert float(res.losses[4]) == float(2 * kept.sum())


def test_lone_step_followed_by_many_tile_mask_off_step(ops, lib):
    """Stream order under programmatic dependent launch: a one-tile-per-CTA step triggers its dependents at its very
    top; a many-tile MASK-OFF step right behind it has no count pre-pass in front and must wait (griddepcontrol.wait)
    before it touches the workspace header or its inputs.  Both results must equal the ones of isolated runs."""
    loc, rot, _, par = O.reference_tensors(torch.float32)
    cam = O.training_camera()
    tree, cams = lib.make_tree(par), lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py)
    loc, rot = loc.to(DEV), rot.to(DEV)
    g = torch.Generator().manual_seed(99)
    base = torch.tensor([1., 0, 0, 0, 1, 0])

    def inputs(n, l):
        d6 = (torch.randn((n, l, 26, 6), generator=g) * 0.3 + base).to(DEV)
        gt3 = torch.randn((n, l, 26, 3), generator=g).to(DEV)
        gt2 = torch.randn((n, l, 26, 3), generator=g).to(DEV)
        sid = torch.randint(0, 4, (n,), generator=g).int().to(DEV)
        return d6, gt2, gt3, sid
    small, big = inputs(256, 16), inputs(4096, 16)
    outs = {k: ops.fused_step_buffers(n, 16, 26, DEV) for k, n in (('small', 256), ('big', 4096))}

    def step(which, data, mask):
        d6, gt2, gt3, sid = data
        r = ops.fused_step(d6, tree, cams, loc, rot, gt2, gt3, ref_index=sid, mask_missing_joints=mask,
                           want_transformed=False, out=outs[which])
        return r
    torch.cuda.synchronize()
    ref_small = step('small', small, True)
    torch.cuda.synchronize()
    keep_small = (ref_small.losses.clone(), ref_small.grad_pose_in.clone())
    ref_big = step('big', big, False)
    torch.cuda.synchronize()
    keep_big = (ref_big.losses.clone(), ref_big.grad_pose_in.clone(), ref_big.absolute_pose_loc.clone())
    for _ in range(20):   # back to back on one stream, no host synchronisation in between
        a = step('small', small, True)
        b = step('big', big, False)
    torch.cuda.synchronize()
    assert torch.equal(a.losses, keep_small[0]) and torch.equal(a.grad_pose_in, keep_small[1])
    assert torch.equal(b.losses, keep_big[0]) and torch.equal(b.grad_pose_in, keep_big[1])
    assert torch.equal(b.absolute_pose_loc, keep_big[2])


def test_extra_loss_on_the_module_tensors_gets_its_gradient(ops, lib):
    """The step's module tensors carry autograd edges: a caller that adds rot_3d on absolute_pose_rot and a penalty on
    projection_2d_transformed (neither fused) gets the exact gradient of the whole sum."""
    c = make_case(6, 8, 515, world=False)
    gt2, gt3 = make_targets(c)
    gt_rot = O.rotation_6d_to_matrix(c['d6_gt'])
    with torch.no_grad():
        gt_abs_rot = oracle_step(c, gt_rot)['absolute_pose_rot']
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
    extra = O.loss_rot_3d(want['absolute_pose_rot'], gt_abs_rot) + \
        0.1 * (want['projection_2d_transformed'][..., :2] ** 2).mean() + 0.5 * (want['relative_pose_rot'] ** 3).mean()
    (want['loc_2d_3d'] + extra).backward()
    cam = c['cam']
    x = cu(c['d6']).requires_grad_(True)
    res = ops.fused_step(x, lib.make_tree(c['par']), lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py),
                         cu(c['loc']), cu(c['rot']), cu(gt2), cu(gt3), ref_index=c['sid'].int().to(DEV))
    got_extra = ((res.absolute_pose_rot - cu(gt_abs_rot)) ** 2).mean() + \
        0.1 * (res.projection_2d_transformed[..., :2] ** 2).mean() + 0.5 * (res.relative_pose_rot ** 3).mean()
    (res.loc_2d_3d + got_extra).backward()
    assert_close('extra loss', got_extra, extra)
    c32 = cast_case(c, torch.float32)
    x32 = c32['d6'].clone().requires_grad_(True)
    o32 = oracle_step(c32, O.rotation_6d_to_matrix(x32), gt=(gt2.float(), gt3.float()))
    (o32['loc_2d_3d'] + O.loss_rot_3d(o32['absolute_pose_rot'], gt_abs_rot.float()) +
     0.1 * (o32['projection_2d_transformed'][..., :2] ** 2).mean() + 0.5 * (o32['relative_pose_rot'] ** 3).mean()).backward()
    assert_close_fp32_aware('grad of fused + extra losses', x.grad, d6.grad, x32.grad)


def test_two_channel_view_targets_take_the_tile_kernel(ops, lib):
    """The reference hands `projection[..., :2]` (a stride-3 view) as the 2-D target: it is passed as the 3-channel
    records it lives in, so the step stays on the tile kernel (one launch at 256x16) and equals the contiguous call."""
    c = make_case(256, 16, 616, world=False)
    gt2, gt3 = make_targets(c)
    cam = c['cam']
    tree, cams = lib.make_tree(c['par']), lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py)
    g3 = cu(gt2)
    view = g3[..., :2]
    assert not view.is_contiguous()
    rec = ops._gt2_records(view)
    assert rec.shape[-1] == 3 and rec.data_ptr() == g3.data_ptr()
    a = ops.fused_step(cu(c['d6']), tree, cams, cu(c['loc']), cu(c['rot']), view, cu(gt3), ref_index=c['sid'].int().to(DEV))
    b = ops.fused_step(cu(c['d6']), tree, cams, cu(c['loc']), cu(c['rot']), g3, cu(gt3), ref_index=c['sid'].int().to(DEV))
    assert torch.equal(a.losses, b.losses) and torch.equal(a.grad_pose_in, b.grad_pose_in)


def test_degenerate_normaliser_poses_stay_within_the_reference_noise(ops, lib):
    """UNconditioned random poses (tests/_util.py::_well_conditioned off): torsos that point at the camera bring the
    projected hips-neck distance below a pixel, the normalised coordinates into the hundreds, and every fp32
    evaluation -- the reference's own included -- to rounding noise far above 1e-4 on those poses.  The step must stay
    finite and within an order of magnitude of the fp32 reference's own deviation from float64, group by group."""
    c = make_case(256, 16, 7, world=False, conditioned=False)
    gt2, gt3 = make_targets(c)
    d6 = c['d6'].clone().requires_grad_(True)
    want = oracle_step(c, O.rotation_6d_to_matrix(d6), gt=(gt2, gt3))
    want['loc_2d_3d'].backward()
    res, grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, 'pose_changes', True, True)
    assert bool(torch.isfinite(grad).all()) and bool(torch.isfinite(res.losses).all())
    ref32 = oracle_grad32(c, gt2, gt3)
    assert_close('absolute_pose_loc', res.absolute_pose_loc, want['absolute_pose_loc'])
    assert_close('projection_2d', res.projection_2d, want['projection_2d'])
    assert_close('loc_3d', res.loc_3d, want['loc_3d'])
    assert_close_fp32_aware('loc_2d', res.loc_2d, want['loc_2d'], ref32.out['loc_2d'], slack=10.0)
    assert_close_fp32_aware('grad', grad, d6.grad, ref32.grad, slack=10.0)


@pytest.mark.parametrize('n,l,tree,w', [(12, 16, 'carla', (1.0, 1.0, 1.0)), (600, 16, 'carla', (0.5, 2.0, 3.0)),
                                        (8, 64, 'carla', (0.0, 0.0, 1.0)), (7, 5, 'carla', (1.0, 1.0, 0.7)),
                                        (6, 8, 'smpl', (1.0, 1.0, 1.0))])
def test_rot_3d_fused_in_the_step(ops, lib, n, l, tree, w):
    """weight_2d*loc_2d + weight_3d*loc_3d + weight_rot*rot_3d from one fused step (rot_3d: loss/rot_3d.py:9-37) --
    values and gradient against the float64 oracle: one tile per CTA, many tiles, two-slice clips, a shape the tile
    kernel declines (35 poses: generic composite with the 9-channel reduction), a non-CARLA tree."""
    c = make_case(n, l, 700 + n + l, par=None if tree == 'carla' else smpl_parents(), world=False)
    gt2, gt3 = make_targets(c)
    with torch.no_grad():
        gt_rot = oracle_step(c, O.rotation_6d_to_matrix(c['d6_gt']))['absolute_pose_rot']
    w2, w3, wr = w

    def oracle(case, dtype):
        x = case['d6'].to(dtype).clone().requires_grad_(True)
        out = oracle_step(case, O.rotation_6d_to_matrix(x), gt=(gt2.to(dtype), gt3.to(dtype)))
        out['rot_3d'] = O.loss_rot_3d(out['absolute_pose_rot'], gt_rot.to(dtype))
        out['total'] = w2 * out['loc_2d'] + w3 * out['loc_3d'] + wr * out['rot_3d']
        out['total'].backward()
        return x.grad, out
    g64, want = oracle(c, torch.float64)
    g32, _ = oracle(cast_case(c, torch.float32), torch.float32)
    cam = c['cam']
    x = cu(c['d6']).requires_grad_(True)
    res = ops.fused_step(x, lib.make_tree(c['par']), lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py),
                         cu(c['loc']), cu(c['rot']), cu(gt2), cu(gt3), ref_index=c['sid'].int().to(DEV), hips=c['hips'],
                         neck=c['neck'], weight_2d=w2, weight_3d=w3, gt_rot=cu(gt_rot), weight_rot=wr)
    res.total.backward()
    assert_close('rot_3d', res.rot_3d, want['rot_3d'])
    assert_close('loc_2d', res.loc_2d, want['loc_2d'])
    assert_close('loc_3d', res.loc_3d, want['loc_3d'])
    assert_close('total', res.total, want['total'])
    assert float(res.losses[10]) == n * l * c['b'] * 9
    assert_close_fp32_aware('grad', x.grad, g64, g32)


def test_one_workspace_through_shape_changes_and_a_zeroed_header(ops, lib):
    """The step's loss totals go through the caller's workspace: per-tile partials behind a ticket (many tiles) or as
    self-validating (value, launch sequence) entries folded by CTA 0 (one tile per CTA).  One workspace is taken
    through a one-tile launch, a many-tile launch, one-tile launches again, and a header that the caller zeroes in
    between (which restarts the launch sequence): every step's totals must be the ones of its own inputs."""
    from pedestrians_video_2_carla_b200 import ops as _ops
    cases = []
    for n, l, seed in ((8, 16, 3), (400, 16, 4), (8, 16, 5), (6, 16, 6), (8, 16, 7)):
        c = make_case(n, l, seed, world=False)
        gt2, gt3 = make_targets(c)
        want = oracle_step(c, O.rotation_6d_to_matrix(c['d6']), gt=(gt2, gt3))
        cases.append((c, gt2, gt3, want))
    for k, (c, gt2, gt3, want) in enumerate(cases):
        if k == 3:   # the caller zeroes the 16 header words again (and only them)
            for ws in _ops._WS.values():
                ws[:16].zero_()
        for _ in range(2):
            res, _grad = run_cuda(ops, lib, c, c['d6'], gt2, gt3, 'pose_changes', True, True)
            assert_close('loc_2d', res.loc_2d, want['loc_2d'])
            assert_close('loc_3d', res.loc_3d, want['loc_3d'])


@pytest.mark.parametrize('n,l', [(256, 16), (100, 64), (37, 8), (2048, 16)])
def test_back_to_back_launches_are_bitwise_repeatable(n, l):
    """Steps launched back to back overlap through programmatic dependent launch and share one workspace; inside a step
    the warps hand shared-memory regions to each other phase by phase.  A missing barrier or an unordered reuse shows
    up as a run-to-run difference: every output of every launch must equal the first launch's bit for bit
    (tools/stress_determinism.py runs the same check for longer)."""
    import bench
    dev = torch.device('cuda', 0)
    sh = bench.FusedShard(dev, n, l, bench.SEED, want_bytes=1, max_sets=1)
    st = sh.sets[0]
    names = ('losses', 'g_d6', 'rel_rot', 'abs_rot', 'abs_loc', 'proj')
    sh.launch(0)
    torch.cuda.synchronize()
    ref = {a: st[a].clone() for a in names}
    for i in range(25):
        for a in names[1:]:
            st[a].fill_(float('nan'))
        for _ in range(4):
            sh.launch(0)
        torch.cuda.synchronize()
        for a in names:
            assert torch.equal(st[a], ref[a]), '%s differs from the first launch at iteration %d' % (a, i)
