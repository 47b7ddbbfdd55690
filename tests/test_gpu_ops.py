"""GPU parity of each C-ABI entry point against the CPU oracle (float64), through the ctypes binding."""
import pytest
import torch

import oracle as O
from _util import assert_close, assert_close_fp32_aware, carla, make_case, random_rotations, smpl_parents

pytestmark = pytest.mark.gpu

DEV = 'cuda:0'


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


def cam_struct(lib, cam):
    return lib.make_camera(cam.R.tolist(), cam.T.tolist(), cam.f, cam.px, cam.py)


@pytest.mark.parametrize('n', [1, 7, 256, 1000])
def test_rotation_6d_to_matrix(ops, n):
    g = torch.Generator().manual_seed(n)
    d6 = torch.randn((n, 6), generator=g, dtype=torch.float64)
    up = torch.randn((n, 3, 3), generator=g, dtype=torch.float64)
    ref_in = d6.clone().requires_grad_(True)
    ref = O.rotation_6d_to_matrix(ref_in)
    (ref * up).sum().backward()
    x = cu(d6).requires_grad_(True)
    out = ops.rotation_6d_to_matrix(x)
    (out * cu(up)).sum().backward()
    assert_close('rot6d fwd', out, ref)
    assert_close('rot6d bwd', x.grad, ref_in.grad)
    assert_close('matrix_to_rotation_6d', ops.matrix_to_rotation_6d(out.detach()), O.matrix_to_rotation_6d(ref.detach()))


def test_rotation_6d_degenerate(ops):
    """F.normalize eps path: zero / parallel inputs must not produce NaN in the forward."""
    d6 = torch.tensor([[0., 0, 0, 0, 1, 0], [1., 0, 0, 2, 0, 0], [1., 2, 3, 4, 5, 6]], dtype=torch.float64)
    assert_close('rot6d degenerate', ops.rotation_6d_to_matrix(cu(d6)), O.rotation_6d_to_matrix(d6))


def test_euler_conversions(ops):
    g = torch.Generator().manual_seed(3)
    e = (torch.rand((513, 3), generator=g, dtype=torch.float64) - 0.5) * 2.0
    m = ops.euler_angles_to_matrix(cu(e), 'XYZ')
    assert_close('euler->matrix', m, O.euler_angles_to_matrix(e))
    assert_close('matrix->euler', ops.matrix_to_euler_angles(m, 'XYZ'), e)
    with pytest.raises(RuntimeError):
        ops.euler_angles_to_matrix(cu(e), 'ZYX')


def test_no_cpu_fallback(ops):
    with pytest.raises(RuntimeError):
        ops.rotation_6d_to_matrix(torch.randn(4, 6))


@pytest.mark.parametrize('mode', ['pose_changes', 'relative_rot'])
@pytest.mark.parametrize('tree', ['carla', 'smpl'])
def test_fk_clip_fwd_bwd(ops, lib, mode, tree):
    """P3dPose.forward looped over frames + P3dPoseProjection, all outputs and all input gradients,
    with NON-orthonormal changes (reference tests/walker_control/test_p3d_pose.py:93-95)."""
    par = None if tree == 'carla' else smpl_parents()
    c = make_case(5, 6, 11, par=par)
    g = c['gen']
    n, l, b = c['n'], c['l'], c['b']
    mats = (O.rotation_6d_to_matrix(c['d6']) + torch.randn((n, l, b, 3, 3), generator=g, dtype=torch.float64) * 0.1)
    ref_loc, ref_rot = c['loc'][c['sid']].clone(), c['rot'][c['sid']].clone()
    wl_abs, wr_abs = O.calculate_world_from_changes((n, l), c['wl'], c['wr'], dtype=torch.float64)
    ins = [t.clone().requires_grad_(True) for t in (mats, ref_loc, ref_rot, wl_abs, wr_abs)]
    if mode == 'pose_changes':
        rl, rr, al, ar = O.abs_from_pose_changes(ins[0], ins[1], ins[2], c['par'])
    else:
        rl, rr, al, ar = O.abs_from_relative_rot(ins[0], ins[1], c['par'])
    proj = torch.stack([O.project(al[:, i], ins[3][:, i], ins[4][:, i], c['cam']) for i in range(l)], 1)
    ups = [torch.randn(t.shape, generator=g, dtype=torch.float64) for t in (rr, ar, al, proj)]
    loss = (ar * ups[1]).sum() + (al * ups[2]).sum() + (proj * ups[3]).sum() * 1e-3
    if mode == 'pose_changes':
        loss = loss + (rr * ups[0]).sum()
    loss.backward()

    cins = [cu(t).requires_grad_(True) for t in (mats, ref_loc, ref_rot, wl_abs, wr_abs)]
    if mode == 'relative_rot':  # reference skeletons are detached in the reference (p3d_pose.py:274-277); the
        cins[1] = cins[1].detach()  # per-pose kernel does not reduce d/d ref_loc over frames
    m = lib.MODE_POSE_CHANGES if mode == 'pose_changes' else lib.MODE_RELATIVE_ROT
    crr, car, cal, cproj = ops.fk_clip(cins[0], cins[1], cins[2] if mode == 'pose_changes' else None, None, cins[3],
                                       cins[4], lib.make_tree(c['par']), cam_struct(lib, c['cam']), m, True)
    closs = (car * cu(ups[1])).sum() + (cal * cu(ups[2])).sum() + (cproj * cu(ups[3])).sum() * 1e-3
    if mode == 'pose_changes':
        closs = closs + (crr * cu(ups[0])).sum()
        assert_close('rel_rot', crr, rr)
    closs.backward()
    assert_close('abs_rot', car, ar)
    assert_close('abs_loc', cal, al)
    assert_close('proj', cproj, proj)
    assert_close('g_pose_in', cins[0].grad, ins[0].grad)
    if mode == 'pose_changes':
        assert_close('g_ref_loc', cins[1].grad, ins[1].grad)
        assert_close('g_ref_rot', cins[2].grad, ins[2].grad)
    assert_close('g_world_loc', cins[3].grad, ins[3].grad)
    assert_close('g_world_rot', cins[4].grad, ins[4].grad)


def test_fk_indexed_reference_and_empty(ops, lib):
    c = make_case(9, 3, 5, world=False)
    mats = O.rotation_6d_to_matrix(c['d6'])
    _, rr, al, ar = O.abs_from_pose_changes(mats, c['loc'][c['sid']], c['rot'][c['sid']], c['par'])
    tree = lib.make_tree(c['par'])
    crr, car, cal, cproj = ops.fk_clip(cu(mats), cu(c['loc']), cu(c['rot']), c['sid'].int().to(DEV), None, None, tree,
                                       None, lib.MODE_POSE_CHANGES, False)
    assert cproj is None
    assert_close('rel_rot', crr, rr)
    assert_close('abs_rot', car, ar)
    assert_close('abs_loc', cal, al)
    # empty batch (reference: zero-length tensors flow through torch ops)
    e = ops.fk_clip(torch.empty((0, 3, 26, 3, 3), device=DEV), cu(c['loc']), cu(c['rot']),
                    torch.empty(0, dtype=torch.int32, device=DEV), None, None, tree, None, lib.MODE_POSE_CHANGES, False)
    assert e[1].shape == (0, 3, 26, 3, 3) and e[2].shape == (0, 3, 26, 3)


def test_bad_tree_is_rejected(ops, lib):
    c = make_case(2, 2, 1, world=False)
    bad = list(c['par'])
    bad[3] = 7  # parent after child: not a DFS order
    with pytest.raises(RuntimeError):
        ops.fk_clip(cu(O.rotation_6d_to_matrix(c['d6'])), cu(c['loc'][c['sid']]), cu(c['rot'][c['sid']]), None, None,
                    None, lib.make_tree(bad), None, lib.MODE_POSE_CHANGES, False)


def test_project_fwd_bwd(ops, lib):
    g = torch.Generator().manual_seed(9)
    m, j = 37, 26
    x = torch.randn((m, j, 3), generator=g, dtype=torch.float64) * 0.5
    loc = torch.randn((m, 3), generator=g, dtype=torch.float64) * 0.1
    rot = O.euler_angles_to_matrix((torch.rand((m, 3), generator=g, dtype=torch.float64) - 0.5) * 0.4)
    up = torch.randn((m, j, 3), generator=g, dtype=torch.float64)
    cam = O.training_camera()
    ins = [t.clone().requires_grad_(True) for t in (x, loc, rot)]
    ref = O.project(*ins, cam)
    (ref * up).sum().backward()
    cins = [cu(t).requires_grad_(True) for t in (x, loc, rot)]
    out = ops.project(*cins, cam_struct(lib, cam))
    (out * cu(up)).sum().backward()
    assert_close('project', out, ref)
    for name, a, b in zip(('g_x', 'g_loc', 'g_rot'), cins, ins):
        assert_close(name, a.grad, b.grad)


def test_world_scan(ops):
    g = torch.Generator().manual_seed(4)
    n, l = 13, 9
    dl = torch.randn((n, l, 3), generator=g, dtype=torch.float64)
    dr = O.euler_angles_to_matrix((torch.rand((n, l, 3), generator=g, dtype=torch.float64) - 0.5)) \
        + torch.randn((n, l, 3, 3), generator=g, dtype=torch.float64) * 0.05
    il = torch.randn((n, 3), generator=g, dtype=torch.float64)
    ir = O.euler_angles_to_matrix(torch.rand((n, 3), generator=g, dtype=torch.float64))
    ul = torch.randn((n, l, 3), generator=g, dtype=torch.float64)
    ur = torch.randn((n, l, 3, 3), generator=g, dtype=torch.float64)
    a, b = dl.clone().requires_grad_(True), dr.clone().requires_grad_(True)
    # functional restatement of utils/world.py:56-61 (the reference writes slices in place, which torch autograd
    # refuses once rot_change requires grad); values are checked against the oracle below
    rots, locs = [ir], [il]
    for i in range(l):
        rots.append(torch.bmm(rots[-1], b[:, i]))
        locs.append(locs[-1] + a[:, i])
    wl, wr = torch.stack(locs[1:], 1), torch.stack(rots[1:], 1)
    owl, owr = O.calculate_world_from_changes((n, l), dl, dr, il, ir, dtype=torch.float64)
    assert torch.equal(owl, wl.detach()) and torch.equal(owr, wr.detach())
    ((wl * ul).sum() + (wr * ur).sum()).backward()
    ca, cb = cu(dl).requires_grad_(True), cu(dr).requires_grad_(True)
    cwl, cwr = ops.world_scan(ca, cb, cu(il), cu(ir))
    ((cwl * cu(ul)).sum() + (cwr * cu(ur)).sum()).backward()
    assert_close('world_loc', cwl, wl)
    assert_close('world_rot', cwr, wr)
    assert_close('g_loc_change', ca.grad, a.grad)
    assert_close('g_rot_change', cb.grad, b.grad)


@pytest.mark.parametrize('c', [2, 3])
def test_hips_neck_normalize(ops, c):
    g = torch.Generator().manual_seed(c)
    s = torch.randn((4, 5, 26, c), generator=g, dtype=torch.float64) * 100 + 300
    if c == 3:
        s[..., 2] = torch.rand((4, 5, 26), generator=g, dtype=torch.float64) * 0.5
        s[0, 0, 3, 2] = 1e-7        # "low confidence": xy zeroed (normalizer.py:33-37)
        s[0, 1, 4, 2] = float('nan')
    s[1, 1, O.NECK] = s[1, 1, O.HIPS]  # zero scale -> inf/nan -> nan_to_zero
    up = torch.randn(s.shape, generator=g, dtype=torch.float64)
    a = s.clone().requires_grad_(True)
    ref = O.hips_neck_normalize(a)
    (ref * up).sum().backward()
    ca = cu(s).requires_grad_(True)
    out = ops.hips_neck_normalize(ca, O.HIPS, O.NECK)
    (out * cu(up)).sum().backward()
    assert_close('normalize', out, ref)
    good = torch.ones(s.shape[:2], dtype=torch.bool)
    good[1, 1] = False  # torch autograd yields NaN grads for the degenerate pose; ours are zero there
    assert_close('normalize bwd', ca.grad.cpu()[good], a.grad[good])
    assert torch.isfinite(ca.grad).all()


@pytest.mark.parametrize('mask', [True, False])
def test_masked_mse(ops, mask):
    g = torch.Generator().manual_seed(8)
    pred = torch.randn((6, 4, 26, 3), generator=g, dtype=torch.float64)
    gt = torch.randn((6, 4, 26, 3), generator=g, dtype=torch.float64)
    gt[torch.rand((6, 4, 26), generator=g) < 0.2] = 0
    a = pred.clone().requires_grad_(True)
    ref = O.loss_loc_2d(a, gt, mask)
    (ref * 3.0).backward()
    ca = cu(pred).requires_grad_(True)
    out = ops.masked_mse(ca, cu(gt), 2, O.HIPS if mask else -1)
    (out * 3.0).backward()
    assert_close('loc_2d', out, ref)
    assert_close('loc_2d bwd', ca.grad, a.grad)
    # loc_3d: plain MSE over 3 channels
    assert_close('loc_3d', ops.masked_mse(cu(pred), cu(gt), 3, -1), O.loss_loc_3d(pred, gt))


@pytest.mark.parametrize('shape,c_gt,mask', [((256, 16, 26), 2, True), ((7, 3, 26), 3, True), ((5, 1, 25), 2, False),
                                             ((8192, 16, 26), 2, True), ((1, 1, 1), 2, True)])
def test_loc2d_step_fused_forward_backward(ops, shape, c_gt, mask):
    """p2c_loc2d_step (the autoencoder flow's step, BASELINE configs[2]): loss and gradient of Loc2DPoseLoss from one
    cooperative launch against the oracle's value and autograd; odd joint counts, 3-channel target records, mask off."""
    g = torch.Generator().manual_seed(sum(shape) + c_gt)
    pred = torch.randn(shape + (2,), generator=g, dtype=torch.float64)
    gt = torch.randn(shape + (c_gt,), generator=g, dtype=torch.float64)
    gt[torch.rand(shape, generator=g) < 0.15] = 0
    hips = 1 % shape[-1]
    p64 = pred.clone().requires_grad_(True)
    want = O.loss_loc_2d(p64, gt, mask, hips)
    (3.0 * want).backward()
    x = cu(pred).requires_grad_(True)
    loss, sums = ops.loc2d_loss(x, cu(gt), hips if mask else -1, with_sums=True)
    assert sums is not None, 'the fused kernel declined a shape it should take'
    (3.0 * loss).backward()
    assert_close('loc_2d', loss, want)
    assert_close('grad', x.grad, p64.grad)
    kept = O.missing_joints_mask(gt[..., :2], hips) if mask else torch.ones(shape, dtype=torch.bool)
    assert float(sums[2]) == float(2 * kept.sum())
    again, _ = ops.loc2d_loss(cu(pred).requires_grad_(True), cu(gt), hips if mask else -1, with_sums=True)
    assert torch.equal(again, loss)     # fixed grid, fixed order: bitwise repeatable


def test_loc2d_step_declines_what_it_cannot_hold(ops):
    """No gradient wanted, 3-channel predictions: the two-stage kernels; same value."""
    g = torch.Generator().manual_seed(12)
    pred = torch.randn((9, 4, 26, 3), generator=g, dtype=torch.float64)
    gt = torch.randn((9, 4, 26, 3), generator=g, dtype=torch.float64)
    loss, sums = ops.loc2d_loss(cu(pred).requires_grad_(True), cu(gt), 1, with_sums=True)
    assert sums is None
    assert_close('loc_2d', loss, O.loss_loc_2d(pred, gt, True, 1))


def _fk_reference(c, mats, mode, ups):
    ins = [t.clone().requires_grad_(True) for t in (mats,)]
    ref_loc, ref_rot = c['loc'][c['sid']], c['rot'][c['sid']]
    n, l = c['n'], c['l']
    wl, wr = O.calculate_world_from_changes((n, l), c['wl'], c['wr'], dtype=mats.dtype)
    if mode == 'pose_changes':
        rl, rr, al, ar = O.abs_from_pose_changes(ins[0], ref_loc, ref_rot, c['par'])
    else:
        rl, rr, al, ar = O.abs_from_relative_rot(ins[0], ref_loc, c['par'])
    proj = torch.stack([O.project(al[:, i], wl[:, i], wr[:, i], c['cam']) for i in range(l)], 1)
    loss = (ar * ups[1]).sum() + (al * ups[2]).sum() + (proj * ups[3]).sum() * 1e-3
    if mode == 'pose_changes':
        loss = loss + (rr * ups[0]).sum()
    loss.backward()
    return dict(rr=rr, ar=ar, al=al, proj=proj, grad=ins[0].grad, wl=wl, wr=wr)


@pytest.mark.parametrize('tile', [True, False])
@pytest.mark.parametrize('mode', ['pose_changes', 'relative_rot'])
@pytest.mark.parametrize('n,l,tree', [(6, 16, 'carla'), (3, 64, 'carla'), (2, 100, 'carla'), (5, 7, 'carla'), (40, 1, 'carla'),
                                      (4, 40, 'smpl'), (300, 16, 'carla')])
def test_fk_module_api_kernels(ops, lib, request, monkeypatch, tile, mode, n, l, tree):
    """p2c_fk_fwd / p2c_fk_bwd as the reference's module API calls them -- 3x3 matrix input that is NOT a rotation
    (reference tests/walker_control/test_p3d_pose.py:93-95: rand * 0.8 'changes'), indexed skeleton table, world
    transform, upstream gradients on every output -- through the pose-per-lane tile kernels (fk_tile.cu: whole clips in
    the lanes, clips longer than 32 frames walked tile by tile with the scan state carried, odd clip lengths declined to
    the generic kernels) and, with P2C_DISABLE_FWD_TILE, through the generic warp-per-clip kernels."""
    if not tile:
        monkeypatch.setenv('P2C_DISABLE_FWD_TILE', '1')
    lib.load().p2c_debug_reload_env()
    request.addfinalizer(lambda: (monkeypatch.undo(), lib.load().p2c_debug_reload_env()))
    par = None if tree == 'carla' else smpl_parents()
    c = make_case(n, l, 13 * n + l, par=par)
    g = c['gen']
    b = c['b']
    spread = 0.02 if l > 32 else 0.1      # (products of 100 non-rotations: keep them from exploding)
    mats = O.rotation_6d_to_matrix(c['d6']) + torch.randn((n, l, b, 3, 3), generator=g, dtype=torch.float64) * spread
    mats = mats.float().double()     # the oracle gets exactly the fp32 matrices the kernels get
    shapes = ((n, l, b, 3, 3), (n, l, b, 3, 3), (n, l, b, 3), (n, l, b, 3))
    ups = [torch.randn(s, generator=g, dtype=torch.float64) for s in shapes]
    ref = _fk_reference(c, mats, mode, ups)
    x = cu(mats).requires_grad_(True)
    m = lib.MODE_POSE_CHANGES if mode == 'pose_changes' else lib.MODE_RELATIVE_ROT
    crr, car, cal, cproj = ops.fk_clip(x, cu(c['loc']), cu(c['rot']) if mode == 'pose_changes' else None,
                                       c['sid'].int().to(DEV), cu(ref['wl']), cu(ref['wr']), lib.make_tree(c['par']),
                                       cam_struct(lib, c['cam']), m, True)
    closs = (car * cu(ups[1])).sum() + (cal * cu(ups[2])).sum() + (cproj * cu(ups[3])).sum() * 1e-3
    if mode == 'pose_changes':
        closs = closs + (crr * cu(ups[0])).sum()
        assert_close('rel_rot', crr, ref['rr'])
    closs.backward()
    assert_close('abs_rot', car, ref['ar'])
    assert_close('abs_loc', cal, ref['al'])
    # pixel coordinates: joints near the image centre are small differences of numbers ~400, where the reference's own
    # fp32 evaluation is as far from float64 as the kernel's
    c32 = dict(c, loc=c['loc'].float(), rot=c['rot'].float(), wl=c['wl'].float(), wr=c['wr'].float())
    ref32 = _fk_reference(c32, mats.float(), mode, [u.float() for u in ups])
    assert_close_fp32_aware('proj', cproj, ref['proj'], ref32['proj'], slack=5.0)
    # (16 - 100 chained products of NON-rotations: the chain rule multiplies rounding like the values themselves)
    assert_close_fp32_aware('g_pose_in', x.grad, ref['grad'], ref32['grad'], slack=5.0)


def test_fk_tile_partial_upstream_and_no_projection(ops, lib):
    """Only some outputs carry an upstream gradient (the usual case: loc_3d alone, or rot_3d alone); no projection."""
    c = make_case(10, 16, 321, world=False)
    mats = O.rotation_6d_to_matrix(c['d6'])
    for which in ('loc', 'rot'):
        m64 = mats.clone().requires_grad_(True)
        _, rr, al, ar = O.abs_from_pose_changes(m64, c['loc'][c['sid']], c['rot'][c['sid']], c['par'])
        ((al ** 2).sum() if which == 'loc' else (ar ** 3).sum()).backward()
        x = cu(mats).requires_grad_(True)
        crr, car, cal, cproj = ops.fk_clip(x, cu(c['loc']), cu(c['rot']), c['sid'].int().to(DEV), None, None,
                                           lib.make_tree(c['par']), None, lib.MODE_POSE_CHANGES, False)
        assert cproj is None
        ((cal ** 2).sum() if which == 'loc' else (car ** 3).sum()).backward()
        assert_close('abs_loc', cal, al)
        assert_close('g_pose_in (%s only)' % which, x.grad, m64.grad)
