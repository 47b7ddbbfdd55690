"""Shared helpers for the parity tests (oracle in float64 on CPU vs the CUDA path in float32)."""
import numpy as np
import torch

import oracle as O

REL_TOL = 1e-4  # BASELINE.json north_star: 1e-4 relative in fp32 for locations, projections, loss, gradients
FP32_SLACK = 3.0  # gradient groups may be this many times further from fp64 than the fp32 reference evaluation is


def _pair(got, ref):
    got = torch.as_tensor(np.asarray(got.detach().cpu() if torch.is_tensor(got) else got)).double()
    ref = torch.as_tensor(np.asarray(ref.detach().cpu() if torch.is_tensor(ref) else ref)).double()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    return got, ref


def scaled_err(got, ref):
    """max|got - ref| / max|ref| over the WHOLE tensor (kept for scalars and bit-level property checks)."""
    got, ref = _pair(got, ref)
    if ref.numel() == 0:
        return 0.0
    return float((got - ref).abs().max() / max(float(ref.abs().max()), 1e-12))


def group_err(got, ref, keep):
    """max|got - ref| / max|ref| per group: both maxima over every axis NOT in `keep`.  Returns a tensor shaped like
    the kept axes (groups whose reference is identically zero report their absolute error)."""
    got, ref = _pair(got, ref)
    red = [a for a in range(ref.dim()) if a not in keep]
    err = (got - ref).abs()
    scale = ref.abs()
    if red:
        err, scale = err.amax(dim=red), scale.amax(dim=red)
    return err / scale.clamp_min(1e-300).where(scale > 0, torch.ones_like(scale))


def _layout(ref):
    """(channel axes, bone axis) of the path's tensors: (..., B, C) records or (..., B, 3, 3) matrices."""
    d = ref.dim()
    if d >= 3 and tuple(ref.shape[-2:]) == (3, 3):
        return (d - 2, d - 1), (d - 3 if d >= 4 else None)
    if d >= 2:
        return (d - 1,), (d - 2 if d >= 3 else None)
    return (), None


def close_report(got, ref, tol=REL_TOL, floor=1e-2):
    """Per-CHANNEL and per-BONE comparison (reference tests/walker_control/test_pose.py:31-52 compares bone by bone):
    every (bone, channel) group must satisfy  max|got - ref| <= tol * max(scale(bone, channel), floor * scale(channel)),
    scales being max|ref| over the group / over the whole channel.  Channels are never mixed (pixels ~400 next to
    1/z ~0.3 in projection_2d; a trunk bone's gradient next to a finger's), and `floor` only keeps a bone whose values
    are ~0 by construction (root / hips location, a lateral coordinate of the spine) from being held to 1e-4 of
    rounding noise.  Returns (worst ratio err / allowed, where)."""
    got, ref = _pair(got, ref)
    if ref.numel() == 0:
        return 0.0, None
    chan, bone = _layout(ref)
    err, mag = (got - ref).abs(), ref.abs()
    keep_bc = tuple(chan) + ((bone,) if bone is not None else ())
    red_bc = [a for a in range(ref.dim()) if a not in keep_bc]
    red_c = [a for a in range(ref.dim()) if a not in chan]
    e = err.amax(dim=red_bc, keepdim=True) if red_bc else err
    s_bc = mag.amax(dim=red_bc, keepdim=True) if red_bc else mag
    s_c = mag.amax(dim=red_c, keepdim=True) if red_c else mag
    allowed = tol * torch.maximum(s_bc, floor * s_c).clamp_min(1e-300)
    ratio = e / allowed
    worst = float(ratio.max())
    idx = np.unravel_index(int(ratio.argmax()), ratio.shape)
    return worst, tuple(int(i) for i in idx)


def assert_close(name, got, ref, tol=REL_TOL, floor=1e-2):
    """See close_report: tol relative to the scale of each (bone, channel) group, not of the whole tensor."""
    worst, where = close_report(got, ref, tol, floor)
    assert worst <= 1.0, '%s: error is %.2f x the allowed %.1e of the group scale at group %s' % (name, worst, tol, where)


def carla(dtype=torch.float64):
    loc, rot, names, par = O.reference_tensors(dtype)
    return loc, rot, names, par


def smpl_parents():
    structure = O.load_yaml('structure.yaml', files_dir=O.FILES_DIR + '/../../smpl/files')['structure']
    return O.parents_of(structure)


def random_rotations(shape, g, dtype=torch.float64, spread=0.3):
    base = torch.tensor([1., 0, 0, 0, 1, 0], dtype=dtype)
    d6 = torch.randn(tuple(shape) + (6,), generator=g, dtype=dtype) * spread + base
    return d6


MIN_HIPS_NECK_PX = 5.0


def _well_conditioned(d6, loc, rot, sid, par, cam, wl, wr, hips, neck, base, mode):
    """The hips/neck normaliser divides by the projected hips-neck distance (~60 px for an upright pedestrian).  A
    random pose whose torso points at the camera can bring it below a pixel: normalised coordinates in the hundreds, and
    every fp32 evaluation of that pose -- the reference's own included -- is rounding noise at the 1e-4 level (one such
    pose in 4096 dominated loss and gradient of a whole batch).  Parity at 1e-4 is only defined for conditioned inputs:
    clips that hold such a frame get their trunk bones (root, hips, spine, spine01) set upright."""
    for _ in range(4):
        with torch.no_grad():
            proj, _ = O.projection_module_forward(O.rotation_6d_to_matrix(d6), loc[sid], rot[sid], par, cam, wl, wr, mode)
        dist = (proj[..., neck, :2] - proj[..., hips, :2]).norm(dim=-1)
        bad = (dist < MIN_HIPS_NECK_PX).any(dim=1)
        if not bool(bad.any()):
            break
        if mode == 'pose_changes':
            d6[bad, :, 0:4] = base
        else:   # relative_rot: the 6-D vector IS the bone's rotation -- upright = the reference skeleton's own rotation
            d6[bad, :, 0:4] = O.matrix_to_rotation_6d(rot[sid][bad][:, 0:4]).unsqueeze(1).to(d6.dtype)
    return d6


def make_case(n, l, seed, par=None, n_types=4, world=True, drop=0.1, dtype=torch.float64, conditioned=True,
              mode='pose_changes'):
    """Synthetic step inputs per SURVEY 8d. Returns dict of float64 CPU tensors."""
    g = torch.Generator().manual_seed(seed)
    loc, rot, _, cpar = carla(dtype)
    if par is None:
        par = cpar
    b = len(par)
    if b != loc.shape[1]:  # other trees (SMPL): random skeleton table
        loc = torch.randn((n_types, b, 3), generator=g, dtype=dtype) * 0.2
        rot = O.rotation_6d_to_matrix(random_rotations((n_types, b), g, dtype))
    sid = torch.randint(0, n_types, (n,), generator=g)
    d6 = random_rotations((n, l, b), g, dtype)
    d6_gt = random_rotations((n, l, b), g, dtype)
    wl = wr = None
    if world:
        wl = torch.randn((n, l, 3), generator=g, dtype=dtype) * 0.02
        wr = O.euler_angles_to_matrix((torch.rand((n, l, 3), generator=g, dtype=dtype) - 0.5) * 0.1)
    cam = O.training_camera()
    hips, neck = (O.HIPS, O.NECK) if b == 26 else (0, min(12, b - 1))
    if conditioned and b == 26:
        base = torch.tensor([1., 0, 0, 0, 1, 0], dtype=dtype)
        d6 = _well_conditioned(d6, loc, rot, sid, par, cam, wl, wr, hips, neck, base, mode)
        d6_gt = _well_conditioned(d6_gt, loc, rot, sid, par, cam, wl, wr, hips, neck, base, mode)
    return dict(n=n, l=l, b=b, par=par, loc=loc, rot=rot, sid=sid, d6=d6, d6_gt=d6_gt, wl=wl, wr=wr, cam=cam,
                hips=hips, neck=neck, drop=drop, gen=g)


def oracle_step(c, pose, mode='pose_changes', normalize=True, mask=True, gt=None):
    """full_step generalised to any hips/neck (the oracle's defaults are the CARLA indices)."""
    proj, out = O.projection_module_forward(pose, c['loc'][c['sid']], c['rot'][c['sid']], c['par'], c['cam'],
                                            c['wl'], c['wr'], mode)
    proj_t = O.hips_neck_normalize(proj, c['hips'], c['neck']) if normalize else proj
    out.update(projection_2d=proj, projection_2d_transformed=proj_t)
    if gt is not None:
        l2 = O.loss_loc_2d(proj_t, gt[0], mask, c['hips'])
        l3 = O.loss_loc_3d(out['absolute_pose_loc'], gt[1])
        out.update(loc_2d=l2, loc_3d=l3, loc_2d_3d=l2 + l3)
    return out


def make_targets(c, mode='pose_changes', normalize=True):
    with torch.no_grad():
        t = oracle_step(c, O.rotation_6d_to_matrix(c['d6_gt']), mode, normalize)
    gt2 = t['projection_2d_transformed'].clone()
    drop = torch.rand(gt2.shape[:-1], generator=c['gen']) < c['drop']
    drop[..., c['hips']] = False
    gt2[drop] = 0
    return gt2, t['absolute_pose_loc'].clone()


def cast_case(c, dtype):
    d = dict(c)
    for k in ('loc', 'rot', 'd6', 'd6_gt', 'wl', 'wr'):
        if d.get(k) is not None:
            d[k] = d[k].to(dtype)
    return d


def assert_close_fp32_aware(name, got, ref64, ref32, tol=REL_TOL, slack=FP32_SLACK):
    """Gradients of the trunk bones are sums of large cancelling terms (all later frames x all descendants): the
    reference's own float32 evaluation deviates from float64 by ~1e-4 of the bone's gradient scale there.  Per bone and
    channel group (close_report), the CUDA path must be within `tol` of the float64 oracle, or no further from it than
    `slack` x the float32 oracle (= the reference PyTorch path in fp32) is IN THE SAME GROUP."""
    got, ref64 = _pair(got, ref64)
    ref32 = _pair(ref32, ref64)[0]
    chan, bone = _layout(ref64)
    keep = tuple(chan) + ((bone,) if bone is not None else ())
    red = [a for a in range(ref64.dim()) if a not in keep]
    red_c = [a for a in range(ref64.dim()) if a not in chan]
    mag = ref64.abs()
    s_bc = mag.amax(dim=red, keepdim=True) if red else mag
    s_c = mag.amax(dim=red_c, keepdim=True) if red_c else mag
    scale = torch.maximum(s_bc, 1e-2 * s_c).clamp_min(1e-300)
    e = ((got - ref64).abs().amax(dim=red, keepdim=True) if red else (got - ref64).abs()) / scale
    e32 = ((ref32 - ref64).abs().amax(dim=red, keepdim=True) if red else (ref32 - ref64).abs()) / scale
    bad = e > torch.maximum(torch.full_like(e, tol), slack * e32)
    if bool(bad.any()):
        i = np.unravel_index(int((e / torch.maximum(torch.full_like(e, tol), slack * e32)).argmax()), e.shape)
        raise AssertionError('%s: group %s: scaled error %.3e > max(%.1e, %.1f x fp32-reference error %.3e)' % (
            name, tuple(int(v) for v in i), float(e[i]), tol, slack, float(e32[i])))
