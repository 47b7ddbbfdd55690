"""Shared helpers for the parity tests (oracle in float64 on CPU vs the CUDA path in float32)."""
import numpy as np
import torch

import oracle as O

REL_TOL = 1e-4  # BASELINE.json north_star: 1e-4 relative in fp32 for locations, projections, loss, gradients


def scaled_err(got, ref):
    got = torch.as_tensor(np.asarray(got.detach().cpu() if torch.is_tensor(got) else got)).double()
    ref = torch.as_tensor(np.asarray(ref.detach().cpu() if torch.is_tensor(ref) else ref)).double()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    if ref.numel() == 0:
        return 0.0
    return float((got - ref).abs().max() / max(float(ref.abs().max()), 1e-12))


def assert_close(name, got, ref, tol=REL_TOL):
    """|got - ref| <= tol * max|ref| (scale-relative: fp32 cannot hold an elementwise 1e-4 on near-zero entries)."""
    e = scaled_err(got, ref)
    assert e <= tol, '%s: scaled max error %.3e > %.1e' % (name, e, tol)


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


def make_case(n, l, seed, par=None, n_types=4, world=True, drop=0.1, dtype=torch.float64):
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


def assert_close_fp32_aware(name, got, ref64, ref32, tol=REL_TOL, slack=3.0):
    """Gradients of the trunk bones are sums of large cancelling terms: the reference's own float32 evaluation
    deviates from float64 by ~1e-4 of the gradient scale there.  The CUDA path must be within `tol` of the float64
    oracle, or no further from it than `slack` x the float32 oracle (= the reference PyTorch path in fp32) is."""
    e = scaled_err(got, ref64)
    e32 = scaled_err(ref32, ref64)
    assert e <= max(tol, slack * e32), '%s: scaled max error %.3e > max(%.1e, %.1f x fp32-reference error %.3e)' % (
        name, e, tol, slack, e32)
