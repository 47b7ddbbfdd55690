"""The oracle against (a) the reference's golden YAML pair and (b) outputs of the reference's own code."""
import numpy as np
import pytest
import torch

import oracle as O


def T(a, dtype=torch.float32):
    return torch.as_tensor(np.asarray(a)).to(dtype)


def close(a, b, rtol=1e-4, atol=1e-5):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol)


def test_golden_yaml_pair():
    """reference tests/walker_control/test_pose.py:31-52: loc atol 1e-5 m, angle atol 1e-2 deg, root skipped."""
    loc, rot, names, par = O.reference_tensors(torch.float32)
    al, ar = O.relative_to_absolute(loc[:1], rot[:1], par)
    absy = O.yaml_to_pose(O.load_yaml('sk_female_absolute.yaml')['transforms'], is_abs=True)
    eul = -np.rad2deg(O.matrix_to_euler_angles(ar[0]).numpy())
    for i, n in enumerate(names):
        if i == 0:
            continue
        (x, y, z), (pitch, yaw, roll) = absy[n]
        assert np.allclose(al[0, i].numpy() * [1, 1, -1], [x, y, z], atol=1e-5)
        for got, want in ((eul[i, 0], roll), (eul[i, 1], pitch), (eul[i, 2], yaw)):
            assert abs((got - want + 180) % 360 - 180) < 1e-2


def test_tree_tables():
    _, _, names, par = O.reference_tensors()
    assert len(names) == 26 and names[O.HIPS] == 'crl_hips__C' and names[O.NECK] == 'crl_neck__C'
    assert par == [-1, 0, 1, 2, 3, 4, 5, 6, 3, 8, 9, 9, 3, 12, 13, 14, 1, 16, 17, 18, 19, 1, 21, 22, 23, 24]


def test_reference_tensors_match_reference_code(golden):
    loc, rot, _, par = O.reference_tensors(torch.float32)
    close(loc, golden['ref_rel_loc'], atol=1e-7)
    close(rot, golden['ref_rel_rot'], atol=1e-6)
    ident = torch.eye(3).reshape(1, 1, 3, 3).repeat(4, 26, 1, 1)
    al, ar, _ = O.p3d_pose_forward(ident, loc, rot, par)
    close(al, golden['ref_abs_loc'], atol=1e-6)
    close(ar, golden['ref_abs_rot'], atol=1e-6)
    proj = O.project(al, torch.zeros(4, 3), torch.eye(3)[None].repeat(4, 1, 1), O.reference_camera())
    close(proj, golden['ref_projections'], atol=1e-3)


def test_p3d_pose_forward_matches_reference_code(golden):
    loc, rot, _, par = O.reference_tensors(torch.float32)
    ch = T(golden['fwd_changes'])
    n = ch.shape[0]
    al, ar, rr = O.p3d_pose_forward(ch, loc[:1].repeat(n, 1, 1), rot[:1].repeat(n, 1, 1, 1), par)
    close(al, golden['fwd_abs_loc'], atol=1e-6)
    close(ar, golden['fwd_abs_rot'], atol=1e-6)
    close(rr, golden['fwd_rel_rot'], atol=1e-6)


def test_world_scan_matches_reference_code(golden):
    wl, wr = O.calculate_world_from_changes((3, 7), T(golden['world_loc_change']), T(golden['world_rot_change']))
    close(wl, golden['world_loc'], atol=1e-6)
    close(wr, golden['world_rot'], atol=1e-6)
    # early-out branch, utils/world.py:32-37 (+ reference tests/utils/test_world.py invariants)
    wl, wr = O.calculate_world_from_changes((2, 5))
    assert wl.shape == (2, 5, 3) and wr.shape == (2, 5, 3, 3)
    assert torch.equal(wl, torch.zeros(2, 5, 3)) and torch.equal(wr, torch.eye(3).expand(2, 5, 3, 3))


def test_projection_matches_reference_code(golden):
    cam = O.training_camera()
    close(cam.R, golden['camera_R'], atol=1e-7)
    close(cam.T, golden['camera_T'], atol=1e-6)
    out = O.project(T(golden['proj_x']), T(golden['proj_loc']), T(golden['proj_rot']), cam)
    close(out, golden['proj_out'], rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize('mode', ['pose_changes', 'relative_rot'])
@pytest.mark.parametrize('dtype', [torch.float32, torch.float64])
def test_full_step_matches_reference_code(golden, mode, dtype):
    """ProjectionModule + Normalizer + loc_2d/loc_3d/loc_2d_3d forward AND autograd gradients."""
    k = mode + '/'
    loc, rot, _, par = O.reference_tensors(dtype)
    sid = torch.as_tensor(golden['skeleton_id']).long()
    d6 = T(golden[k + 'd6'], dtype).requires_grad_(True)
    mats = O.rotation_6d_to_matrix(d6)
    mats.retain_grad()
    out = O.full_step(mats, loc[sid], rot[sid], par, O.training_camera(),
                      T(golden[k + 'gt_projection_2d_transformed'], dtype), T(golden[k + 'gt_absolute_pose_loc'], dtype),
                      T(golden[k + 'world_loc_change'], dtype), T(golden[k + 'world_rot_change'], dtype), mode=mode)
    out['loc_2d_3d'].backward()
    close(mats.detach(), golden[k + 'mats'], atol=1e-5)
    for key in ('relative_pose_loc', 'relative_pose_rot', 'absolute_pose_loc', 'absolute_pose_rot', 'world_loc',
                'world_rot'):
        close(out[key].detach(), golden[k + key], atol=5e-6)
    close(out['projection_2d'].detach(), golden[k + 'projection_2d'], rtol=1e-5, atol=2e-3)
    close(out['projection_2d_transformed'].detach(), golden[k + 'projection_2d_transformed'], rtol=1e-4, atol=1e-4)
    for key in ('loc_2d', 'loc_3d', 'loc_2d_3d'):
        close(out[key].detach(), golden[k + key], rtol=1e-4, atol=1e-6)
    gm, gd = golden[k + 'grad_mats'], golden[k + 'grad_d6']
    close(mats.grad, gm, rtol=1e-3, atol=1e-4 * np.abs(gm).max())
    close(d6.grad, gd, rtol=1e-3, atol=1e-4 * np.abs(gd).max())


def test_euler_matches_scipy():
    from scipy.spatial.transform import Rotation
    e = torch.rand(64, 3, dtype=torch.float64) * 2 - 1
    m = O.euler_angles_to_matrix(e, 'XYZ').numpy()
    close(m, Rotation.from_euler('XYZ', e.numpy()).as_matrix(), atol=1e-12)
    close(O.matrix_to_euler_angles(torch.as_tensor(m)), e, atol=1e-10)


def test_rotation_6d_roundtrip():
    d6 = torch.randn(128, 6, dtype=torch.float64)
    m = O.rotation_6d_to_matrix(d6)
    close(m @ m.transpose(-1, -2), torch.eye(3, dtype=torch.float64).expand(128, 3, 3), atol=1e-12)
    close(torch.linalg.det(m), torch.ones(128), atol=1e-12)
    close(O.rotation_6d_to_matrix(O.matrix_to_rotation_6d(m)), m, atol=1e-12)


def test_oracle_gradient_finite_differences():
    """the oracle's autograd is the gradient reference -> check it with float64 central differences."""
    g = torch.Generator().manual_seed(1)
    dt = torch.float64
    loc, rot, _, par = O.reference_tensors(dt)
    N, L = 2, 3
    d6 = (torch.randn((N, L, 26, 6), generator=g, dtype=dt) * 0.2 + torch.tensor([1., 0, 0, 0, 1, 0], dtype=dt))
    gt2 = torch.randn((N, L, 26, 3), generator=g, dtype=dt)
    gt2[0, 1, 5] = 0
    gt3 = torch.randn((N, L, 26, 3), generator=g, dtype=dt)
    cam = O.training_camera()

    def f(x):
        return O.full_step(O.rotation_6d_to_matrix(x), loc[:N], rot[:N], par, cam, gt2, gt3)['loc_2d_3d']
    x = d6.clone().requires_grad_(True)
    f(x).backward()
    idx = [(0, 0, 1, 2), (1, 2, 8, 4), (0, 1, 20, 0), (1, 0, 3, 5), (0, 2, 0, 1)]
    for i in idx:
        h = 1e-6
        xp, xm = d6.clone(), d6.clone()
        xp[i] += h
        xm[i] -= h
        fd = (f(xp) - f(xm)) / (2 * h)
        assert abs(float(fd) - float(x.grad[i])) <= 1e-6 * max(1.0, abs(float(fd)))


def test_generator_stream_known_answers():
    """The device generator's stream is Philox4x32-10 (Salmon et al., Random123): its published known-answer vectors
    pin the oracle's restatement, which in turn pins the kernel bit for bit (tests/test_gpu_widen.py)."""
    import numpy as np
    kat = (((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)))
    for ctr, key, want in kat:
        got = O.philox4x32_10(*[[c] for c in ctr], *key)
        assert tuple(int(g[0]) for g in got) == want
    e, w, s = O.draw_sparse_changes(22742, 3, 40, 7, 26, 3, 0.0873, 0.03, 0.5)
    assert ((np.abs(e).sum(-1) > 0).sum(-1) == 3).all()            # exactly 3 distinct bones per frame
    assert float(np.abs(e).max()) <= 0.0873 and (w[..., :2] == 0).all()
    assert float(np.abs(w[:, 1:, 2]).max()) <= 0.03 and float(np.abs(w[:, 0, 2]).max()) <= 0.5
    assert set(s.tolist()) <= {0, 1, 2, 3} and len(set(s.tolist())) == 4
    # every bone gets chosen, about uniformly (40*7*3 = 840 picks over 26 bones)
    counts = (np.abs(e).sum(-1) > 0).sum((0, 1))
    assert counts.min() > 10 and counts.max() < 60
