"""torch.autograd.Function wrappers over the C ABI (include/p2c_b200.h).

PyTorch here is plumbing only (device memory, streams, autograd graph); every forward and
backward below is a hand-written sm_100a kernel.  There is no CPU fallback: CPU tensors raise.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check, ptr, stream


def _f32c(t, name):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError('p2c_b200: %s must be a CUDA tensor (no CPU fallback); got %s' % (name, t.device))
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


class _Rot6dToMatrix(torch.autograd.Function):
    """pytorch3d rotation_6d_to_matrix at modules/movements/movements.py:113-116."""

    @staticmethod
    def forward(ctx, d6):
        d6 = _f32c(d6, 'd6')
        out = torch.empty(d6.shape[:-1] + (3, 3), device=d6.device, dtype=torch.float32)
        n = d6.numel() // 6
        check(_lib.load().p2c_rot6d_to_matrix_fwd(ptr(d6), ptr(out), n, stream()), 'rot6d_to_matrix_fwd')
        ctx.save_for_backward(d6)
        return out

    @staticmethod
    def backward(ctx, g):
        (d6,) = ctx.saved_tensors
        g = _f32c(g, 'grad')
        gd = torch.empty_like(d6)
        check(_lib.load().p2c_rot6d_to_matrix_bwd(ptr(d6), ptr(g), ptr(gd), d6.numel() // 6, stream()),
              'rot6d_to_matrix_bwd')
        return gd


def rotation_6d_to_matrix(d6):
    if d6.shape[-1] != 6:
        raise RuntimeError('rotation_6d_to_matrix expects (..., 6)')
    return _Rot6dToMatrix.apply(d6)


class _QuaternionToMatrix(torch.autograd.Function):
    """pytorch3d quaternion_to_matrix (real part first; the quaternion need not be unit: M = I + 2/|q|^2 A(q))."""

    @staticmethod
    def forward(ctx, q):
        q = _f32c(q, 'quaternions')
        out = torch.empty(q.shape[:-1] + (3, 3), device=q.device, dtype=torch.float32)
        check(_lib.load().p2c_quaternion_to_matrix_fwd(ptr(q), ptr(out), q.numel() // 4, stream()),
              'quaternion_to_matrix_fwd')
        ctx.save_for_backward(q)
        return out

    @staticmethod
    def backward(ctx, g):
        (q,) = ctx.saved_tensors
        g = _f32c(g, 'grad')
        gq = torch.empty_like(q)
        check(_lib.load().p2c_quaternion_to_matrix_bwd(ptr(q), ptr(g), ptr(gq), q.numel() // 4, stream()),
              'quaternion_to_matrix_bwd')
        return gq


def quaternion_to_matrix(quaternions):
    """(..., 4) quaternions, real part first -> (..., 3, 3) rotation matrices acting like the rest of the path."""
    if quaternions.shape[-1] != 4:
        raise RuntimeError('quaternion_to_matrix expects (..., 4)')
    return _QuaternionToMatrix.apply(quaternions)


def matrix_to_rotation_6d(matrix):
    """pytorch3d matrix_to_rotation_6d (modules/movements/seq2seq.py:330): first two rows; a view + copy."""
    return matrix[..., :2, :].clone().reshape(matrix.shape[:-2] + (6,))


def euler_angles_to_matrix(euler, convention='XYZ'):
    """pytorch3d euler_angles_to_matrix for the only convention the reference uses ("XYZ")."""
    if convention != 'XYZ':
        raise RuntimeError('only the "XYZ" convention is implemented (the only one the reference uses)')
    e = _f32c(euler.detach(), 'euler')
    out = torch.empty(e.shape[:-1] + (3, 3), device=e.device, dtype=torch.float32)
    check(_lib.load().p2c_euler_xyz_to_matrix(ptr(e), ptr(out), e.numel() // 3, stream()), 'euler_xyz_to_matrix')
    return out


def matrix_to_euler_angles(matrix, convention='XYZ'):
    if convention != 'XYZ':
        raise RuntimeError('only the "XYZ" convention is implemented (the only one the reference uses)')
    m = _f32c(matrix.detach(), 'matrix')
    out = torch.empty(m.shape[:-2] + (3,), device=m.device, dtype=torch.float32)
    check(_lib.load().p2c_matrix_to_euler_xyz(ptr(m), ptr(out), m.numel() // 9, stream()), 'matrix_to_euler_xyz')
    return out


class _FkClip(torch.autograd.Function):
    """P3dPose.forward / relative_to_absolute over (N, L) with optional fused projection."""

    @staticmethod
    def forward(ctx, pose_in, ref_loc, ref_rot, ref_index, world_loc, world_rot, tree, camera, mode, want_proj):
        pose_in = _f32c(pose_in, 'pose_in')
        ref_loc = _f32c(ref_loc, 'ref_loc')
        ref_rot = _f32c(ref_rot, 'ref_rot')
        world_loc = _f32c(world_loc, 'world_loc')
        world_rot = _f32c(world_rot, 'world_rot')
        N, L, B = pose_in.shape[:3]
        dev = pose_in.device
        rel_rot = torch.empty((N, L, B, 3, 3), device=dev) if mode == _lib.MODE_POSE_CHANGES else None
        abs_rot = torch.empty((N, L, B, 3, 3), device=dev)
        abs_loc = torch.empty((N, L, B, 3), device=dev)
        proj = torch.empty((N, L, B, 3), device=dev) if want_proj else None
        a = _lib.FkFwd()
        a.tree, a.n_clips, a.clip_len, a.mode = tree, N, L, mode
        a.pose_in, a.ref_loc, a.ref_rot, a.ref_index = ptr(pose_in), ptr(ref_loc), ptr(ref_rot), ptr(ref_index)
        a.world_loc, a.world_rot = ptr(world_loc), ptr(world_rot)
        if camera is not None:
            a.camera = camera
        a.rel_rot, a.abs_rot, a.abs_loc, a.proj = ptr(rel_rot), ptr(abs_rot), ptr(abs_loc), ptr(proj)
        check(_lib.load().p2c_fk_fwd(a, stream()), 'fk_fwd')
        ctx.save_for_backward(pose_in, ref_loc, ref_rot, ref_index, world_loc, world_rot, rel_rot, abs_rot, abs_loc)
        ctx.tree, ctx.camera, ctx.mode, ctx.want_proj = tree, camera, mode, want_proj
        ctx.set_materialize_grads(False)   # outputs nobody differentiates stay None (no zero tensors for the kernel to read)
        if rel_rot is None:
            rel_rot = pose_in.new_empty(0)
        if proj is None:
            proj = pose_in.new_empty(0)
        ctx.mark_non_differentiable(*[t for t, w in ((rel_rot, mode != _lib.MODE_POSE_CHANGES), (proj, not want_proj)) if w])
        return rel_rot, abs_rot, abs_loc, proj

    @staticmethod
    def backward(ctx, g_rel_rot, g_abs_rot, g_abs_loc, g_proj):
        pose_in, ref_loc, ref_rot, ref_index, world_loc, world_rot, rel_rot, abs_rot, abs_loc = ctx.saved_tensors
        N, L, B = pose_in.shape[:3]
        dev = pose_in.device
        need = ctx.needs_input_grad
        if ctx.mode != _lib.MODE_POSE_CHANGES:
            g_rel_rot = None
        if not ctx.want_proj:
            g_proj = None
        a = _lib.FkBwd()
        a.tree, a.n_clips, a.clip_len, a.mode = ctx.tree, N, L, ctx.mode
        a.pose_in, a.ref_loc, a.ref_rot, a.ref_index = ptr(pose_in), ptr(ref_loc), ptr(ref_rot), ptr(ref_index)
        a.world_loc, a.world_rot = ptr(world_loc), ptr(world_rot)
        if ctx.camera is not None:
            a.camera = ctx.camera
        a.rel_rot, a.abs_rot, a.abs_loc = ptr(rel_rot), ptr(abs_rot), ptr(abs_loc)
        g_rel_rot, g_abs_rot = _f32c(g_rel_rot, 'g_rel_rot'), _f32c(g_abs_rot, 'g_abs_rot')
        g_abs_loc, g_proj = _f32c(g_abs_loc, 'g_abs_loc'), _f32c(g_proj, 'g_proj')
        a.g_rel_rot, a.g_abs_rot, a.g_abs_loc, a.g_proj = ptr(g_rel_rot), ptr(g_abs_rot), ptr(g_abs_loc), ptr(g_proj)
        g_in = torch.empty_like(pose_in)
        a.g_pose_in = ptr(g_in)
        g_ref_loc = g_ref_rot = g_wl = g_wr = None
        if need[1]:
            if ref_index is not None:
                raise RuntimeError('gradient w.r.t. an indexed reference table is not supported')
            g_ref_loc = torch.empty_like(ref_loc)
            a.g_ref_loc = ptr(g_ref_loc)
        if need[2] and ref_rot is not None and ctx.mode == _lib.MODE_POSE_CHANGES:
            if ref_index is not None:
                raise RuntimeError('gradient w.r.t. an indexed reference table is not supported')
            g_ref_rot = torch.empty_like(ref_rot)
            a.g_ref_rot = ptr(g_ref_rot)
        if g_proj is not None and world_loc is not None and need[4]:
            g_wl = torch.empty_like(world_loc)
            a.g_world_loc = ptr(g_wl)
        if g_proj is not None and world_rot is not None and need[5]:
            g_wr = torch.empty_like(world_rot)
            a.g_world_rot = ptr(g_wr)
        check(_lib.load().p2c_fk_bwd(a, stream()), 'fk_bwd')
        return g_in, g_ref_loc, g_ref_rot, None, g_wl, g_wr, None, None, None, None


def fk_clip(pose_in, ref_loc, ref_rot, ref_index, world_loc, world_rot, tree, camera, mode, want_proj):
    rel_rot, abs_rot, abs_loc, proj = _FkClip.apply(pose_in, ref_loc, ref_rot, ref_index, world_loc, world_rot, tree,
                                                    camera, mode, want_proj)
    return (rel_rot if mode == _lib.MODE_POSE_CHANGES else None), abs_rot, abs_loc, (proj if want_proj else None)


class _Project(torch.autograd.Function):
    """P3dPoseProjection.forward (walker_control/p3d_pose_projection.py:115-152)."""

    @staticmethod
    def forward(ctx, x, loc, rot, camera):
        x, loc, rot = _f32c(x, 'x'), _f32c(loc, 'loc'), _f32c(rot, 'rot')
        out = torch.empty_like(x)
        m, j = x.shape[0], x.shape[1]
        check(_lib.load().p2c_project_fwd(ptr(x), ptr(loc), ptr(rot), camera, ptr(out), m, j, stream()), 'project_fwd')
        ctx.save_for_backward(x, loc, rot)
        ctx.camera = camera
        return out

    @staticmethod
    def backward(ctx, g):
        x, loc, rot = ctx.saved_tensors
        g = _f32c(g, 'grad')
        need = ctx.needs_input_grad
        gx = torch.empty_like(x) if need[0] else None
        gl = torch.empty_like(loc) if need[1] else None
        gr = torch.empty_like(rot) if need[2] else None
        check(_lib.load().p2c_project_bwd(ptr(x), ptr(loc), ptr(rot), ctx.camera, ptr(g), ptr(gx), ptr(gl), ptr(gr),
                                          x.shape[0], x.shape[1], stream()), 'project_bwd')
        return gx, gl, gr, None


def project(x, loc, rot, camera):
    return _Project.apply(x, loc, rot, camera)


class _WorldScan(torch.autograd.Function):
    """utils/world.py:16-63 (the cumulative part)."""

    @staticmethod
    def forward(ctx, loc_change, rot_change, init_loc, init_rot):
        loc_change, rot_change = _f32c(loc_change, 'world_loc_change'), _f32c(rot_change, 'world_rot_change')
        init_loc, init_rot = _f32c(init_loc, 'initial_world_loc'), _f32c(init_rot, 'initial_world_rot')
        n, l = loc_change.shape[:2]
        wl = torch.empty((n, l, 3), device=loc_change.device)
        wr = torch.empty((n, l, 3, 3), device=loc_change.device)
        check(_lib.load().p2c_world_scan_fwd(ptr(loc_change), ptr(rot_change), ptr(init_loc), ptr(init_rot), ptr(wl),
                                             ptr(wr), n, l, stream()), 'world_scan_fwd')
        ctx.save_for_backward(rot_change, wr, init_rot)
        return wl, wr

    @staticmethod
    def backward(ctx, g_wl, g_wr):
        rot_change, wr, init_rot = ctx.saved_tensors
        n, l = rot_change.shape[:2]
        need = ctx.needs_input_grad
        g_wl, g_wr = _f32c(g_wl, 'g_world_loc'), _f32c(g_wr, 'g_world_rot')
        g_dl = torch.empty((n, l, 3), device=wr.device) if need[0] else None
        g_dr = torch.empty_like(rot_change) if need[1] else None
        check(_lib.load().p2c_world_scan_bwd(ptr(rot_change), ptr(wr), ptr(init_rot), ptr(g_wl), ptr(g_wr), ptr(g_dl),
                                             ptr(g_dr), n, l, stream()), 'world_scan_bwd')
        return g_dl, g_dr, None, None


def world_scan(loc_change, rot_change, init_loc=None, init_rot=None):
    return _WorldScan.apply(loc_change, rot_change, init_loc, init_rot)


class _HipsNeckNormalize(torch.autograd.Function):
    """Normalizer(HipsNeckExtractor) (transforms/pose/normalization/normalizer.py:20-41)."""

    @staticmethod
    def forward(ctx, sample, hips, neck, near_zero):
        s = _f32c(sample, 'sample')
        out = torch.empty_like(s)
        j, c = s.shape[-2], s.shape[-1]
        m = s.numel() // (j * c)
        check(_lib.load().p2c_hips_neck_normalize_fwd(ptr(s), ptr(out), m, j, c, hips, neck, near_zero, stream()),
              'hips_neck_normalize_fwd')
        ctx.save_for_backward(s)
        ctx.cfg = (hips, neck, near_zero)
        return out

    @staticmethod
    def backward(ctx, g):
        (s,) = ctx.saved_tensors
        g = _f32c(g, 'grad')
        hips, neck, near_zero = ctx.cfg
        gs = torch.empty_like(s)
        j, c = s.shape[-2], s.shape[-1]
        check(_lib.load().p2c_hips_neck_normalize_bwd(ptr(s), ptr(g), ptr(gs), s.numel() // (j * c), j, c, hips, neck,
                                                      near_zero, stream()), 'hips_neck_normalize_bwd')
        return gs, None, None, None


def hips_neck_normalize(sample, hips, neck, near_zero=1e-5):
    return _HipsNeckNormalize.apply(sample, hips, neck, near_zero)


def normalize_extractor(sample, hips, neck, extractor, near_zero=1e-5):
    """Normalizer(BBoxExtractor) (extractor 1) / Normalizer(HipsNeckBBoxFallbackExtractor) (extractor 2) over
    (..., J, 2|3) samples; forward only (dataset inputs carry no gradient)."""
    s = _f32c(sample.detach(), 'sample')
    out = torch.empty_like(s)
    j, c = s.shape[-2], s.shape[-1]
    check(_lib.load().p2c_normalize_extractor_fwd(ptr(s), ptr(out), s.numel() // (j * c), j, c, hips, neck, extractor,
                                                  near_zero, stream()), 'normalize_extractor_fwd')
    return out


def extractor_shift_scale(sample, dim, hips, neck, extractor=0, near_zero=1e-5):
    """Extractor.get_shift_scale (extractor.py:23-36) of (..., J, C) samples: shift (..., dim), scale (...,).
    extractor 0 = hips/neck, 1 = bbox, 2 = hips_neck_bbox.  No gradient (reported values)."""
    s = _f32c(sample.detach(), 'sample')
    j, c = s.shape[-2], s.shape[-1]
    lead = s.shape[:-2]
    m = s.numel() // (j * c) if s.numel() else 0
    shift = torch.empty(lead + (dim,), device=s.device)
    scale = torch.empty(lead, device=s.device)
    check(_lib.load().p2c_extractor_shift_scale(ptr(s), ptr(shift), ptr(scale), m, j, c, dim, hips, neck, extractor,
                                                near_zero, stream()), 'extractor_shift_scale')
    return shift, scale


class _MaskedMSE(torch.autograd.Function):
    """nn.MSELoss('mean') over [..., :c_used] of the joints kept by utils/tensors.py:29-40."""

    @staticmethod
    def forward(ctx, pred, gt, c_used, mask_hips):
        pred, gt = _f32c(pred, 'pred'), _f32c(gt.detach(), 'gt')
        j, cp, cg = pred.shape[-2], pred.shape[-1], gt.shape[-1]
        m = pred.numel() // (j * cp)
        if gt.numel() // (j * cg) != m or gt.shape[-2] != j:
            raise RuntimeError('pred %s and gt %s do not match' % (tuple(pred.shape), tuple(gt.shape)))
        lib = _lib.load()
        ws = torch.empty(lib.p2c_mse_workspace_floats(m), device=pred.device)
        out = torch.empty(3, device=pred.device)
        check(lib.p2c_masked_mse_fwd(ptr(pred), ptr(gt), m, j, cp, cg, c_used, mask_hips, ptr(ws), ptr(out), stream()),
              'masked_mse_fwd')
        ctx.save_for_backward(pred, gt, out)
        ctx.cfg = (c_used, mask_hips)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        pred, gt, out = ctx.saved_tensors
        c_used, mask_hips = ctx.cfg
        g = _f32c(g, 'grad').reshape(1)
        gp = torch.empty_like(pred)
        j, cp, cg = pred.shape[-2], pred.shape[-1], gt.shape[-1]
        check(_lib.load().p2c_masked_mse_bwd(ptr(pred), ptr(gt), ptr(out), ptr(g), ptr(gp), pred.numel() // (j * cp), j,
                                             cp, cg, c_used, mask_hips, stream()), 'masked_mse_bwd')
        return gp, None, None, None


def masked_mse(pred, gt, c_used, mask_hips=-1):
    """mask_hips < 0: plain MSE; otherwise joints whose gt[..., :c_used] has a zero are dropped, hips kept."""
    return _MaskedMSE.apply(pred, gt, c_used, mask_hips)


class _Loc2dStep(torch.autograd.Function):
    """Loc2DPoseLoss forward + its gradient in one cooperative launch (p2c_loc2d_step): the autoencoder flow's step
    (modules/flow/autoencoder.py:102-112, loss/loc_2d.py:69-89).  backward only scales the stored gradient."""

    @staticmethod
    def forward(ctx, pred, gt, mask_hips, exchange, out):
        lib = _lib.load()
        j = pred.shape[-2]
        m = pred.numel() // (j * 2)
        out = out or {}
        g = out.get('g_pred') if out.get('g_pred') is not None else torch.empty_like(pred)
        ws = out.get('workspace')
        if ws is None:
            ws = _loc2d_workspace(pred.device)
        res = out.get('losses') if out.get('losses') is not None else torch.empty(4, device=pred.device)
        xb, xr, xw = (C.c_void_p(exchange.block), exchange.rank, exchange.world) if exchange is not None else (None, 0, 0)
        check(lib.p2c_loc2d_step(ptr(pred), ptr(gt), m, j, gt.shape[-1], mask_hips, 1.0, ptr(g), ptr(ws), ptr(res), xb, xr,
                                 xw, stream()), 'loc2d_step')
        if exchange is not None:
            exchange.steps += 1
        ctx.save_for_backward(g)
        return res[0].clone(), res

    @staticmethod
    def backward(ctx, g_loss, _):
        (g,) = ctx.saved_tensors
        return g * g_loss, None, None, None, None


_LOC2D_WS = {}


def _loc2d_workspace(dev):
    key = (dev.type, dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _LOC2D_WS.get(key)
    if ws is None:
        ws = torch.zeros(int(_lib.load().p2c_loc2d_step_workspace_floats()), device=dev)
        _LOC2D_WS[key] = ws
    return ws


def loc2d_loss(pred, gt, mask_hips=-1, exchange=None, out=None, with_sums=False):
    """loc_2d of (..., J, >=2) predictions against (..., J, >=2) targets over the first two channels (mask_hips < 0: no
    missing-joints mask).  Contiguous 2-channel predictions that want a gradient take the one-launch fused
    forward + backward (p2c_loc2d_step); everything else the two-stage reduction with its separate backward."""
    p = pred
    fused = (p.is_cuda and p.dtype == torch.float32 and p.shape[-1] == 2 and p.is_contiguous() and p.requires_grad
             and p.numel() > 0 and p.data_ptr() % 16 == 0)
    if fused:
        g = _gt2_records(gt.detach())
        if g.shape[-1] not in (2, 3) or g.data_ptr() % (16 if g.shape[-1] == 2 else 8) != 0:
            fused = False
    if fused:
        j = p.shape[-2]
        fused = bool(_lib.load().p2c_loc2d_step_supported(p.numel() // (j * 2), j, g.shape[-1]))
    if not fused:
        loss = masked_mse(pred, gt, 2, mask_hips)
        return (loss, None) if with_sums else loss
    loss, sums = _Loc2dStep.apply(p, g, mask_hips, exchange, out)
    return (loss, sums) if with_sums else loss


class _CumChanges(torch.autograd.Function):
    """Per-bone cumulative product of pose changes over frames (loss/cum_pose_changes.py:26-48)."""

    @staticmethod
    def forward(ctx, changes):
        c = _f32c(changes, 'pose_changes')
        if c.dim() != 5 or c.shape[-2:] != (3, 3):
            raise RuntimeError('cumulative_changes expects (N, L, B, 3, 3), got %s' % (tuple(c.shape),))
        out = torch.empty_like(c)
        n, l, b = c.shape[:3]
        check(_lib.load().p2c_cum_changes_fwd(ptr(c), ptr(out), n, l, b, stream()), 'cum_changes_fwd')
        ctx.save_for_backward(c, out)
        return out

    @staticmethod
    def backward(ctx, g):
        c, out = ctx.saved_tensors
        g = _f32c(g, 'grad')
        gc = torch.empty_like(c)
        n, l, b = c.shape[:3]
        check(_lib.load().p2c_cum_changes_bwd(ptr(c), ptr(out), ptr(g), ptr(gc), n, l, b, stream()),
              'cum_changes_bwd')
        return gc


def cumulative_changes(changes):
    return _CumChanges.apply(changes)


class _Renormalize(torch.autograd.Function):
    """Normalizer(dim) and/or DeNormalizer.from_reference per clip (normalizer.py:20-41, denormalizer.py:13-37,
    reference_skeletons_denormalizer.py:41-91) in one pass."""

    @staticmethod
    def forward(ctx, sample, dim, hips, neck, autonormalize, near_zero, ref_shift, ref_scale, ref_index):
        s = _f32c(sample, 'sample')
        if s.dim() < 3:
            raise RuntimeError('renormalize expects (..., L, J, C), got %s' % (tuple(s.shape),))
        out = torch.empty_like(s)
        l, j, c = s.shape[-3:]
        m = s.numel() // (j * c)
        ref_shift, ref_scale = _f32c(ref_shift, 'ref_shift'), _f32c(ref_scale, 'ref_scale')
        if ref_scale is not None:
            rows = ref_scale.numel()
            if ref_shift.numel() != rows * dim:
                raise RuntimeError('ref_shift %s does not match ref_scale %s for dim %d'
                                   % (tuple(ref_shift.shape), tuple(ref_scale.shape), dim))
            if ref_index is None and rows != m // l:
                raise RuntimeError('per-clip reference rows (%d) do not match the clips (%d)' % (rows, m // l))
            if ref_index is not None and ref_index.numel() != m // l:
                raise RuntimeError('ref_index has %d entries for %d clips' % (ref_index.numel(), m // l))
        check(_lib.load().p2c_renormalize_fwd(ptr(s), ptr(out), m, j, c, dim, hips, neck, 1 if autonormalize else 0,
                                              near_zero, ptr(ref_shift), ptr(ref_scale), ptr(ref_index), l, stream()),
              'renormalize_fwd')
        ctx.save_for_backward(s, ref_scale, ref_index)
        ctx.cfg = (dim, hips, neck, autonormalize, near_zero)
        return out

    @staticmethod
    def backward(ctx, g):
        s, ref_scale, ref_index = ctx.saved_tensors
        dim, hips, neck, autonormalize, near_zero = ctx.cfg
        g = _f32c(g, 'grad')
        gs = torch.empty_like(s)
        l, j, c = s.shape[-3:]
        check(_lib.load().p2c_renormalize_bwd(ptr(s), ptr(g), ptr(gs), s.numel() // (j * c), j, c, dim, hips, neck,
                                              1 if autonormalize else 0, near_zero, ptr(ref_scale), ptr(ref_index), l,
                                              stream()), 'renormalize_bwd')
        return gs, None, None, None, None, None, None, None, None


def renormalize(sample, dim, hips, neck, autonormalize=True, near_zero=1e-5, ref_shift=None, ref_scale=None,
                ref_index=None):
    """sample (N, L, J, C).  autonormalize: hips -> 0, |neck - hips| -> 1 over the first `dim` channels; ref_shift
    (rows, dim) / ref_scale (rows): then `y * scale + shift` with the row of each clip (ref_index (N,) int32, or one
    row per clip).  The reference tables carry no gradient (they are constants of the reference skeletons)."""
    return _Renormalize.apply(sample, dim, hips, neck, autonormalize, near_zero,
                              None if ref_shift is None else ref_shift.detach(),
                              None if ref_scale is None else ref_scale.detach(), ref_index)


def draw_sparse_changes(seed, offset, n_clips, clip_len, n_bones, changes_per_frame, max_change, max_world_change=0.0,
                        max_initial_world_change=0.0, n_types=4, device=None):
    """The synthetic generator's random draws on the device (p2c_draw_sparse_changes; reference
    carla_2d3d_dataset.py:157-182): Euler changes (N,L,B,3), world Euler changes (N,L,3), skeleton type (N,) int32.
    Counter-based: (seed, offset) name the batch, nothing is kept between calls."""
    dev = torch.device(device if device is not None else 'cuda')
    if dev.type != 'cuda':
        raise RuntimeError('p2c_b200: the generator draws on a CUDA device (no CPU fallback)')
    euler = torch.empty((n_clips, clip_len, n_bones, 3), device=dev)
    world = torch.empty((n_clips, clip_len, 3), device=dev)
    skeleton = torch.empty((n_clips,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().p2c_draw_sparse_changes(int(seed) & (2 ** 64 - 1), int(offset) & (2 ** 64 - 1), n_clips, clip_len,
                                                  n_bones, changes_per_frame, float(max_change), float(max_world_change),
                                                  float(max_initial_world_change), n_types, ptr(euler), ptr(world),
                                                  ptr(skeleton), stream()), 'draw_sparse_changes')
    return euler, world, skeleton


_WS = {}


def _workspace(dev, nfloats):
    """Per-device scratch for p2c_fused_step.  Its first 16 floats are counters that must be zero before the first
    call and that every call leaves zero again (include/p2c_b200.h), so the buffer is zero-filled once and reused.
    Calls on one device are stream-ordered by PyTorch's current stream."""
    key = (dev.type, dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() < nfloats:
        ws = torch.zeros(max(nfloats, 1), device=dev)
        _WS[key] = ws
    return ws


class FusedStepResult(object):
    """Outputs of one fused pose_lifting step (see fused_step)."""
    __slots__ = ('loc_2d', 'loc_3d', 'loc_2d_3d', 'rot_3d', 'total', 'losses', 'relative_pose_rot', 'absolute_pose_rot',
                 'absolute_pose_loc', 'projection_2d', 'projection_2d_transformed', 'pose_in', 'grad_pose_in')

    def backward(self):
        """Same effect as `self.loc_2d_3d.backward()`, without a pass through the autograd engine for the loss node:
        the kernel has already left d(loc_2d_3d)/d(pose_in) in `grad_pose_in`.  A leaf input receives it in `.grad`
        (accumulated like autograd does); a model output continues into the model's graph from there.  With
        preallocated `out=` buffers the gradient tensor is the preallocated one (overwritten by the next step)."""
        x, g = self.pose_in, self.grad_pose_in
        if x.grad_fn is not None:
            torch.autograd.backward(x, g)
        elif x.requires_grad:
            x.grad = g if x.grad is None else x.grad + g


def _gt2_records(gt2):
    """The 2-D target as the kernels want it: contiguous (N,L,B,C) fp32 records.  The reference hands over
    `projection[..., :2]` (data/base/mixins/dataset/projection_2d_mixin.py:224-228), a stride-3 VIEW of 3-channel
    records: that is passed as the 3-channel tensor it lives in (no copy, and the layout the tile kernel takes)."""
    if (gt2 is not None and gt2.is_cuda and gt2.dtype == torch.float32 and gt2.dim() >= 2 and gt2.shape[-1] == 2
            and not gt2.is_contiguous() and gt2.numel() > 0):
        want = [1] * gt2.dim()
        want[-1], acc = 1, 3
        for d in range(gt2.dim() - 2, -1, -1):
            want[d] = acc
            acc *= gt2.shape[d]
        if list(gt2.stride()) == want and \
                gt2.storage_offset() + acc <= gt2.untyped_storage().nbytes() // gt2.element_size():
            return gt2.as_strided(tuple(gt2.shape[:-1]) + (3,), gt2.stride())
    return _f32c(gt2, 'gt_2d')


class _FusedStep(torch.autograd.Function):
    """ProjectionModule.forward -> Normalizer -> loc_2d / loc_3d / loc_2d_3d with the gradient w.r.t. the pose
    input produced in the SAME launch sequence (modules/flow/pose_lifting.py:121-195, flow/base.py:440-469).
    Outputs: the three losses and the module's tensors (relative_pose_rot, absolute_pose_rot, absolute_pose_loc,
    projection_2d, projection_2d_transformed), ALL with an autograd edge back to pose_in.  backward: the fused sum
    only scales the gradient the step kernel already produced; an upstream gradient on any of the module tensors
    (a caller adding another loss on them, e.g. rot_3d on absolute_pose_rot) is chained through the exact backward
    kernels (p2c_hips_neck_normalize_bwd, p2c_fk_bwd, p2c_rot6d_to_matrix_bwd) and added."""

    @staticmethod
    def forward(ctx, pose_in, cfg):
        lib = _lib.load()
        pose_in = _f32c(pose_in, 'pose_in')
        N, L, B = pose_in.shape[:3]
        dev = pose_in.device
        gt2, gt3 = _gt2_records(cfg['gt_2d']), _f32c(cfg['gt_3d'], 'gt_3d')
        gt_rot = _f32c(cfg.get('gt_rot'), 'gt_rot')
        n_losses = 12 if gt_rot is not None else 8
        ref_index = cfg.get('ref_index')
        out = cfg.get('out') or {}
        input_6d = 1 if pose_in.shape[-1] == 6 and pose_in.dim() == 4 else 0
        # a caller that passes the same `out` dict every step (preallocated buffers, see fused_step_buffers) gets the
        # argument block of the previous call back: only the per-step pointers are refreshed
        key = (N, L, B, input_6d, gt2.shape[-1], cfg['mode'], cfg['hips'], cfg['neck'], bool(cfg['normalize']),
               bool(cfg['mask_missing_joints']), float(cfg.get('weight_2d', 1.0)), float(cfg.get('weight_3d', 1.0)),
               bool(cfg.get('want_transformed', True)), id(cfg['tree']), id(cfg['camera']),
               cfg['ref_loc'].data_ptr(), cfg['ref_rot'].data_ptr(),
               None if cfg.get('world_loc') is None else cfg['world_loc'].data_ptr(),
               None if cfg.get('world_rot') is None else cfg['world_rot'].data_ptr(), _lib.raw_stream(),
               None if cfg.get('exchange') is None else cfg['exchange'].block,
               gt_rot is not None, float(cfg.get('weight_rot', 1.0)))
        plan = out.get('_plan')
        if plan is not None and plan[0] == key:
            _, a, rel_rot, abs_rot, abs_loc, proj, proj_norm, g_in, losses, keep = plan
            a.pose_in, a.gt_2d, a.gt_3d, a.ref_index = ptr(pose_in), ptr(gt2), ptr(gt3), ptr(ref_index)
            a.gt_abs_rot = ptr(gt_rot)
            ws, ref_loc, ref_rot, world_loc, world_rot = keep[:5]
        else:
            a = _lib.FusedStep()
            a.tree, a.hips, a.neck = cfg['tree'], cfg['hips'], cfg['neck']
            a.n_clips, a.clip_len, a.mode = N, L, cfg['mode']
            a.input_6d = input_6d
            a.normalize = 1 if cfg['normalize'] else 0
            a.mask_missing_joints = 1 if cfg['mask_missing_joints'] else 0
            ref_loc, ref_rot = _f32c(cfg['ref_loc'], 'ref_loc'), _f32c(cfg['ref_rot'], 'ref_rot')
            world_loc, world_rot = _f32c(cfg.get('world_loc'), 'world_loc'), _f32c(cfg.get('world_rot'), 'world_rot')
            a.pose_in, a.ref_loc, a.ref_rot, a.ref_index = ptr(pose_in), ptr(ref_loc), ptr(ref_rot), ptr(ref_index)
            a.world_loc, a.world_rot, a.camera = ptr(world_loc), ptr(world_rot), cfg['camera']
            a.gt_2d, a.gt_2d_channels, a.gt_3d = ptr(gt2), gt2.shape[-1], ptr(gt3)
            a.weight_2d, a.weight_3d = float(cfg.get('weight_2d', 1.0)), float(cfg.get('weight_3d', 1.0))
            a.gt_abs_rot, a.weight_rot = ptr(gt_rot), float(cfg.get('weight_rot', 1.0))
            if cfg.get('exchange') is not None:
                cfg['exchange'].attach(a)
            rel_rot = out.get('rel_rot') if out.get('rel_rot') is not None else torch.empty((N, L, B, 3, 3), device=dev)
            abs_rot = out.get('abs_rot') if out.get('abs_rot') is not None else torch.empty((N, L, B, 3, 3), device=dev)
            abs_loc = out.get('abs_loc') if out.get('abs_loc') is not None else torch.empty((N, L, B, 3), device=dev)
            proj = out.get('proj') if out.get('proj') is not None else torch.empty((N, L, B, 3), device=dev)
            proj_norm = None
            if cfg.get('want_transformed', True) and cfg['normalize']:
                proj_norm = (out.get('proj_norm') if out.get('proj_norm') is not None
                             else torch.empty((N, L, B, 3), device=dev))
            g_in = out.get('g_pose_in') if out.get('g_pose_in') is not None else torch.empty_like(pose_in)
            nws = lib.p2c_fused_step_workspace_floats(N * L, B, a.input_6d)
            ws = out.get('workspace')
            if ws is None or ws.numel() < nws:
                ws = _workspace(dev, int(nws))
            losses = out.get('losses')
            if losses is None or losses.numel() < n_losses:
                losses = torch.empty(n_losses, device=dev)
            a.rel_rot, a.abs_rot, a.abs_loc, a.proj, a.proj_norm = (ptr(rel_rot), ptr(abs_rot), ptr(abs_loc),
                                                                    ptr(proj), ptr(proj_norm))
            a.g_pose_in, a.workspace, a.losses = ptr(g_in), ptr(ws), ptr(losses)
            if all(out.get(k) is not None for k in ('rel_rot', 'abs_rot', 'abs_loc', 'proj', 'g_pose_in')) \
                    and out.get('losses') is losses \
                    and (proj_norm is None or out.get('proj_norm') is not None):
                # `keep` pins the tensors whose pointers live in the argument block
                out['_plan'] = (key, a, rel_rot, abs_rot, abs_loc, proj, proj_norm, g_in, losses,
                                (ws, ref_loc, ref_rot, world_loc, world_rot, cfg['tree'], cfg['camera']))
        check(lib.p2c_fused_step(a, stream()), 'fused_step')
        if cfg.get('exchange') is not None:
            cfg['exchange'].steps += 1
        # fresh tensor objects over the (possibly preallocated, reused) buffers: autograd attaches this call's node to
        # the objects it is handed
        o_rel, o_absr, o_absl, o_proj = rel_rot.detach(), abs_rot.detach(), abs_loc.detach(), proj.detach()
        o_pn = proj_norm.detach() if proj_norm is not None else pose_in.new_empty(0)
        ctx.save_for_backward(g_in, pose_in, ref_loc, ref_rot, ref_index, world_loc, world_rot, o_rel, o_absr, o_absl,
                              o_proj)
        ctx.cfg = (cfg['tree'], cfg['camera'], cfg['mode'], cfg['hips'], cfg['neck'], input_6d, proj_norm is not None)
        ctx.set_materialize_grads(False)
        res = cfg.setdefault('_result', FusedStepResult())
        res.losses, res.grad_pose_in = losses, g_in
        # loc_2d / loc_3d are reported values (views of the losses buffer, no autograd edge: for one of them alone call
        # with weight_3d = 0 or weight_2d = 0); the fused sum and the module tensors are differentiable
        l2, l3 = losses[0], losses[1]
        ctx.mark_non_differentiable(l2, l3)
        if proj_norm is None:
            ctx.mark_non_differentiable(o_pn)
        return l2, l3, losses[2].clone(), o_rel, o_absr, o_absl, o_proj, o_pn

    @staticmethod
    def backward(ctx, g2d, g3d, g23, g_rel, g_absr, g_absl, g_proj, g_pn):
        g_in, pose_in, ref_loc, ref_rot, ref_index, world_loc, world_rot, rel_rot, abs_rot, abs_loc, proj = \
            ctx.saved_tensors
        tree, camera, mode, hips, neck, input_6d, has_pn = ctx.cfg
        if not has_pn:
            g_pn = None
        # g_pose_in holds d(w2*loc_2d + w3*loc_3d)/d pose_in: for the fused sum the backward is a pure scale
        total = g_in * g23 if g23 is not None else None
        if any(g is not None for g in (g_rel, g_absr, g_absl, g_proj, g_pn)):
            # another loss on the module tensors: the exact chain rule through the standalone backward kernels
            lib = _lib.load()
            N, L, B = pose_in.shape[:3]
            g_proj = _f32c(g_proj, 'g_proj')
            if g_pn is not None:
                gp = torch.empty_like(proj)
                check(lib.p2c_hips_neck_normalize_bwd(ptr(proj), ptr(_f32c(g_pn, 'g_proj_norm')), ptr(gp), N * L, B, 3,
                                                      hips, neck, 1e-5, stream()), 'hips_neck_normalize_bwd')
                g_proj = gp if g_proj is None else g_proj + gp
            mats = pose_in
            if input_6d:
                mats = torch.empty((N, L, B, 3, 3), device=pose_in.device)
                check(lib.p2c_rot6d_to_matrix_fwd(ptr(pose_in), ptr(mats), N * L * B, stream()), 'rot6d_to_matrix_fwd')
            a = _lib.FkBwd()
            a.tree, a.n_clips, a.clip_len, a.mode = tree, N, L, mode
            a.pose_in, a.ref_loc, a.ref_index = ptr(mats), ptr(ref_loc), ptr(ref_index)
            a.ref_rot = ptr(ref_rot) if mode == _lib.MODE_POSE_CHANGES else None
            a.world_loc, a.world_rot, a.camera = ptr(world_loc), ptr(world_rot), camera
            a.rel_rot = ptr(rel_rot) if mode == _lib.MODE_POSE_CHANGES else None
            a.abs_rot, a.abs_loc = ptr(abs_rot), ptr(abs_loc)
            g_rel = _f32c(g_rel, 'g_rel_rot')
            g_absr, g_absl = _f32c(g_absr, 'g_abs_rot'), _f32c(g_absl, 'g_abs_loc')
            a.g_rel_rot = ptr(g_rel) if mode == _lib.MODE_POSE_CHANGES else None
            a.g_abs_rot, a.g_abs_loc, a.g_proj = ptr(g_absr), ptr(g_absl), ptr(g_proj)
            g_mats = torch.empty_like(mats)
            a.g_pose_in = ptr(g_mats)
            check(lib.p2c_fk_bwd(a, stream()), 'fk_bwd')
            if mode != _lib.MODE_POSE_CHANGES and g_rel is not None:
                g_mats = g_mats + g_rel   # relative_rot mode: rel_rot IS the input matrix (no scan)
            extra = g_mats
            if input_6d:
                extra = torch.empty_like(pose_in)
                check(lib.p2c_rot6d_to_matrix_bwd(ptr(pose_in), ptr(g_mats), ptr(extra), N * L * B, stream()),
                      'rot6d_to_matrix_bwd')
            total = extra if total is None else total + extra
        return total, None


def fused_step_buffers(n_clips, clip_len, n_bones, device, input_6d=True, transformed=False, rot=False):
    """Preallocated outputs for fused_step(out=...): a training loop with static shapes passes the same dict every
    step, which also lets fused_step reuse its argument block (no per-step allocations; the outputs of step i are
    overwritten by step i+1)."""
    n, l, b = n_clips, clip_len, n_bones
    out = dict(rel_rot=torch.empty((n, l, b, 3, 3), device=device), abs_rot=torch.empty((n, l, b, 3, 3), device=device),
               abs_loc=torch.empty((n, l, b, 3), device=device), proj=torch.empty((n, l, b, 3), device=device),
               g_pose_in=torch.empty((n, l, b, 6) if input_6d else (n, l, b, 3, 3), device=device),
               losses=torch.empty(12 if rot else 8, device=device))
    if transformed:
        out['proj_norm'] = torch.empty((n, l, b, 3), device=device)
    return out


def fused_step(pose_in, tree, camera, ref_loc, ref_rot, gt_2d, gt_3d, ref_index=None, world_loc=None, world_rot=None,
               mode=_lib.MODE_POSE_CHANGES, hips=1, neck=8, normalize=True, mask_missing_joints=True, weight_2d=1.0,
               weight_3d=1.0, want_transformed=True, out=None, exchange=None, gt_rot=None, weight_rot=1.0):
    """One pose_lifting step of the hot path in one call.  pose_in: (N,L,B,6) rotation_6d model output or
    (N,L,B,3,3) matrices.  Returns FusedStepResult; .loc_2d_3d (= weight_2d*loc_2d + weight_3d*loc_3d) and the module
    tensors (.relative_pose_rot, .absolute_pose_rot, .absolute_pose_loc, .projection_2d, .projection_2d_transformed)
    carry autograd edges back to pose_in (.loc_2d / .loc_3d are reported values).  `out` may hold preallocated
    buffers (CUDA-graph friendly; the outputs of step i are then overwritten by step i+1, so any extra loss on them
    must be back-propagated before the next call).  `exchange`: a LossExchange (exchange.py) -- the step then also
    publishes its four loss sums to every data-parallel rank.  `gt_rot` = targets['absolute_pose_rot']: rot_3d
    (loss/rot_3d.py) is fused as well -- `.rot_3d`, and `.total` = weight_2d*loc_2d + weight_3d*loc_3d + weight_rot*rot_3d
    carries the gradient (the loc_rot_3d / loc_2d_loc_rot_3d / weighted_loc_2d_loc_rot_3d modes of loss/__init__.py:28-48
    are weightings of it)."""
    cfg = dict(tree=tree, camera=camera, ref_loc=ref_loc, ref_rot=ref_rot, gt_2d=gt_2d, gt_3d=gt_3d,
               ref_index=ref_index, world_loc=world_loc, world_rot=world_rot, mode=mode, hips=hips, neck=neck,
               normalize=normalize, mask_missing_joints=mask_missing_joints, weight_2d=weight_2d,
               weight_3d=weight_3d, want_transformed=want_transformed, out=out, exchange=exchange, gt_rot=gt_rot,
               weight_rot=weight_rot)
    l2, l3, l23, rel_rot, abs_rot, abs_loc, proj, proj_norm = _FusedStep.apply(pose_in, cfg)
    res = cfg['_result']
    res.loc_2d, res.loc_3d, res.loc_2d_3d, res.pose_in = l2, l3, l23, pose_in
    # with gt_rot (targets['absolute_pose_rot']) the differentiable sum is weight_2d*loc_2d + weight_3d*loc_3d +
    # weight_rot*rot_3d (`total`; `loc_2d_3d` names the same tensor); rot_3d alone is a reported value
    res.total = l23
    res.rot_3d = res.losses[8] if gt_rot is not None else None
    res.relative_pose_rot, res.absolute_pose_rot, res.absolute_pose_loc = rel_rot, abs_rot, abs_loc
    res.projection_2d = proj
    res.projection_2d_transformed = proj_norm if (want_transformed and normalize) else None
    return res
