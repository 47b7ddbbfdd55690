"""p2c_loc2d_step_static: the loc_2d step with its scratch in a file-scope device array (no allocation inside the call),
called through the C ABI and compared with p2c_loc2d_step (caller-allocated workspace) and with the oracle.

Collected LAST on purpose (the file name sorts after every other test file): the entry point has no counterpart in the
reference and a failure here must not hide the parity tests under `pytest -x`.  The tests run the same alone in a fresh
process and after the rest of the suite (the scratch array is shared by every call on the device and keeps the last
call's contents; nothing here depends on what it held before)."""
import pytest
import torch

import oracle as O
from _util import assert_close

pytestmark = pytest.mark.gpu

DEV = 'cuda:0'
P2C_E_TOO_LARGE = -5


@pytest.fixture(scope='module')
def lib():
    from pedestrians_video_2_carla_b200 import _lib
    _lib.load()
    return _lib


def _case(shape, c_gt, seed):
    g = torch.Generator().manual_seed(seed)
    pred = torch.randn(shape + (2,), generator=g, dtype=torch.float64)
    gt = torch.randn(shape + (c_gt,), generator=g, dtype=torch.float64)
    gt[torch.rand(shape, generator=g) < 0.15] = 0
    return pred, gt


def _both(lib, pred, gt, hips, upstream):
    """One call of each entry point on the same device tensors; every buffer is allocated BEFORE the static call."""
    L = lib.load()
    m, j, c_gt = pred.numel() // (pred.shape[-2] * 2), pred.shape[-2], gt.shape[-1]
    p, t = pred.float().to(DEV).contiguous(), gt.float().to(DEV).contiguous()
    ws = torch.zeros(int(L.p2c_loc2d_step_workspace_floats()), device=DEV)
    g_ref, g_new = torch.empty_like(p), torch.full_like(p, float('nan'))
    out_ref, out_new = torch.zeros(4, device=DEV), torch.full((4,), float('nan'), device=DEV)
    stream = lib.stream()
    lib.check(L.p2c_loc2d_step(lib.ptr(p), lib.ptr(t), m, j, c_gt, hips, upstream, lib.ptr(g_ref), lib.ptr(ws),
                               lib.ptr(out_ref), None, 0, 1, stream), 'loc2d_step')
    torch.cuda.synchronize()
    held = torch.cuda.memory_allocated(DEV), torch.cuda.memory_reserved(DEV)
    lib.check(L.p2c_loc2d_step_static(lib.ptr(p), lib.ptr(t), m, j, c_gt, hips, upstream, lib.ptr(g_new),
                                      lib.ptr(out_new), stream), 'loc2d_step_static')
    torch.cuda.synchronize()
    assert (torch.cuda.memory_allocated(DEV), torch.cuda.memory_reserved(DEV)) == held
    return (out_ref.cpu(), g_ref.cpu()), (out_new.cpu(), g_new.cpu())


def test_static_scratch_matches_workspace_entry_two_calls_in_a_row(lib):
    """Two calls in a row on different batches (the second smaller, so its grid leaves stale partial sums of the first
    in the scratch): loss, sums and gradient equal p2c_loc2d_step's and the oracle's."""
    for shape, c_gt, seed, ups in (((8192, 16, 26), 2, 5, 1.0), ((7, 3, 26), 3, 6, 3.0), ((256, 16, 26), 2, 7, 0.5)):
        pred, gt = _case(shape, c_gt, seed)
        (out_ref, g_ref), (out_new, g_new) = _both(lib, pred, gt, 1, ups)
        assert_close('loss vs p2c_loc2d_step', out_new[0], out_ref[0])
        assert_close('sum_sq vs p2c_loc2d_step', out_new[1], out_ref[1])
        assert float(out_new[2]) == float(out_ref[2])
        assert float(out_new[3]) == 0.0
        assert_close('grad vs p2c_loc2d_step', g_new, g_ref)
        assert torch.equal(out_new, out_ref) and torch.equal(g_new, g_ref)   # same kernel, same grid, same order
        p64 = pred.clone().requires_grad_(True)
        want = O.loss_loc_2d(p64, gt, True, 1)
        (ups * want).backward()
        assert_close('loss vs oracle', out_new[0], want)
        assert_close('grad vs oracle', g_new, p64.grad)
        kept = O.missing_joints_mask(gt[..., :2], 1)
        assert float(out_new[2]) == float(2 * kept.sum())


def test_static_scratch_same_batch_twice_is_bitwise_repeatable(lib):
    pred, gt = _case((300, 16, 26), 2, 11)
    _, (out_a, g_a) = _both(lib, pred, gt, 1, 1.0)
    _, (out_b, g_b) = _both(lib, pred, gt, 1, 1.0)
    assert torch.equal(out_a, out_b) and torch.equal(g_a, g_b)


def test_static_scratch_rejects_sizes_beyond_its_maximum(lib):
    """Beyond the stated maximum: P2C_E_TOO_LARGE before anything is launched (the buffers are never touched)."""
    L = lib.load()
    cap = int(L.p2c_loc2d_step_static_max_joints())
    assert cap == 148 * 25600
    small = torch.zeros(64, device=DEV)
    out = torch.full((4,), 7.0, device=DEV)
    stream = lib.stream()
    for m, j in ((cap // 26 + 1, 26), (cap + 1, 1), (1 << 40, 26)):
        rc = L.p2c_loc2d_step_static(lib.ptr(small), lib.ptr(small), m, j, 2, -1, 1.0, lib.ptr(small), lib.ptr(out), stream)
        assert rc == P2C_E_TOO_LARGE
    torch.cuda.synchronize()
    assert torch.equal(out.cpu(), torch.full((4,), 7.0))
    with pytest.raises(RuntimeError, match='P2C_E_TOO_LARGE'):
        lib.check(P2C_E_TOO_LARGE, 'loc2d_step_static')
    # the largest batch it takes still runs (all of a B200's SMs, 200 KB of errors each)
    if torch.cuda.get_device_properties(0).multi_processor_count >= 148:
        n = cap // 26
        pred, gt = _case((n, 1, 26), 2, 13)
        (out_ref, g_ref), (out_new, g_new) = _both(lib, pred, gt, 1, 1.0)
        assert_close('loss at the maximum', out_new[0], out_ref[0])
        assert_close('grad at the maximum', g_new, g_ref)
