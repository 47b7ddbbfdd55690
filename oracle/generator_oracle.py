"""CPU restatement of the device draw kernel of the synthetic-target generator (csrc/generator.cu), bit for bit.

TEST INFRASTRUCTURE.  The reference draws with numpy / CPU torch generators on the host
(data/carla/datasets/carla_2d3d_dataset.py:157-182); those streams have no device form, so the product defines its own
counter-based stream (Philox4x32-10, Salmon et al. 2011) and this file pins it: integer work, compared bit-exactly.
What is pinned to the REFERENCE is the distribution (exactly k distinct bones per frame, angles uniform in +-max, yaw
only, a uniform skeleton type per clip) and everything computed FROM the draws (tests/test_gpu_widen.py)."""
import numpy as np

__all__ = ['philox4x32_10', 'draw_sparse_changes']

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint32 arrays (counters) with scalar keys."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & MASK for c in (c0, c1, c2, c3))
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def _symmetric(r, amp):
    u = (r >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return (u * np.float32(2.0) - np.float32(1.0)) * np.float32(amp)


def draw_sparse_changes(seed, offset, n_clips, clip_len, n_bones, k, max_change, max_world=0.0, max_world_initial=0.0,
                        n_types=4):
    """-> euler (N,L,B,3) float32, world_euler (N,L,3) float32, skeleton_index (N,) int32."""
    poses = n_clips * clip_len
    m = np.arange(poses, dtype=np.uint32)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    o0, o1 = offset & 0xFFFFFFFF, (offset >> 32) & 0xFFFFFFFF
    euler = np.zeros((poses, n_bones, 3), dtype=np.float32)
    chosen = np.zeros((poses, n_bones), dtype=bool)
    for i in range(k):
        r = philox4x32_10(m, np.full(poses, i, np.uint32), np.full(poses, o0, np.uint32), np.full(poses, o1, np.uint32), k0, k1)
        pick = (r[0] % np.uint32(n_bones - i)).astype(np.int64)
        # the pick-th bone not chosen yet
        free_rank = np.cumsum(~chosen, axis=1) - 1
        bone = np.argmax((~chosen) & (free_rank == pick[:, None]), axis=1)
        chosen[np.arange(poses), bone] = True
        for ch in range(3):
            euler[np.arange(poses), bone, ch] = _symmetric(r[1 + ch], max_change)
    w = philox4x32_10(m, np.full(poses, 0xFFFFFFFF, np.uint32), np.full(poses, o0 ^ 0x5bd1e995, np.uint32),
                      np.full(poses, o1, np.uint32), k0, k1)
    frame = (np.arange(poses) % clip_len)
    yaw = np.zeros(poses, dtype=np.float32)
    if max_world_initial > 0:
        yaw = np.where(frame == 0, _symmetric(w[0], max_world_initial), yaw)
    if max_world != 0:
        yaw = np.where(frame > 0, _symmetric(w[1], max_world), yaw)
    world = np.zeros((poses, 3), dtype=np.float32)
    world[:, 2] = yaw
    skeleton = (w[2][frame == 0] % np.uint32(n_types)).astype(np.int32)
    return (euler.reshape(n_clips, clip_len, n_bones, 3), world.reshape(n_clips, clip_len, 3), skeleton)
