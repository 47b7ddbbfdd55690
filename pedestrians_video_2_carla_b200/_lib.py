"""ctypes binding of include/p2c_b200.h.  The product path has NO CPU fallback: if the library is
missing or a tensor is not a CUDA fp32 tensor, calls raise RuntimeError."""
import ctypes as C
import os

import torch

from . import _build

MAX_BONES = 32
MODE_POSE_CHANGES, MODE_RELATIVE_ROT = 0, 1


class Tree(C.Structure):
    _fields_ = [('n_bones', C.c_int32), ('parent', C.c_int32 * MAX_BONES)]


class Camera(C.Structure):
    _fields_ = [('R', C.c_float * 9), ('T', C.c_float * 3), ('f', C.c_float), ('px', C.c_float), ('py', C.c_float)]


_P = C.c_void_p


class FkFwd(C.Structure):
    _fields_ = [('tree', Tree), ('n_clips', C.c_int32), ('clip_len', C.c_int32), ('mode', C.c_int32),
                ('pose_in', _P), ('ref_loc', _P), ('ref_rot', _P), ('ref_index', _P), ('world_loc', _P),
                ('world_rot', _P), ('camera', Camera), ('rel_rot', _P), ('abs_rot', _P), ('abs_loc', _P),
                ('proj', _P)]


class FkBwd(C.Structure):
    _fields_ = [('tree', Tree), ('n_clips', C.c_int32), ('clip_len', C.c_int32), ('mode', C.c_int32),
                ('pose_in', _P), ('ref_loc', _P), ('ref_rot', _P), ('ref_index', _P), ('world_loc', _P),
                ('world_rot', _P), ('camera', Camera), ('rel_rot', _P), ('abs_rot', _P), ('abs_loc', _P),
                ('g_rel_rot', _P), ('g_abs_rot', _P), ('g_abs_loc', _P), ('g_proj', _P), ('g_pose_in', _P),
                ('g_ref_rot', _P), ('g_ref_loc', _P), ('g_world_loc', _P), ('g_world_rot', _P)]


class FusedStep(C.Structure):
    _fields_ = [('tree', Tree), ('hips', C.c_int32), ('neck', C.c_int32),
                ('n_clips', C.c_int32), ('clip_len', C.c_int32), ('mode', C.c_int32), ('input_6d', C.c_int32),
                ('normalize', C.c_int32), ('mask_missing_joints', C.c_int32),
                ('pose_in', _P), ('ref_loc', _P), ('ref_rot', _P), ('ref_index', _P), ('world_loc', _P),
                ('world_rot', _P), ('camera', Camera), ('gt_2d', _P), ('gt_2d_channels', C.c_int32),
                ('gt_3d', _P), ('weight_2d', C.c_float), ('weight_3d', C.c_float),
                ('rel_rot', _P), ('abs_rot', _P), ('abs_loc', _P), ('proj', _P), ('proj_norm', _P),
                ('g_pose_in', _P), ('workspace', _P), ('losses', _P),
                ('exchange', _P), ('exchange_rank', C.c_int32), ('exchange_world', C.c_int32),
                ('gt_abs_rot', _P), ('weight_rot', C.c_float)]


_SIGS = {
    'p2c_fk_fwd': (C.c_int, [C.POINTER(FkFwd), _P]),
    'p2c_fk_bwd': (C.c_int, [C.POINTER(FkBwd), _P]),
    'p2c_project_fwd': (C.c_int, [_P, _P, _P, C.POINTER(Camera), _P, C.c_int64, C.c_int32, _P]),
    'p2c_project_bwd': (C.c_int, [_P, _P, _P, C.POINTER(Camera), _P, _P, _P, _P, C.c_int64, C.c_int32, _P]),
    'p2c_world_scan_fwd': (C.c_int, [_P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32, _P]),
    'p2c_world_scan_bwd': (C.c_int, [_P, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32, _P]),
    'p2c_rot6d_to_matrix_fwd': (C.c_int, [_P, _P, C.c_int64, _P]),
    'p2c_rot6d_to_matrix_bwd': (C.c_int, [_P, _P, _P, C.c_int64, _P]),
    'p2c_euler_xyz_to_matrix': (C.c_int, [_P, _P, C.c_int64, _P]),
    'p2c_matrix_to_euler_xyz': (C.c_int, [_P, _P, C.c_int64, _P]),
    'p2c_hips_neck_normalize_fwd': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                              C.c_float, _P]),
    'p2c_hips_neck_normalize_bwd': (C.c_int, [_P, _P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                              C.c_float, _P]),
    'p2c_mse_workspace_floats': (C.c_int64, [C.c_int64]),
    'p2c_masked_mse_fwd': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P,
                                     _P, _P]),
    'p2c_masked_mse_bwd': (C.c_int, [_P, _P, _P, _P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_int32, _P]),
    'p2c_loc2d_step_workspace_floats': (C.c_int64, []),
    'p2c_loc2d_step_supported': (C.c_int, [C.c_int64, C.c_int32, C.c_int32]),
    'p2c_loc2d_step': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_float, _P, _P, _P, _P,
                                 C.c_int32, C.c_int32, _P]),
    'p2c_loc2d_step_static_max_joints': (C.c_int64, []),
    'p2c_loc2d_step_static': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_float, _P, _P, _P]),
    'p2c_fused_step_workspace_floats': (C.c_int64, [C.c_int64, C.c_int32, C.c_int32]),
    'p2c_fused_step': (C.c_int, [C.POINTER(FusedStep), _P]),
    'p2c_fused_step_launches': (C.c_int, [C.POINTER(FusedStep)]),
    'p2c_exchange_block_bytes': (C.c_int64, []),
    'p2c_exchange_max_ranks': (C.c_int32, []),
    'p2c_exchange_create': (C.c_int, [C.POINTER(_P), C.c_char_p]),
    'p2c_exchange_open': (C.c_int, [C.c_char_p, C.POINTER(_P)]),
    'p2c_exchange_close': (C.c_int, [_P]),
    'p2c_exchange_destroy': (C.c_int, [_P]),
    'p2c_exchange_set_peers': (C.c_int, [_P, C.POINTER(_P), C.c_int32]),
    'p2c_exchange_read': (C.c_int, [_P, C.c_int32, C.c_uint32, _P, _P]),
    'p2c_exchange_flush': (C.c_int, [_P, C.c_int32, C.c_int32, _P]),
    'p2c_quaternion_to_matrix_fwd': (C.c_int, [_P, _P, C.c_int64, _P]),
    'p2c_quaternion_to_matrix_bwd': (C.c_int, [_P, _P, _P, C.c_int64, _P]),
    'p2c_cum_changes_fwd': (C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, _P]),
    'p2c_cum_changes_bwd': (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P]),
    'p2c_renormalize_fwd': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_float, _P, _P, _P, C.c_int32, _P]),
    'p2c_renormalize_bwd': (C.c_int, [_P, _P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_float, _P, _P, C.c_int32, _P]),
    'p2c_normalize_extractor_fwd': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                              C.c_int32, C.c_float, _P]),
    'p2c_extractor_shift_scale': (C.c_int, [_P, _P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                            C.c_int32, C.c_float, _P]),
    'p2c_metric_workspace_floats': (C.c_int64, []),
    'p2c_mpjpe_sum': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, _P, _P, C.c_int32, _P]),
    'p2c_mrpe_sum': (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, _P, _P, C.c_int32, _P]),
    'p2c_pck_count': (C.c_int, [_P, _P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                C.c_int32, C.c_int32, C.c_float, C.c_float, _P, _P, C.c_int32, _P]),
    'p2c_draw_sparse_changes': (C.c_int, [C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float,
                                          C.c_float, C.c_float, C.c_int32, _P, _P, _P, _P]),
    'p2c_version': (C.c_char_p, []),
    'p2c_debug_reload_env': (None, []),
}

_lib = None


def lib_path():
    """The in-tree library; P2C_LIB names an experiment build of the same sources (tools/variants.py)."""
    return os.environ.get('P2C_LIB') or _build.LIB_PATH


def load():
    """dlopen libp2c_b200.so (built in-tree by __graft_entry__.build()); fail loudly if absent."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError('p2c_b200: %s is missing -- run `python __graft_entry__.py` (build()) first; '
                           'there is no CPU fallback' % path)
    lib = C.CDLL(path)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)      # AttributeError if the header and the library disagree
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def exported_symbols():
    return sorted(_SIGS)


def check(rc, what):
    if rc != 0:
        kind = 'bad argument' if rc < 0 else 'CUDA error'
        if rc == -5:
            kind = 'size beyond the maximum of this entry point (P2C_E_TOO_LARGE)'
        raise RuntimeError('p2c_b200.%s failed: %s %d' % (what, kind, rc))


def ptr(t, name='tensor'):
    """Device pointer of a contiguous fp32 (or int32) CUDA tensor; None -> NULL."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError('p2c_b200: %s must be a CUDA tensor (no CPU fallback); got device %s' % (name, t.device))
    if t.dtype not in (torch.float32, torch.int32, torch.float64):
        raise RuntimeError('p2c_b200: %s must be float32/int32 (float64: metric accumulators), got %s' % (name, t.dtype))
    if not t.is_contiguous():
        raise RuntimeError('p2c_b200: %s must be contiguous' % name)
    return C.c_void_p(t.data_ptr())


def raw_stream():
    """cudaStream_t of PyTorch's current stream on the current device as an int (the raw query: the Stream object
    torch.cuda.current_stream() builds costs ~5 us per call, a third of the fused step's kernel time)."""
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


def stream():
    return C.c_void_p(raw_stream())


def make_tree(parents):
    t = Tree()
    t.n_bones = len(parents)
    if len(parents) > MAX_BONES:
        raise RuntimeError('p2c_b200: at most %d bones are supported, got %d' % (MAX_BONES, len(parents)))
    for i in range(MAX_BONES):
        t.parent[i] = parents[i] if i < len(parents) else -1
    return t


def make_camera(R, T, f, px, py):
    c = Camera()
    flat = [float(v) for row in R for v in row]
    for i in range(9):
        c.R[i] = flat[i]
    for i in range(3):
        c.T[i] = float(T[i])
    c.f, c.px, c.py = float(f), float(px), float(py)
    return c
