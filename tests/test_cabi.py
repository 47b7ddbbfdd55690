"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads, and exports every symbol that
include/p2c_b200.h declares; the ctypes structs mirror the header; CPU tensors are refused (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'p2c_b200.h')


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(p2c_[a-z0-9_]+)\s*\(', src)))


@pytest.fixture(scope='module')
def lib():
    from pedestrians_video_2_carla_b200 import _build, _lib
    _build.build()
    return _lib


def test_header_and_binding_agree(lib):
    assert declared_functions() == lib.exported_symbols()


def test_library_exports_every_symbol(lib):
    cdll = ctypes.CDLL(lib.lib_path())
    for name in declared_functions():
        assert hasattr(cdll, name), name
    assert b'sm_100a' in lib.load().p2c_version()


def test_struct_sizes_match_header(lib, tmp_path):
    """Compile a tiny C program against the header and compare sizeof() with the ctypes mirrors."""
    import subprocess
    src = tmp_path / 'sz.c'
    src.write_text('#include <stdio.h>\n#include "p2c_b200.h"\nint main(void){printf("%zu %zu %zu %zu %zu\\n",'
                   'sizeof(p2c_tree_t),sizeof(p2c_camera_t),sizeof(p2c_fk_fwd_t),sizeof(p2c_fk_bwd_t),'
                   'sizeof(p2c_fused_step_t));return 0;}\n')
    exe = tmp_path / 'sz'
    subprocess.check_call(['gcc', '-I', os.path.join(ROOT, 'include'), str(src), '-o', str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert sizes == [ctypes.sizeof(t) for t in (lib.Tree, lib.Camera, lib.FkFwd, lib.FkBwd, lib.FusedStep)]


def test_bad_arguments_return_codes_without_gpu(lib):
    """Argument validation happens before any CUDA call, so it is checkable on a CPU-only box."""
    l = lib.load()
    assert l.p2c_fk_fwd(None, None) == -3            # P2C_E_NULL
    a = lib.FkFwd()
    a.tree = lib.make_tree([-1, 0, 1])
    a.n_clips, a.clip_len, a.mode = 1, 0, 0
    assert l.p2c_fk_fwd(a, None) == -1               # P2C_E_BADARG (clip_len < 1)
    a.tree = lib.make_tree([-1, 2, 1])
    a.clip_len = 1
    assert l.p2c_fk_fwd(a, None) == -2               # P2C_E_BONES (parent after child)
    assert l.p2c_rot6d_to_matrix_fwd(None, None, 4, None) == -3


def test_cpu_tensors_are_refused(lib):
    from pedestrians_video_2_carla_b200 import ops
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ops.rotation_6d_to_matrix(torch.zeros(2, 6))
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ops.hips_neck_normalize(torch.zeros(2, 26, 3), 1, 8)


def test_widened_entry_points_validate_arguments_without_gpu(lib):
    """SURVEY 8f entry points: argument validation precedes any CUDA call; empty batches are a no-op (rc 0)."""
    l = lib.load()
    assert l.p2c_cum_changes_fwd(None, None, 4, 0, 26, None) == -1          # clip_len < 1
    assert l.p2c_cum_changes_fwd(None, None, 4, 16, 26, None) == -3         # NULL tensors
    assert l.p2c_cum_changes_fwd(None, None, 0, 16, 26, None) == 0          # empty batch
    assert l.p2c_cum_changes_bwd(None, None, None, None, 0, 16, 26, None) == 0
    assert l.p2c_renormalize_fwd(None, None, 8, 26, 3, 4, 1, 8, 1, 1e-5, None, None, None, 16, None) == -1   # dim 4
    assert l.p2c_renormalize_fwd(None, None, 8, 26, 2, 3, 1, 8, 1, 1e-5, None, None, None, 16, None) == -1   # c < dim
    assert l.p2c_renormalize_fwd(None, None, 8, 26, 3, 3, 1, 40, 1, 1e-5, None, None, None, 16, None) == -1  # neck >= J
    assert l.p2c_renormalize_fwd(None, None, 8, 26, 3, 3, 1, 8, 1, 1e-5, None, None, None, 16, None) == -3
    assert l.p2c_renormalize_fwd(None, None, 0, 26, 3, 3, 1, 8, 1, 1e-5, None, None, None, 16, None) == 0
    assert l.p2c_quaternion_to_matrix_fwd(None, None, -1, None) == -1
    assert l.p2c_quaternion_to_matrix_fwd(None, None, 5, None) == -3
    assert l.p2c_normalize_extractor_fwd(None, None, 8, 26, 3, 1, 8, 0, 1e-5, None) == -1   # extractor 0 is the plain op
    assert l.p2c_normalize_extractor_fwd(None, None, 8, 26, 3, 1, 8, 2, 1e-5, None) == -3
    assert l.p2c_mpjpe_sum(None, None, 8, 0, 3, None, None, 1, None) == -1
    assert l.p2c_pck_count(None, None, None, 8, 26, 3, 0, 1, 1, 8, 2, 0.05, 1e-5, None, None, 1, None) == -1  # norm kind
    assert l.p2c_metric_workspace_floats() > 0
